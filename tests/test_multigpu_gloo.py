"""world_size-2 test of the multi-GPU plumbing on the CPU (gloo): BVH broadcast, ray sharding,
hit-record gather and counter reduction -- the host logic bench.py uses with NCCL at N > 1."""
import importlib
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, ws, port, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    mg = importlib.import_module("gpu-ray-tracing_b200.multigpu")
    rng = np.random.RandomState(5)
    truth = {"layout": 5, "nodes": rng.rand(16 * 37).astype(np.float32), "woop": rng.rand(4 * 101).astype(np.float32),
             "triIndex": rng.randint(0, 1000, 101).astype(np.int32)}
    got = mg.broadcast_bvh(truth if rank == 0 else None, "cpu")
    for k in ("nodes", "woop", "triIndex"):
        assert got[k].tobytes() == truth[k].tobytes(), k
    # shard 1001 "rays", each rank fabricates its slice of hit records, gather restores the whole
    total = 1001
    lo, hi = mg.shard_range(total, rank, ws)
    full = torch.arange(total * 4, dtype=torch.int32).reshape(total, 4)
    gathered = mg.gather_results(full[lo:hi].clone(), total, "cpu")
    assert torch.equal(gathered, full)
    s = mg.reduce_stats([hi - lo, 1.0], "cpu", op="sum")
    m = mg.reduce_stats([float(rank + 1)], "cpu", op="max")
    assert s == [float(total), float(ws)] and m == [float(ws)]
    dist.barrier()
    dist.destroy_process_group()
    open(os.path.join(out_dir, "ok%d" % rank), "w").close()


def test_broadcast_shard_gather_world2(tmp_path):
    ws = 2
    mp.spawn(_worker, args=(ws, _free_port(), str(tmp_path)), nprocs=ws, join=True)
    assert all((tmp_path / ("ok%d" % r)).exists() for r in range(ws))
