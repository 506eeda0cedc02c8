"""Multi-GPU plumbing: one process per GPU, BVH replicated, ray batches sharded.

The reference is single-GPU (no NCCL, no peer access -- SURVEY.md 2a).  Rays are independent, so
the path shards with no exchange step inside a trace:
  * scene: rank 0 builds the BVH on its host and the three Compact2 arrays travel to every other
    rank with torch.distributed.broadcast (NCCL over NVLink on the GPU box, gloo in CPU tests);
  * rays: each rank takes the contiguous slice [lo, hi) of a batch -- after the Morton sort, so a
    slice stays spatially coherent -- or generates its own slice on the device;
  * hits: all_gather of the 16-byte result records (padded to equal slice sizes), and an
    all_reduce of the hit / ray counters and of the per-rank kernel milliseconds (max).
No collective runs inside the timed trace.
"""
import numpy as np
import torch
import torch.distributed as dist


def world():
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def shard_range(n, rank, world_size):
    """Contiguous slice of n units for `rank`: sizes differ by at most one, union is [0, n)."""
    base, rem = divmod(int(n), int(world_size))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def _comm_device(device):
    """Tensors must live where the backend can see them: GPU for nccl, CPU for gloo."""
    if dist.get_backend() == "nccl":
        return torch.device(device)
    return torch.device("cpu")


def broadcast_bvh(arrays, device, src=0):
    """Replicate the Compact2 arrays of rank `src` on every rank.  `arrays` is a dict with numpy
    nodes / woop / triIndex on `src` and may be None elsewhere.  Returns the dict on every rank: device tensors under
    nccl, numpy arrays under gloo and for a single rank."""
    rank, ws = world()
    if ws == 1:
        return arrays
    cd = _comm_device(device)
    sizes = torch.zeros(3, dtype=torch.int64, device=cd)
    if rank == src:
        sizes = torch.tensor([arrays["nodes"].size, arrays["woop"].size, arrays["triIndex"].size],
                             dtype=torch.int64, device=cd)
    dist.broadcast(sizes, src)
    n_nodes, n_woop, n_tri = (int(v) for v in sizes.tolist())
    out = {"layout": 5}
    for key, n, dt, npdt in (("nodes", n_nodes, torch.float32, np.float32), ("woop", n_woop, torch.float32, np.float32),
                             ("triIndex", n_tri, torch.int32, np.int32)):
        if rank == src:
            t = torch.from_numpy(np.ascontiguousarray(arrays[key], npdt)).to(cd)
        else:
            t = torch.empty(n, dtype=dt, device=cd)
        dist.broadcast(t, src)
        # nccl: the arrays stay on the device -- CudaBVH takes device tensors as they are, so no rank pulls the BVH back to
        # the host to upload it again (1.2 GB on the 6.5 M-triangle scene); gloo (CPU tests): numpy, as on rank `src`
        out[key] = t if t.is_cuda else t.numpy()
    return out


def gather_results(local_results, total, device):
    """all_gather the per-rank slices of 16-byte result records (int32 [n_r, 4]) into [total, 4],
    in rank order.  Slices are padded to the largest slice for the collective."""
    rank, ws = world()
    if ws == 1:
        return local_results
    cd = _comm_device(device)
    per = max(shard_range(total, r, ws)[1] - shard_range(total, r, ws)[0] for r in range(ws))
    pad = torch.zeros((per, 4), dtype=torch.int32, device=cd)
    pad[: local_results.shape[0]] = local_results.to(cd)
    out = torch.empty((ws * per, 4), dtype=torch.int32, device=cd)
    dist.all_gather_into_tensor(out, pad) if dist.get_backend() == "nccl" else \
        dist.all_gather(list(out.view(ws, per, 4).unbind(0)), pad)
    parts = []
    for r in range(ws):
        lo, hi = shard_range(total, r, ws)
        parts.append(out[r * per: r * per + (hi - lo)])
    return torch.cat(parts, 0)


def reduce_stats(values, device, op="sum"):
    """all_reduce a list of python numbers (sum or max) -- hit counts, ray counts, kernel ms."""
    rank, ws = world()
    if ws == 1:
        return list(values)
    t = torch.tensor(values, dtype=torch.float64, device=_comm_device(device))
    dist.all_reduce(t, dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX)
    return t.tolist()


def max_over_ranks(value, device):
    return int(reduce_stats([float(value)], device, op="max")[0])


def all_sizes(value, device):
    """The integer `value` of every rank, in rank order."""
    rank, ws = world()
    if ws == 1:
        return [int(value)]
    t = torch.zeros(ws, dtype=torch.float64, device=_comm_device(device))
    t[rank] = float(value)
    dist.all_reduce(t, dist.ReduceOp.SUM)
    return [int(v) for v in t.tolist()]


def local_core_slice(cores, local_rank, local_world):
    """The share of the allowed host cores that rank `local_rank` of `local_world` ranks on this node keeps: a contiguous
    block, every core in exactly one block, sizes differ by at most one.  With fewer cores than ranks everybody keeps all."""
    cores = sorted(cores)
    if local_world <= 1 or len(cores) < local_world:
        return cores
    lo, hi = shard_range(len(cores), local_rank, local_world)
    return cores[lo:hi]


def bind_to_local_cores(local_rank, local_world):
    """Pin this process to its GPU's NUMA-local host cores, and inside them to its own block, BEFORE it allocates pinned
    memory: cudaHostAlloc pages are first-touched by the calling thread, so they land on the NUMA node it runs on, and the
    runtime's copy-submission threads of eight ranks no longer share cores.  (Round 1: end-to-end throughput through host
    buffers scaled 1 -> 2.15x at 8 GPUs with every rank free to run anywhere.)  Returns a description for the bench line."""
    import os
    try:
        allowed = sorted(os.sched_getaffinity(0))
    except Exception:
        return {"bound": False, "why": "no sched_getaffinity"}
    local = allowed
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(vis.split(",")[local_rank]) if vis else local_rank
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        words = (max(allowed) // 64) + 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        near = [c for c in allowed if (mask[c // 64] >> (c % 64)) & 1]
        if near:
            local = near
    except Exception:
        pass
    mine = local_core_slice(local, local_rank, local_world)
    try:
        os.sched_setaffinity(0, mine)
    except Exception as e:
        return {"bound": False, "why": str(e)}
    return {"bound": True, "cores": len(mine), "first": mine[0], "last": mine[-1], "gpu_local_cores": len(local)}


class DeviceGroup:
    """include/b2rt_mg.h from Python: ONE process driving several GPUs through the C ABI (b2rt_mg_create /
    b2rt_mg_broadcast_bvh / b2rt_mg_trace / b2rt_mg_trace_host).  bench.py's N > 1 runs are one process per GPU under torchrun
    (the driver's contract) and use the functions above; this class is what a C++ host holding all devices calls."""

    def __init__(self, devices):
        import ctypes as C
        from ._lib import lib
        self._C, self._lib = C, lib()
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = self._lib.b2rt_mg_create(len(devices), devs, C.byref(h))
        if rc != 0:
            raise RuntimeError("b2rt_mg_create: %s" % self._lib.b2rt_mg_last_error(None).decode())
        self.handle, self.devices = h, list(devices)

    def close(self):
        if self.handle:
            self._lib.b2rt_mg_destroy(self.handle)
            self.handle = None

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError("b2rt_mg: %s" % self._lib.b2rt_mg_last_error(self.handle).decode())

    def broadcast_bvh(self, arrays):
        C = self._C
        n = np.ascontiguousarray(arrays["nodes"], np.float32)
        w = np.ascontiguousarray(arrays["woop"], np.float32)
        t = np.ascontiguousarray(arrays["triIndex"], np.int32)
        self._check(self._lib.b2rt_mg_broadcast_bvh(self.handle, n.ctypes.data_as(C.c_void_p), n.nbytes, w.ctypes.data_as(C.c_void_p), w.nbytes,
                                                    t.ctypes.data_as(C.c_void_p), t.nbytes, int(arrays.get("layout", 5)), 0))

    def trace(self, rays, results, gathered=None, any_hit=False):
        """rays[r], results[r]: torch tensors on device r ([n_r, 8] float32, [n_r, 4] int32); gathered[r]: [sum n_r, 4] int32 or None."""
        from ._lib import MgStats, TRACE_ANY_HIT
        C = self._C
        k = len(self.devices)
        rp = (C.c_void_p * k)(*[r.data_ptr() if r.numel() else None for r in rays])
        sp = (C.c_void_p * k)(*[r.data_ptr() if r.numel() else None for r in results])
        cnt = (C.c_int64 * k)(*[int(r.shape[0]) for r in rays])
        gp = (C.c_void_p * k)(*[g.data_ptr() for g in gathered]) if gathered is not None else None
        st = MgStats()
        self._check(self._lib.b2rt_mg_trace(self.handle, rp, sp, cnt, TRACE_ANY_HIT if any_hit else 0, gp, C.byref(st)))
        return st

    def trace_host(self, rays_host, results_host, any_hit=False):
        from ._lib import MgStats, TRACE_ANY_HIT
        C = self._C
        st = MgStats()
        self._check(self._lib.b2rt_mg_trace_host(self.handle, C.c_void_p(rays_host.ctypes.data), C.c_void_p(results_host.ctypes.data),
                                                 len(rays_host), TRACE_ANY_HIT if any_hit else 0, C.byref(st)))
        return st
