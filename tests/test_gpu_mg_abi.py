"""include/b2rt_mg.h through the C ABI: one process, min(2, #GPUs) devices -- BVH broadcast (built once on the first device, the
private forms ncclBroadcast into the others), sharded trace, gather of the hit records into every device, all-reduced hit
count, and the host-buffer variant.  Checked against the single-device entry point b2rt_trace on the same rays.  On a
one-GPU box the group has one member (NCCL with a single rank): the same code path, trivial collectives."""
import importlib
import os
import subprocess

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("kind,ntris,wide", [("interior", 30000, False), ("knot", 700000, True)])
def test_device_group_matches_single_device(b2rt, oracle, ctx, kind, ntris, wide):
    mgm = importlib.import_module("gpu-ray-tracing_b200.multigpu")
    ndev = min(2, torch.cuda.device_count())
    scene = b2rt.Scene.generate(kind, ntris, 20260104)
    bvh = b2rt.BVH(scene)
    arrays = bvh.arrays()
    assert (arrays["nodes"].nbytes + arrays["woop"].nbytes > (48 << 20)) == wide        # the knot takes the 8-wide form
    w, h = 512, 384
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    i2p, _ = oracle.pixel_table(w, h)
    rays, _, _ = oracle.raygen_primary(origin, m, w, h, scene.camera["far"], i2p)
    # single device, the existing entry point
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    rb = b2rt.RayBuffer()
    rb.setRays(rays)
    tracer.traceBatch(rb)
    want = rb.getResults().copy()
    group = mgm.DeviceGroup(list(range(ndev)))
    try:
        group.broadcast_bvh(arrays)
        # unequal shards, one of them ragged
        cut = [0, len(rays) * 3 // 5 + 7, len(rays)] if ndev == 2 else [0, len(rays)]
        flat = torch.from_numpy(rays.view(np.float32).reshape(-1, 8).copy())
        d_rays = [flat[cut[r]:cut[r + 1]].to("cuda:%d" % r) for r in range(ndev)]
        d_res = [torch.zeros((cut[r + 1] - cut[r], 4), dtype=torch.int32, device="cuda:%d" % r) for r in range(ndev)]
        d_all = [torch.full((len(rays), 4), -5, dtype=torch.int32, device="cuda:%d" % r) for r in range(ndev)]
        st = group.trace(d_rays, d_res, d_all)
        assert st.rays == len(rays) and st.hits == int((want["id"] >= 0).sum())
        assert st.ms_trace_max > 0 and st.ms_with_gather_max >= st.ms_trace_max
        for r in range(ndev):
            got = d_all[r].cpu().numpy().view(oracle.RESULT_DTYPE).reshape(-1)
            same = got["id"] == want["id"]
            assert same.mean() >= 0.9999, "device %d: id agreement %.6f" % (r, same.mean())
            assert (got["t"][same] == want["t"][same]).all()
        # host buffers in, host records out
        out = np.zeros(len(rays), oracle.RESULT_DTYPE)
        st2 = group.trace_host(rays, out)
        assert st2.hits == st.hits
        same = out["id"] == want["id"]
        assert same.mean() >= 0.9999 and (out["t"][same] == want["t"][same]).all()
        assert (out["padA"] == 0).all() and (out["padB"] == 0).all()
        # an empty shard on one device is fine
        if ndev == 2:
            e_rays = [flat.to("cuda:0"), flat[:0].to("cuda:1")]
            e_res = [torch.zeros((len(rays), 4), dtype=torch.int32, device="cuda:0"), torch.zeros((0, 4), dtype=torch.int32, device="cuda:1")]
            st3 = group.trace(e_rays, e_res, None)
            assert st3.hits == st.hits
    finally:
        group.close()


def test_cpp_caller_of_the_multi_gpu_abi(b2rt, oracle, ctx):
    """tests/cpp/test_mg.cpp: a C++ host driving the group through include/b2rt_mg.h (what SURVEY 8b asks a C++ caller to have)."""
    pkg = os.path.join(ROOT, "gpu-ray-tracing_b200")
    exe = os.path.join(ROOT, "tests", "cpp", "test_mg")
    src = os.path.join(ROOT, "tests", "cpp", "test_mg.cpp")
    cuda = "/usr/local/cuda"
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", cuda + "/include", src, "-o", exe,
           "-L", pkg, "-lb2rt", "-lb2rt_scenes", "-L", cuda + "/lib64", "-lcudart", "-Wl,-rpath," + pkg]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "test_mg ok" in r.stdout, r.stdout + r.stderr
    print(r.stdout.strip())
