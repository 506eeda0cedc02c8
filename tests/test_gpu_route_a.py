"""Drop-in route A, proved by linking it: oracle/_ref/route_a is the reference's UNCHANGED
CudaTracer.cc, RayGen.cc, RayBuffer.cc, PixelTable.cc, CudaBVH.cc, SplitBVHBuilder.cc ... (and its own fetch_*.cu dump kernels)
linked against libb2rt_refabi.so + libb2rt.so instead of the reference's four kernel files, with the real CUDA runtime.
It runs the reference's own call sequence -- RayGen::primary, CudaTracer::traceBatch (rt/cuda/CudaTracer.cc:119-178),
RayGen::ao, RayBuffer::mortonSort (RayBuffer.cc:256-324), traceBatch again -- on the B200 and dumps what it saw; this test
traces the same rays through b2rt_trace with the same (reference-built) BVH and compares, and checks the sort the
reference's host code produced from the new key / reorder kernels."""
import importlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "route_a")


def test_reference_host_classes_run_on_the_new_kernels(b2rt, oracle, ctx, tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/route_a was not built (needs /root/reference at build time)")
    sys.path.insert(0, ROOT)
    import bench
    # C2's interior: CudaTracer.cc's hard-coded triangle dump needs >= 750 000 Woop rows (route_a_driver.cc)
    obj, sig, camera, cfg = bench._workload_obj("c2", "/tmp/b2rt_bench")
    w, h, samples = 512, 384, 2
    prefix = str(tmp_path / "ra")
    r = subprocess.run([EXE, "--obj", obj, "--camera-sig", sig, "--size", str(w), str(h), "--samples", str(samples), "--out", prefix],
                       cwd=str(tmp_path), capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-800:] + r.stderr[-800:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["primary_rays"] == w * h and line["primary_ms"] > 0 and line["secondary_ms"] > 0
    # the owner's dump code ran too (CudaTracer.cc:181-637 with the reference's fetch_tri.cu): twelve hex files per traceBatch
    assert os.path.exists(tmp_path / "v00_x.txt") or any(f.endswith(".txt") for f in os.listdir(tmp_path))

    arrays = oracle.read_bvh(prefix + ".bvh")                      # the REFERENCE's SplitBVHBuilder tree, CudaBVH::serialize
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(arrays))
    for tag in ("prim", "sec"):
        rays = oracle.read_rays("%s.%s.rays" % (prefix, tag))
        theirs = oracle.read_results("%s.%s.results" % (prefix, tag))
        rb = b2rt.RayBuffer()
        rb.setRays(rays)
        tracer.traceBatch(rb)
        mine = rb.getResults()
        same = mine["id"] == theirs["id"]
        assert same.mean() >= 0.9999, "%s: id agreement %.6f between CudaTracer::traceBatch (route A) and b2rt_trace" % (tag, same.mean())
        assert (mine["t"][same] == theirs["t"][same]).all()
        print("route A %s batch: %d rays, id agreement %.6f, t bit-identical on equal ids" % (tag, len(rays), same.mean()))
    # primary rays are the reference generator's (RayGen::primary on rayGenPrimaryKernel): bit-exact against the oracle's restatement
    cam = b2rt.decode_camera_signature(sig)
    origin, m = b2rt.camera_matrices(cam, w, h)
    i2p, _ = oracle.pixel_table(w, h)
    want, _, _ = oracle.raygen_primary(origin, m, w, h, cam["far"], i2p)
    assert oracle.read_rays(prefix + ".prim.rays").tobytes() == want.tobytes()
    # RayBuffer::mortonSort (reference host code: AABB kernel, key kernel, D2H, FW::sort, H2D, reorder kernel):
    # the sorted buffer is a permutation of the unsorted one by slotToID, and its key sequence is non-decreasing
    unsorted = oracle.read_rays(prefix + ".sec.unsorted.rays")
    srt = oracle.read_rays(prefix + ".sec.rays")
    s2i, i2s = oracle.read_ints(prefix + ".sec.slot2id"), oracle.read_ints(prefix + ".sec.id2slot")
    n = len(srt)
    assert np.array_equal(np.sort(s2i), np.arange(n)) and np.array_equal(i2s[s2i], np.arange(n))
    assert srt.tobytes() == unsorted[s2i].tobytes()
    lo, hi = oracle.find_aabb(unsorted)
    keys = oracle.morton_keys(srt, lo, hi)["hash"].astype(np.uint64)
    ok, decided = np.ones(n - 1, bool), np.zeros(n - 1, bool)
    for wd in range(5, -1, -1):
        a, b = keys[:-1, wd], keys[1:, wd]
        ok &= decided | (a <= b)
        decided |= a != b
    assert ok.all()
