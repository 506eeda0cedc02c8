"""GPU parity of the traversal kernel against the oracle, through the C ABI (host.py -> libb2rt.so).

Acceptance rule (BASELINE.json north_star): closest-hit rays -- >= 99.99 % identical
hit triangle id, t within 1e-5 relative where ids match, residual mismatches near-ties; any-hit
rays -- hit / no-hit per ray.  Misses return id = -1 and t = tmax (kepler_dynamic_fetch.cu:407).
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

T_RTOL = 1.0e-5
ID_AGREE = 0.9999


def _primary_rays(b2rt, oracle, scene, w, h):
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    i2p, _ = oracle.pixel_table(w, h)
    rays, _, _ = oracle.raygen_primary(origin, m, w, h, scene.camera["far"], i2p)
    return rays


def _check_closest(oracle, bvh_arrays, rays, got, ref=None, agree=ID_AGREE):
    if ref is None:
        ref = oracle.trace_woop(bvh_arrays, rays)
    same = got["id"] == ref["id"]
    assert same.mean() >= agree, "hit-id agreement %.6f" % same.mean()
    hit = same & (ref["id"] >= 0)
    rel = np.abs(got["t"][hit] - ref["t"][hit]) / np.abs(ref["t"][hit])
    assert rel.max(initial=0.0) <= T_RTOL, "max relative t error %g" % rel.max()
    miss = same & (ref["id"] < 0)
    assert (got["t"][miss] == rays["tmax"][miss]).all(), "a miss must return t = tmax"
    # residual mismatches must be near-ties: same t within tolerance
    bad = ~same
    if bad.any():
        both = bad & (got["id"] >= 0) & (ref["id"] >= 0)
        assert both.sum() == bad.sum(), "a mismatch is hit-vs-miss, not a tie between triangles"
        assert (np.abs(got["t"][bad] - ref["t"][bad]) <= 1e-4 * np.abs(ref["t"][bad])).all()
    return same.mean()


@pytest.mark.parametrize("kind,ntris", [("blob", 20000), ("interior", 30000), ("knot", 40000), ("strands", 30000)])
@pytest.mark.parametrize("variant", [1, 8, 9, 10, 11])
def test_primary_closest_hit(b2rt, oracle, ctx, kind, ntris, variant):
    scene = b2rt.Scene.generate(kind, ntris, 20260101)
    bvh = b2rt.BVH(scene)
    rays = _primary_rays(b2rt, oracle, scene, 256, 192)
    ctx.set_trace_config(variant=variant)
    tracer = b2rt.CudaTracer(ctx)
    tracer.writeBarycentrics = True
    tracer.setBVH(b2rt.CudaBVH(bvh))
    rb = b2rt.RayBuffer()
    rb.setRays(rays)
    ms = tracer.traceBatch(rb)
    assert ms > 0.0
    got = rb.getResults()
    ref = oracle.trace_woop(bvh.arrays(), rays)
    _check_closest(oracle, bvh.arrays(), rays, got, ref)
    # barycentrics agree with the oracle's on identical hits
    same = (got["id"] == ref["id"]) & (ref["id"] >= 0)
    du = np.abs(got["padA"].view(np.float32)[same] - ref["padA"].view(np.float32)[same])
    dv = np.abs(got["padB"].view(np.float32)[same] - ref["padB"].view(np.float32)[same])
    assert max(du.max(initial=0), dv.max(initial=0)) < 1e-3
    ctx.set_trace_config()


def test_any_hit_matches_hit_or_miss(b2rt, oracle, ctx):
    scene = b2rt.Scene.generate("interior", 30000, 3)
    bvh = b2rt.BVH(scene)
    rays = _primary_rays(b2rt, oracle, scene, 256, 192)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    rb = b2rt.RayBuffer(closestHit=False)
    rb.setRays(rays)
    rb.setNeedClosestHit(False)
    tracer.traceBatch(rb)
    got = rb.getResults()
    ref = oracle.trace_woop(bvh.arrays(), rays, any_hit=False)
    assert ((got["id"] >= 0) == (ref["id"] >= 0)).mean() >= ID_AGREE


def test_empty_and_degenerate_rays(b2rt, oracle, ctx):
    scene = b2rt.Scene.generate("blob", 5000, 5)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    rb = b2rt.RayBuffer()
    assert tracer.traceBatch(rb) == 0.0                       # CudaTracer.cc:123-125
    rays = _primary_rays(b2rt, oracle, scene, 64, 48).copy()
    rays["tmax"][::3] = -1.0                                  # degenerate: tmax < tmin (RayGenKernels.cu:223)
    rays["d"][1::7, 0] = 0.0                                  # axis-parallel components
    rb.setRays(rays)
    tracer.traceBatch(rb)
    got = rb.getResults()
    assert (got["id"][::3] == -1).all() and (got["t"][::3] == -1.0).all()
    _check_closest(oracle, bvh.arrays(), rays, got)


def test_no_bvh_and_wrong_layout_fail(b2rt, ctx):
    tracer = b2rt.CudaTracer(b2rt.Context(0))
    rb = b2rt.RayBuffer(4)
    with pytest.raises(RuntimeError, match="No BVH"):
        tracer.traceBatch(rb)


def test_reference_launcher_abi(b2rt, oracle, ctx):
    """libb2rt_refabi.so: bind_CudaBVHTexture + launch_tracingKernel + launch_countHitsKernel with the reference's
    argument lists (CudaTracerKernels.hh:42-52, RendererKernels.hh:88-92) give the same hit records as b2rt_trace."""
    import ctypes as C
    import os
    import torch
    lib = C.CDLL(os.path.join(os.path.dirname(b2rt.LIB_PATH), "libb2rt_refabi.so"))
    lib.launch_tracingKernel.restype = C.c_float
    lib.launch_tracingKernel.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_bool] + [C.c_void_p] * 10
    lib.bind_CudaBVHTexture.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
    scene = b2rt.Scene.generate("interior", 20000, 21)
    bvh = b2rt.BVH(scene)
    cb = b2rt.CudaBVH(bvh)
    rays = _primary_rays(b2rt, oracle, scene, 128, 96)
    rb = b2rt.RayBuffer()
    rb.setRays(rays)
    n, w, t = cb.getNodeBuffer_dev(), cb.getTriWoopBuffer_dev(), cb.getTriIndexBuffer_dev()
    torch.cuda.set_device(0)
    lib.bind_CudaBVHTexture(n.data_ptr(), n.numel() * 4, w.data_ptr(), w.numel() * 4, t.data_ptr(), t.numel() * 4)
    res = torch.zeros((len(rays), 4), dtype=torch.int32, device="cuda")
    block = (C.c_int * 2)(32, 4)
    ms = lib.launch_tracingKernel(23040, block, len(rays), False, rb.m_rays.data_ptr(), res.data_ptr(), n.data_ptr(), n.data_ptr(),
                                  n.data_ptr(), n.data_ptr(), w.data_ptr(), w.data_ptr(), w.data_ptr(), t.data_ptr())
    assert ms > 0.0
    got = res.cpu().numpy().view(oracle.RESULT_DTYPE).reshape(-1)
    _check_closest(oracle, bvh.arrays(), rays, got)

    class CountHitsInput(C.Structure):
        _fields_ = [("numRays", C.c_int32), ("rayResults", C.c_void_p), ("raysPerThread", C.c_int32)]
    inp = CountHitsInput(len(rays), res.data_ptr(), 32)
    lib.launch_countHitsKernel.restype = C.c_int32
    hits = lib.launch_countHitsKernel(C.c_int32(0), block, C.byref(inp))
    assert hits == int((got["id"] >= 0).sum())


@pytest.mark.parametrize("variant", [108, 114])
@pytest.mark.parametrize("kind,ntris", [("interior", 30000), ("strands", 30000), ("blob", 20000)])
def test_wide_kernel_closest_hit(b2rt, oracle, ctx, kind, ntris, variant):
    """The 8-wide compressed private form (csrc/trace_wide.cu): same acceptance rule as the binary kernel, and --
    because both run the Woop test on the caller's rows -- bit-identical t, u, v wherever the ids agree."""
    scene = b2rt.Scene.generate(kind, ntris, 20260101)
    bvh = b2rt.BVH(scene)
    rays = _primary_rays(b2rt, oracle, scene, 256, 192).copy()
    rays["tmax"][5::97] = -1.0                                  # a few degenerate rays
    tracer = b2rt.CudaTracer(ctx)
    tracer.writeBarycentrics = True
    tracer.setBVH(b2rt.CudaBVH(bvh))
    rb = b2rt.RayBuffer()
    rb.setRays(rays)
    ctx.set_trace_config(variant=1)
    tracer.traceBatch(rb)
    binary = rb.getResults().copy()
    ctx.set_trace_config(variant=variant)
    ms = tracer.traceBatch(rb)
    got = rb.getResults().copy()
    ctx.set_trace_config()
    assert ms > 0.0
    _check_closest(oracle, bvh.arrays(), rays, got)
    same = got["id"] == binary["id"]
    assert same.mean() >= ID_AGREE
    hit = same & (got["id"] >= 0)
    assert (got["t"][hit] == binary["t"][hit]).all() and (got["padA"][hit] == binary["padA"][hit]).all()
    assert (got["id"][5::97] == -1).all() and (got["t"][5::97] == -1.0).all()


@pytest.mark.parametrize("n", [1, 31, 127, 128, 129, 4097, 100003])
def test_block_classification_ragged_batches_and_runs_of_degenerate_rays(b2rt, oracle, ctx, n):
    """Variant 114 (trace_wide.cu, SCAN): a warp claims 128 rays, answers the degenerate ones (tmax < tmin,
    RayGenKernels.cu:223) with all lanes at once and hands out the others from a list.  Batches that are not a multiple of
    the block, runs of degenerate rays as an AO / diffuse batch has them (all samples of a primary ray that hit nothing) --
    shorter than, equal to and longer than a block -- any-hit and closest-hit: every record written, degenerate rays get
    (-1, tmax), the others what the binary kernel gives, bit for bit wherever the ids agree."""
    scene = b2rt.Scene.generate("knot", 40000, 7)
    bvh = b2rt.BVH(scene)
    base = _primary_rays(b2rt, oracle, scene, 400, 300)
    rng = np.random.default_rng(n)
    rays = base[rng.integers(0, len(base), n)].copy()
    dead = np.zeros(n, bool)
    pos = 0
    while pos < n:                                              # runs: live 1..200 rays, then dead 1..300 rays
        pos += int(rng.integers(1, 200))
        run = int(rng.integers(1, 300))
        dead[pos:pos + run] = True
        pos += run
    if n > 300:
        dead[128:384] = True                                    # two whole blocks of misses: a claim that finds nothing to hand out
    rays["tmax"][dead] = rays["tmin"][dead] - 0.25
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    for closest in (True, False):
        rb = b2rt.RayBuffer(closestHit=closest)
        rb.setRays(rays)
        rb.setNeedClosestHit(closest)
        out = {}
        for variant in (1, 114):
            rb.getResultBuffer_dev().fill_(0x7E7E7E7E)          # a record nobody writes would keep this
            ctx.set_trace_config(variant=variant)
            try:
                tracer.traceBatch(rb)
            finally:
                ctx.set_trace_config()
            out[variant] = rb.getResults().copy()
        got, want = out[114], out[1]
        assert (got["id"] != 0x7E7E7E7E).all(), "%d records were never written" % int((got["id"] == 0x7E7E7E7E).sum())
        assert (got["id"][dead] == -1).all() and (got["t"][dead] == rays["tmax"][dead]).all()
        live = ~dead
        if closest:
            same = got["id"] == want["id"]
            assert same[live].mean() >= ID_AGREE if live.sum() >= 20000 else (~same[live]).sum() <= 2
            assert (got["t"][same] == want["t"][same]).all()
        else:
            assert ((got["id"] >= 0) == (want["id"] >= 0)).all()


@pytest.mark.parametrize("variant", [1, 9])
@pytest.mark.parametrize("kind,ntris", [("interior", 30000), ("knot", 40000)])
@pytest.mark.parametrize("count", [1, 37, 4097, 0])
def test_any_hit_subtree_sharing(b2rt, oracle, ctx, kind, ntris, variant, count):
    """The any-hit kernels hand stack entries to idle lanes once the work counter is dry (trace.cu SHARE) and merge
    the helpers' hits into the result record with an atomicMin; small batches run almost entirely in that mode.
    Hit / no-hit must equal the oracle's for every ray, a reported hit must be a real one (t inside the ray's
    interval, not in front of the closest hit), and with barycentrics on, u and v must be the values the plain
    closest-hit kernel computes for the same triangle."""
    scene = b2rt.Scene.generate(kind, ntris, 7)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.writeBarycentrics = True
    tracer.setBVH(b2rt.CudaBVH(bvh))
    origin, m = b2rt.camera_matrices(scene.camera, 128, 96)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, 128, 96, scene.camera["far"])
    tracer.traceBatch(prim)
    sec = b2rt.RayBuffer()
    rg.ao(sec, prim, scene, 4, 2.0 * float(np.linalg.norm(scene.verts.max(0) - scene.verts.min(0))) / 10.0, True)
    rays = sec.getRays().copy()
    if count:
        rays = rays[np.linspace(0, len(rays) - 1, count).astype(np.int64)].copy()
    rb = b2rt.RayBuffer(closestHit=False)
    rb.setRays(rays)
    rb.setNeedClosestHit(False)
    try:
        ctx.set_trace_config(variant=variant)
        tracer.traceBatch(rb)
        got = rb.getResults().copy()
    finally:
        ctx.set_trace_config()
    ref = oracle.trace_woop(bvh.arrays(), rays, any_hit=False)          # closest hit: the strongest statement about hit / miss
    agree = (got["id"] >= 0) == (ref["id"] >= 0)
    assert agree.mean() >= (ID_AGREE if len(rays) > 10000 else 0.97), "hit/miss agreement %.6f" % agree.mean()
    hit = (got["id"] >= 0) & (ref["id"] >= 0)
    assert (got["t"][hit] > rays["tmin"][hit]).all() and (got["t"][hit] < rays["tmax"][hit]).all()
    assert (got["t"][hit] >= ref["t"][hit] * (1 - 1e-5)).all()
    miss = got["id"] < 0
    assert (got["t"][miss] == rays["tmax"][miss]).all()
    same = hit & (got["id"] == ref["id"])
    if same.any():
        rel = np.abs(got["t"][same] - ref["t"][same]) / np.abs(ref["t"][same])
        assert rel.max() <= T_RTOL
        # barycentrics of the merged record come from the fix-up pass: same bits as the closest-hit kernel's
        rb.setNeedClosestHit(True)
        tracer.traceBatch(rb)
        closest = rb.getResults()
        both = same & (closest["id"] == got["id"]) & (closest["t"] == got["t"])
        assert both.sum() >= 0.9 * same.sum()
        assert (closest["padA"][both] == got["padA"][both]).all() and (closest["padB"][both] == got["padB"][both]).all()


def test_wide_kernel_any_hit_and_secondary(b2rt, oracle, ctx):
    scene = b2rt.Scene.generate("interior", 30000, 3)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    origin, m = b2rt.camera_matrices(scene.camera, 256, 192)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, 256, 192, scene.camera["far"])
    ctx.set_trace_config(variant=108)
    try:
        tracer.traceBatch(prim)
        sec = b2rt.RayBuffer()
        rg.ao(sec, prim, scene, 4, 1.5, True)                   # any hit
        tracer.traceBatch(sec)
        rays, got = sec.getRays(), sec.getResults()
        ref = oracle.trace_woop(bvh.arrays(), rays, any_hit=False)
        assert ((got["id"] >= 0) == (ref["id"] >= 0)).mean() >= ID_AGREE
        hit = got["id"] >= 0
        assert (got["t"][hit] > 0).all() and (got["t"][hit] < 1.5).all()
        sec.setNeedClosestHit(True)                             # the same rays as a diffuse (closest-hit) batch
        tracer.traceBatch(sec)
        _check_closest(oracle, bvh.arrays(), rays, sec.getResults(), ref)
    finally:
        ctx.set_trace_config()


def _host_batches(b2rt, oracle, ctx, scene, tracer, sizes):
    """Primary-ray batches of the given pixel sizes, as (device RayBuffer traced on the device, pinned host rays)."""
    import torch
    out = []
    for (w, h) in sizes:
        rays = _primary_rays(b2rt, oracle, scene, w, h)
        rb = b2rt.RayBuffer()
        rb.setRays(rays)
        tracer.traceBatch(rb)
        hr = torch.from_numpy(rays.view(np.float32).reshape(-1, 8).copy()).pin_memory()
        out.append((rb, hr))
    return out


def test_trace_host_matches_device_path(b2rt, oracle, ctx):
    """b2rt_trace_host (host rays in, host hit records out; chunked upload / trace / download pipeline) returns the
    records b2rt_trace leaves on the device, pads zeroed; pageable buffers work too; an empty batch is a no-op."""
    import torch
    scene = b2rt.Scene.generate("interior", 30000, 5)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    (rb, hr), = _host_batches(b2rt, oracle, ctx, scene, tracer, [(1024, 600)])      # 614 400 rays: two chunks
    dev = rb.getResults()
    for pinned in (True, False):
        rays_h = hr if pinned else hr.clone()
        res_h = torch.full((hr.shape[0], 4), 7, dtype=torch.int32)
        if pinned:
            res_h = res_h.pin_memory()
        k, t = tracer.traceHost(rays_h, res_h)
        assert 0.0 < k <= t
        got = res_h.numpy().view(oracle.RESULT_DTYPE).reshape(-1)
        assert (got["id"] == dev["id"]).mean() >= ID_AGREE
        same = got["id"] == dev["id"]
        assert (got["t"][same] == dev["t"][same]).all()
        assert (got["padA"] == 0).all() and (got["padB"] == 0).all()
    k, t = tracer.traceHost(hr[:0], torch.empty((0, 4), dtype=torch.int32))
    assert k == 0.0 and t == 0.0


def test_trace_host_begin_end_keeps_batches_apart(b2rt, oracle, ctx):
    """Three batches begun back to back (the third has to wait for the first slot), one wait: every batch gets its own
    records, identical to the device path's."""
    import torch
    scene = b2rt.Scene.generate("knot", 40000, 5)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    batches = _host_batches(b2rt, oracle, ctx, scene, tracer, [(640, 480), (203, 77), (800, 700)])
    outs = [torch.full((hr.shape[0], 4), -5, dtype=torch.int32).pin_memory() for _, hr in batches]
    for (rb, hr), res_h in zip(batches, outs):
        tracer.traceHostBegin(hr, res_h)
    k, t = tracer.traceHostEnd()
    assert 0.0 < k <= t
    for (rb, hr), res_h in zip(batches, outs):
        dev = rb.getResults()
        got = res_h.numpy().view(oracle.RESULT_DTYPE).reshape(-1)
        same = got["id"] == dev["id"]
        assert same.mean() >= ID_AGREE and (got["t"][same] == dev["t"][same]).all()
    assert tracer.traceHostEnd() == (0.0, 0.0)          # nothing in flight


def test_traces_on_two_streams_overlap_safely(b2rt, oracle, ctx):
    """Launch slots (one work counter and one shared-ray list slice per launch): asynchronous traces issued on different
    streams give each batch the records the synchronous path gives it -- closest-hit and any-hit, several rounds."""
    import torch
    scene = b2rt.Scene.generate("knot", 40000, 9)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    origin, m = b2rt.camera_matrices(scene.camera, 320, 240)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, 320, 240, scene.camera["far"])
    tracer.traceBatch(prim)
    batches = []
    for samples, closest in ((4, True), (8, False), (2, True), (16, False), (1, True), (6, False)):
        sec = b2rt.RayBuffer()
        rg.ao(sec, prim, scene, samples, 3.0, True)
        sec.setNeedClosestHit(closest)
        tracer.traceBatch(sec)
        batches.append((sec, sec.getResults().copy()))
    streams = [torch.cuda.Stream() for _ in range(3)]
    for _ in range(5):
        for rb, _ in batches:
            rb.m_results.fill_(-7)
        torch.cuda.synchronize()
        for k, (rb, _) in enumerate(batches):
            with torch.cuda.stream(streams[k % 3]):
                tracer.traceBatchAsync(rb)
        torch.cuda.synchronize()
        for rb, want in batches:
            got = rb.getResults()
            assert ((got["id"] >= 0) == (want["id"] >= 0)).all()
            if rb.getNeedClosestHit():
                same = got["id"] == want["id"]
                assert same.mean() >= 0.999 and (got["t"][same] == want["t"][same]).all()


@pytest.mark.parametrize("pair", [(1, 10), (1, 11), (10, 11)])
def test_lookahead_leaf_loop_returns_the_same_bits(b2rt, oracle, ctx, pair):
    """LEAF = 1 / 2 (row 0 of the next triangle requested ahead, Woop test without early-outs; 2 = that loop unrolled twice:
    variants 11 and 10; trace.cu) performs the same
    floating-point operations in the same order as the reference-shaped loop: wherever both kernels report the same
    triangle, t, u and v must be identical bit for bit, and the ids may differ only between triangles at one distance."""
    scene = b2rt.Scene.generate("interior", 30000, 7)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.writeBarycentrics = True
    tracer.setBVH(b2rt.CudaBVH(bvh))
    origin, m = b2rt.camera_matrices(scene.camera, 320, 240)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, 320, 240, scene.camera["far"])
    tracer.traceBatch(prim)
    sec = b2rt.RayBuffer()
    rg.ao(sec, prim, scene, 2, scene.camera["far"], True)
    sec.setNeedClosestHit(True)
    out = []
    try:
        for variant in pair:
            ctx.set_trace_config(variant=variant)
            per = []
            for rb in (prim, sec):
                tracer.traceBatch(rb)
                per.append(rb.getResults().copy())
            out.append(per)
    finally:
        ctx.set_trace_config()
    for a, b in zip(out[0], out[1]):
        same = a["id"] == b["id"]
        assert same.mean() >= ID_AGREE
        assert a[same].tobytes() == b[same].tobytes()
        assert (a["t"][~same] == b["t"][~same]).all()


def test_launch_attributes_survive_bvhs_of_different_sizes_and_a_second_context(b2rt, oracle, ctx):
    """The binary kernel's dynamic shared memory depends on the BVH (staged top of the tree), and the limit a kernel may use is
    a per-process property of (device, kernel): big BVH, small BVH, a SECOND context on the same device with a small BVH, then
    the big one again -- every launch must go through (a cache keyed on the size once lowered the limit: 'invalid argument')."""
    big = b2rt.Scene.generate("interior", 30000, 3)
    small = b2rt.Scene(np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 1, 0.5]], np.float32), np.array([[0, 1, 2], [1, 3, 2]], np.int32))
    big_bvh, small_bvh = b2rt.CudaBVH(b2rt.BVH(big)), b2rt.CudaBVH(b2rt.BVH(small))
    rays = _primary_rays(b2rt, oracle, big, 64, 48)
    other = b2rt.Context(0)
    want = None
    for which, c in (("big", ctx), ("small", ctx), ("small", other), ("big", ctx), ("big", other), ("small", ctx), ("big", ctx)):
        tracer = b2rt.CudaTracer(c)
        tracer.setBVH(big_bvh if which == "big" else small_bvh)
        rb = b2rt.RayBuffer()
        rb.setRays(rays)
        c.set_trace_config(variant=8)
        try:
            tracer.traceBatch(rb)
        finally:
            c.set_trace_config()
        if which == "big":
            got = rb.getResults().copy()
            if want is None:
                want = got
            assert (got["id"] == want["id"]).mean() >= 0.9999
