"""GPU parity of ray generation, Morton sort, hit count and reconstruction against the oracle."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _scene_and_primary(b2rt, oracle, ctx, kind="interior", ntris=20000, w=200, h=152):
    scene = b2rt.Scene.generate(kind, ntris, 11)
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, w, h, scene.camera["far"])
    return scene, rg, prim, origin, m


@pytest.mark.parametrize("w,h", [(64, 48), (200, 152), (203, 77), (8, 8), (5, 3)])
def test_primary_rays_bit_exact(b2rt, oracle, ctx, w, h):
    scene = b2rt.Scene.generate("blob", 2000, 1)
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    rg = b2rt.RayGen(1 << 21, ctx)
    rb = b2rt.RayBuffer()
    rg.primary(rb, origin, m, w, h, scene.camera["far"])
    i2p, _ = oracle.pixel_table(w, h)
    assert (rg.m_pixelTable.indexToPixel == i2p).all()
    rays, id2slot, slot2id = oracle.raygen_primary(origin, m, w, h, scene.camera["far"], i2p)
    assert rb.getSize() == w * h and rb.getNeedClosestHit()
    assert rb.getRays().tobytes() == rays.tobytes()               # -fmad=false + IEEE div/sqrt: bit exact
    assert (rb.getSlotToIDBuffer_dev().cpu().numpy() == slot2id).all()
    assert (rb.getIDToSlotBuffer_dev().cpu().numpy() == id2slot).all()


@pytest.mark.parametrize("samples", [1, 8, 32])
def test_ao_rays(b2rt, oracle, ctx, samples):
    scene, rg, prim, _, _ = _scene_and_primary(b2rt, oracle, ctx)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    tracer.traceBatch(prim)
    in_rays, in_res = prim.getRays(), prim.getResults()
    sec = b2rt.RayBuffer()
    cont, nb = rg.ao(sec, prim, scene, samples, 5.0, True)
    assert cont and not nb and not sec.getNeedClosestHit()
    n_in = min(prim.getSize(), (1 << 21) // samples)
    ref, id2slot, slot2id = oracle.raygen_ao(0, n_in, samples, 5.0, rg.lastSeed, in_rays, in_res, scene.normals)
    got = sec.getRays()
    assert len(got) == len(ref)
    assert (got["o"] == ref["o"]).all() and (got["tmin"] == ref["tmin"]).all() and (got["tmax"] == ref["tmax"]).all()
    # directions pass through sin/cos: CUDA sincosf vs glibc, tolerance 2e-6 absolute on unit vectors
    assert np.abs(got["d"] - ref["d"]).max() <= 2e-6
    assert (sec.getSlotToIDBuffer_dev().cpu().numpy() == slot2id).all()
    miss = np.repeat(in_res["id"][:n_in] == -1, samples)
    assert (got["tmax"][miss] == -1.0).all() and (got["tmax"][~miss] == 5.0).all()


def test_ao_batching_follows_reference(b2rt, oracle, ctx):
    """RayGen::batching (RayGen.cc:124-142): slices of maxBatchSize/numSamples primaries, libc rand() seeds."""
    scene, _, prim, _, _ = _scene_and_primary(b2rt, oracle, ctx, w=64, h=48)
    prim.m_results.zero_()
    rg = b2rt.RayGen(4096, ctx)
    sec = b2rt.RayBuffer()
    sizes, seeds, newBatch = [], [], True
    while True:
        cont, newBatch = rg.ao(sec, prim, scene, 3, 1.0, newBatch)
        if not cont:
            break
        sizes.append(sec.getSize())
        seeds.append(rg.lastSeed)
    per = 4096 // 3
    n = 64 * 48
    assert sizes == [min(per, n - i) * 3 for i in range(0, n, per)]
    assert seeds[:3] == [1804289383, 846930886, 1681692777]       # glibc rand() after srand(1)


def test_find_aabb_and_keys_bit_exact(b2rt, oracle, ctx):
    scene, rg, prim, _, _ = _scene_and_primary(b2rt, oracle, ctx)
    rays = prim.getRays().copy()
    rng = np.random.default_rng(5)
    rays["o"] += rng.normal(0, 1, rays["o"].shape).astype(np.float32)
    rays["tmax"][::5] = -1.0
    rb = b2rt.RayBuffer()
    rb.setRays(rays)
    lo, hi = b2rt.findAABB(rb.getRayBuffer_dev(), ctx)
    rlo, rhi = oracle.find_aabb(rays)
    assert (lo == rlo).all() and (hi == rhi).all()
    keys = b2rt.mortonKeys(rb.getRayBuffer_dev(), lo, hi, ctx)
    ref = oracle.morton_keys(rays, rlo, rhi)
    assert keys.tobytes() == ref.tobytes()


@pytest.mark.parametrize("n", [1, 31, 4096, 4097, 50000, 300000])
def test_morton_sort(b2rt, oracle, ctx, n):
    rng = np.random.default_rng(n)
    rays = np.zeros(n, b2rt.RAY_DTYPE)
    rays["o"] = rng.uniform(-3, 3, (n, 3)).astype(np.float32)
    # clusters of identical origins (AO rays share their origin) and some exact duplicates
    rays["o"][: n // 2] = rays["o"][: n // 2][rng.integers(0, max(n // 64, 1), n // 2)]
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["tmax"] = rng.uniform(0.1, 4, n).astype(np.float32)
    if n > 100:
        rays[n // 3: n // 3 + 20] = rays[0]
    ids = rng.permutation(n).astype(np.int32)
    rb = b2rt.RayBuffer()
    rb.setRays(rays, ids)
    rb.mortonSort(ctx)
    out = rb.getRays()
    slot2id = rb.getSlotToIDBuffer_dev().cpu().numpy()
    id2slot = rb.getIDToSlotBuffer_dev().cpu().numpy()
    # permutation bookkeeping (reorderRaysKernel, RayBufferKernels.cu:183-200)
    assert sorted(slot2id.tolist()) == list(range(n))
    assert (id2slot[slot2id] == np.arange(n)).all()
    inv = np.empty(n, np.int64)
    inv[ids] = np.arange(n)
    assert out.tobytes() == rays[inv[slot2id]].tobytes()
    # key sequence == the oracle's sorted key sequence (order inside runs of equal keys is free)
    lo, hi = oracle.find_aabb(rays)
    got_keys = oracle.morton_keys(out, lo, hi)["hash"]
    ref_keys = oracle.sort_keys(oracle.morton_keys(rays, lo, hi))["hash"]
    assert got_keys.tobytes() == ref_keys.tobytes()


def _sorted_like_the_oracle(b2rt, oracle, ctx, rays, mode):
    """Sorts `rays` (identity ids) on the device with the given sort path; returns (fell_back, device rays, device
    slotToID, oracle rays, oracle slotToID).  The oracle sorts stably and so do both device paths, so the permutation
    itself must agree, not only the key sequence."""
    rb = b2rt.RayBuffer()
    rb.setRays(rays, np.arange(len(rays), dtype=np.int32))
    ctx.set_sort_mode(mode)
    try:
        rb.mortonSort(ctx)
        fell_back = ctx.sort_fell_back()
    finally:
        ctx.set_sort_mode(0)
    ref_rays, _, ref_slot2id = oracle.morton_sort(rays)[:3]
    return fell_back, rb.getRays(), rb.getSlotToIDBuffer_dev().cpu().numpy(), ref_rays, ref_slot2id


@pytest.mark.parametrize("mode", [0, 1])
def test_morton_sort_settles_prefix_ties_by_the_full_key(b2rt, oracle, ctx, mode):
    """Rays that agree in the 56 key bits the first stage sorts by (same origin cell, directions 1e-5 apart) and differ
    below them, mixed with exact duplicates (runs of 40 <= kMaxRun) and ordinary rays: both sort paths must return the
    oracle's stable order."""
    rng = np.random.default_rng(77)
    n = 20000
    rays = np.zeros(n, b2rt.RAY_DTYPE)
    rays["o"] = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    # groups of 12 rays: one origin, directions inside a 1e-5 cone
    for g in range(600):
        lo = 12 * g
        rays["o"][lo:lo + 12] = rays["o"][lo]
        d[lo:lo + 12] = d[lo] + rng.normal(0, 1e-5, (12, 3)).astype(np.float32)
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["tmax"] = rng.uniform(0.1, 1.0, n).astype(np.float32)
    rays[9000:9040] = rays[8999]                    # 41 identical rays: one run, ordered by slot
    fell_back, got, slot2id, ref, ref_slot2id = _sorted_like_the_oracle(b2rt, oracle, ctx, rays, mode)
    assert fell_back == (mode == 1)
    assert got.tobytes() == ref.tobytes()
    assert np.array_equal(slot2id, ref_slot2id)


def test_morton_sort_falls_back_on_long_runs(b2rt, oracle, ctx):
    """3 000 identical rays form one run of equal prefixes, longer than a thread settles: the call takes the full
    192-bit sort by itself and still returns the oracle's order."""
    rng = np.random.default_rng(78)
    n = 30000
    rays = np.zeros(n, b2rt.RAY_DTYPE)
    rays["o"] = rng.uniform(-2, 2, (n, 3)).astype(np.float32)
    d = rng.normal(0, 1, (n, 3)).astype(np.float32)
    rays["d"] = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays["tmax"] = 1.0
    rays[5000:8000] = rays[4999]
    fell_back, got, slot2id, ref, ref_slot2id = _sorted_like_the_oracle(b2rt, oracle, ctx, rays, 0)
    assert fell_back
    assert got.tobytes() == ref.tobytes()
    assert np.array_equal(slot2id, ref_slot2id)


def test_count_hits_and_reconstruct(b2rt, oracle, ctx):
    scene, rg, prim, _, _ = _scene_and_primary(b2rt, oracle, ctx)
    bvh = b2rt.BVH(scene)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    tracer.traceBatch(prim)
    res = prim.getResults()
    assert b2rt.countHits(prim.getResultBuffer_dev(), ctx) == oracle.count_hits(res) == int((res["id"] >= 0).sum())
    assert b2rt.countHits(prim.getResultBuffer_dev()[:0], ctx) == 0
    nt = scene.getNumTriangles()
    rng = np.random.default_rng(0)
    mat = rng.integers(0, 2**32, nt, dtype=np.uint64).astype(np.uint32)
    shd = rng.integers(0, 2**32, nt, dtype=np.uint64).astype(np.uint32)
    dev = prim.device
    mat_d = torch.from_numpy(mat.view(np.int32)).to(dev)
    shd_d = torch.from_numpy(shd.view(np.int32)).to(dev)
    w, h = 200, 152
    # primary
    pix = torch.zeros(w * h, dtype=torch.int32, device=dev)
    b2rt.reconstruct("primary", 1, 0, w * h, prim, prim, mat_d, shd_d, pix, ctx)
    ref = np.zeros(w * h, np.uint32)
    s2i = prim.getSlotToIDBuffer_dev().cpu().numpy()
    i2s = prim.getIDToSlotBuffer_dev().cpu().numpy()
    oracle.reconstruct(1, 0, w * h, "primary", s2i, res, i2s, res, mat, shd, ref)
    assert (pix.cpu().numpy().view(np.uint32) == ref).all()
    # diffuse with 4 samples
    sec = b2rt.RayBuffer()
    rg.ao(sec, prim, scene, 4, scene.camera["far"], True)
    sec.setNeedClosestHit(True)
    tracer.traceBatch(sec)
    sres = sec.getResults()
    for mode in ("ao", "diffuse"):
        pix.zero_()
        b2rt.reconstruct(mode, 4, 0, w * h, prim, sec, mat_d, shd_d, pix, ctx)
        ref[:] = 0
        oracle.reconstruct(4, 0, w * h, mode, s2i, res, sec.getIDToSlotBuffer_dev().cpu().numpy(), sres, mat, shd, ref)
        assert (pix.cpu().numpy().view(np.uint32) == ref).all()


@pytest.mark.gpu
@pytest.mark.parametrize("kind,target", [("blob", 20000), ("interior", 30000), ("strands", 30000), ("knot", 40000)])
def test_device_woopify_is_bit_identical(b2rt, oracle, ctx, kind, target):
    """the Woop rows computed on the device (b2rt_woopify) equal the host-built ones bit for bit
    -- and therefore the oracle's woopifyTri restatement, which pins the host builder (test_host_logic)."""
    import torch
    scene = b2rt.Scene.generate(kind, target, 20260101)
    bvh = b2rt.BVH(scene)
    a = bvh.arrays()
    dev = b2rt.CudaBVH.fromSceneOnDevice(scene, {"nodes": a["nodes"], "triIndex": a["triIndex"]}, ctx)
    got = dev.getTriWoopBuffer_dev().cpu().numpy().view(np.uint32)
    ref = np.ascontiguousarray(a["woop"], np.float32).reshape(-1).view(np.uint32)
    assert got.shape == ref.shape
    assert (got == ref).all(), "%d of %d words differ" % ((got != ref).sum(), ref.size)
    # and the tracer accepts it like any other CudaBVH
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(dev)
    origin, m = b2rt.camera_matrices(scene.camera, 64, 48)
    rg = b2rt.RayGen(1 << 20, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, 64, 48, scene.camera["far"])
    tracer.traceBatch(prim)
    ref_hits = oracle.trace_woop(a, prim.getRays())
    assert (prim.getResults()["id"] == ref_hits["id"]).mean() >= 0.999
