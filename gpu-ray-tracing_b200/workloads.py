"""The benchmark configurations of BASELINE.json as seeded synthetic workloads.

  c1  144 500-tri blob,      1024x768 primary rays                      (the CPU-runnable case)
  c2  350 949-tri interior,  1024x768 primary + 1-bounce diffuse        (the 1-GPU headline)
  c3  910 348-tri knot,      AO 32 rays/pixel, any-hit, Morton-sorted 2^21-ray batches
  c4  6 469 561-tri strands, primary + incoherent diffuse (divergence stress)
  c5  910 348-tri knot,      one 64 Mi-ray diffuse batch, BVH replicated, rays sharded over ranks
Seeds are 20260101 + config number.  Everything here calls the product's public interface
(host.py -> libb2rt.so); nothing touches oracle/.
"""
import os
import time

import numpy as np
import torch

from .host import BVH, CudaBVH, CudaTracer, Context, RayBuffer, RayGen, Scene, camera_matrices, countHits

CONFIGS = {
    "c1": dict(kind="blob", tris=144500, seed=20260102, size=(1024, 768), ray_type="primary", samples=1, sort=False),
    "c2": dict(kind="interior", tris=350949, seed=20260103, size=(1024, 768), ray_type="diffuse", samples=1, sort=False,
               with_primary=True),
    "c3": dict(kind="knot", tris=910348, seed=20260104, size=(1024, 768), ray_type="ao", samples=32, sort=True),
    "c4": dict(kind="strands", tris=6469561, seed=20260105, size=(1024, 768), ray_type="diffuse", samples=1, sort=False),
    "c5": dict(kind="knot", tris=910348, seed=20260104, size=(1024, 768), ray_type="diffuse", samples=86, sort=False,
               total_rays=64 * 1024 * 1024),
    # small stand-in used by smoke() and the tests
    "tiny": dict(kind="interior", tris=20000, seed=7, size=(160, 120), ray_type="diffuse", samples=2, sort=False),
}

CACHE_DIR = os.environ.get("B2RT_BVH_CACHE", "/tmp/b2rt_bvhcache")


def describe(name):
    c = CONFIGS[name]
    return "%s: procedural %s %d tris, %dx%d %s%s%s" % (
        name, c["kind"], c["tris"], c["size"][0], c["size"][1], c["ray_type"],
        " x%d" % c["samples"] if c["ray_type"] != "primary" else "", ", Morton-sorted" if c["sort"] else "")


def cache_path(scene, split_alpha=1.0e-5, min_leaf=1, max_leaf=8, cache_dir=None):
    """File a reference build would use for this scene's CudaBVH: <dir>/%08x.dat with the key of Renderer::getCudaBVH
    (b2rt_bvhcache_key: hashes of the scene's arrays, the platform, the build parameters and the layout)."""
    import ctypes as C
    from ._lib import LAYOUT_COMPACT2, check, lib, np_ptr
    key = C.c_uint32()
    check(lib().b2rt_bvhcache_key(np_ptr(scene.verts), len(scene.verts), np_ptr(scene.tris), len(scene.tris), None,
                                  C.c_float(split_alpha), int(min_leaf), int(max_leaf), LAYOUT_COMPACT2, C.byref(key)))
    buf = C.create_string_buffer(1024)
    check(lib().b2rt_bvhcache_path((cache_dir or CACHE_DIR).encode(), key.value, buf, 1024))
    return buf.value.decode()


def describe_step(name):
    """What one bench step traces (identical text in both arms of bench.py)."""
    c = CONFIGS[name]
    if c.get("total_rays"):
        return "closest-hit trace of one %d-ray diffuse batch (%d samples per primary hit; degenerate rays of primary misses launched, not counted)" % (
            c["total_rays"], c["samples"])
    if c["ray_type"] == "primary":
        return "closest-hit trace of the frame's primary batch"
    kind = "closest-hit" if c["ray_type"] == "diffuse" else "any-hit"
    return "%s%s trace of the frame's %s batches (%d samples per primary hit)%s" % (
        "closest-hit trace of the primary batch + " if c.get("with_primary") else "", kind, c["ray_type"], c["samples"],
        ", each Morton-sorted first (untimed)" if c["sort"] else "")


def build_scene_and_bvh(name, use_cache=True, log=None):
    """Scene + host BVH for a config; the BVH goes through a bvhcache file (Renderer.cc:176-211)."""
    c = CONFIGS[name]
    t0 = time.time()
    scene = Scene.generate(c["kind"], c["tris"], c["seed"])
    t1 = time.time()
    path = cache_path(scene)                # bvhcache/%08x.dat, the reference's naming (Renderer.cc:176-186)
    bvh = None
    if use_cache and os.path.exists(path):
        try:
            bvh = BVH.load(path, scene)
        except Exception:
            bvh = None
    if bvh is None:
        bvh = BVH(scene)
        if use_cache:
            try:
                os.makedirs(CACHE_DIR, exist_ok=True)
                bvh.serialize(path)
            except Exception:
                pass
    if log:
        log("scene %s: %d tris in %.2fs, BVH (%d inner nodes, %d Woop rows) in %.2fs" % (
            scene.name, scene.getNumTriangles(), t1 - t0, bvh.nodes.size // 16, bvh.woop.size // 4, time.time() - t1))
    return scene, bvh


class Frame:
    """One frame of a config on one device: primary rays traced, then the list of secondary
    batches (as Renderer::beginFrame / nextBatch produce them, Renderer.cc:112-291), all resident
    in HBM.  `batches` holds (RayBuffer, counted_rays) pairs ready to be traced and timed."""

    def __init__(self, name, scene, bvh_arrays, device, ctx=None, shard=(0, 1), max_batch=1 << 21):
        c = CONFIGS[name]
        self.name, self.cfg, self.scene = name, c, scene
        self.device = torch.device(device)
        self.ctx = ctx or Context.get(self.device.index)
        self.tracer = CudaTracer(self.ctx)
        self.cuda_bvh = CudaBVH(bvh_arrays, device=self.device)
        self.tracer.setBVH(self.cuda_bvh)
        self.raygen = RayGen(max_batch, self.ctx)
        w, h = c["size"]
        origin, m = camera_matrices(scene.camera, w, h)
        self.primary = RayBuffer(0, True, self.device)
        self.raygen.primary(self.primary, origin, m, w, h, scene.camera["far"])
        self.batches = []
        rank, ws = shard
        if c["ray_type"] == "primary":
            self.batches.append((self.primary, w * h))
            return
        self.tracer.traceBatch(self.primary)
        if c.get("with_primary"):
            # "primary + 1-bounce diffuse" (BASELINE.json configs[1]): the primary batch is a timed batch too
            self.batches.append((self.primary, w * h))
        hits = countHits(self.primary.getResultBuffer_dev(), self.ctx)
        samples = c["samples"]
        far = scene.camera["far"]
        if c["ray_type"] == "ao":
            lo, hi = scene.verts.min(0), scene.verts.max(0)
            max_dist = 0.05 * float(np.linalg.norm(hi - lo))        # aoRadius = 5 % of the scene diagonal
        else:
            max_dist = far
        total = c.get("total_rays")
        if total:
            # C5: ONE buffer of `total` diffuse rays (primaries x samples).  Sharded over ranks in interleaved chunks of
            # 64 primaries = one 8x8 pixel block (contiguous slices of an image-ordered buffer have unequal cost -- sky vs
            # geometry -- so each rank takes every ws-th chunk); the last rank's slice is cut so the total is exact.
            n_prim_needed = (total + samples - 1) // samples
            big = RayGen(1 << 30, self.ctx)
            chunk = 64
            idx = torch.arange(n_prim_needed, device=self.device)
            mine = idx[((idx // chunk) % ws) == rank]
            sub = _gather_buffer(self.primary, mine)
            sec = RayBuffer(0, True, self.device)
            big.ao(sec, sub, scene, samples, max_dist, True)
            sec.setNeedClosestHit(True)
            excess = n_prim_needed * samples - total
            if excess and mine.numel() and int(mine[-1]) == n_prim_needed - 1:
                sec = _slice_buffer(sec, 0, sec.getSize() - excess, closest=True)
            counted = int((sec.getRayBuffer_dev()[:, 7] >= 0).sum().item())
            self.batches.append((sec, counted))
            return
        newBatch = True
        while True:
            sec = RayBuffer(0, True, self.device)
            cont, newBatch = self.raygen.ao(sec, self.primary, scene, samples, max_dist, newBatch)
            if not cont:
                break
            sec.setNeedClosestHit(c["ray_type"] == "diffuse")
            if c["sort"]:
                sec.mortonSort(self.ctx)
            counted = int((sec.getRayBuffer_dev()[:, 7] >= 0).sum().item())     # degenerate rays have tmax = -1
            self.batches.append((sec, counted))
        assert sum(b[1] for b in self.batches if b[0] is not self.primary) == hits * samples   # Renderer.cc:221-238

    def counted_rays(self):
        return sum(b[1] for b in self.batches)

    def launched_rays(self):
        return sum(b[0].getSize() for b in self.batches)

    def trace_all(self):
        """Trace every batch once; returns the summed kernel milliseconds (App.cc:194-204)."""
        return sum(self.tracer.traceBatch(rb) for rb, _ in self.batches)

    def trace_all_overlapped(self, n_streams=2):
        """The frame's batches issued round-robin on `n_streams` streams without waiting in between: batches that do not
        depend on each other (the secondary batches of a frame all depend on the primary batch only) may overlap, so the
        idle tail of one launch is filled by the next.  Returns the device milliseconds from the first launch's start to
        the last one's end.  The primary batch, when it is part of the frame, is traced first and waited for."""
        if not hasattr(self, "_streams") or len(self._streams) != n_streams:
            self._streams = [torch.cuda.Stream(device=self.device) for _ in range(n_streams)]
        cur = torch.cuda.current_stream(self.device)
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        start.record(cur)
        batches = list(self.batches)
        if batches and batches[0][0] is self.primary:
            self.tracer.traceBatchAsync(self.primary)           # on the current stream; the others wait for `ready`
            batches = batches[1:]
        ready = torch.cuda.Event()
        ready.record(cur)
        for k, (rb, _) in enumerate(batches):
            s = self._streams[k % n_streams]
            if k < n_streams:
                s.wait_event(ready)
            with torch.cuda.stream(s):
                self.tracer.traceBatchAsync(rb)
        for s in self._streams:
            cur.wait_stream(s)
        end.record(cur)
        end.synchronize()
        return start.elapsed_time(end)


def _gather_buffer(rb, index):
    """RayBuffer holding the rays and results of the given slots (identity id maps)."""
    out = RayBuffer(0, rb.getNeedClosestHit(), rb.device)
    n = int(index.numel())
    out.resize(n)
    if n:
        out.m_rays[:n].copy_(rb.m_rays[index])
        out.m_results[:n].copy_(rb.m_results[index])
        ar = torch.arange(n, dtype=torch.int32, device=rb.device)
        out.m_IDToSlot[:n].copy_(ar)
        out.m_slotToID[:n].copy_(ar)
    return out


def _slice_buffer(rb, lo, hi, closest=None):
    out = RayBuffer(0, rb.getNeedClosestHit() if closest is None else closest, rb.device)
    n = hi - lo
    out.resize(n)
    out.m_rays[:n].copy_(rb.m_rays[lo:hi])
    out.m_results[:n].copy_(rb.m_results[lo:hi])
    ar = torch.arange(n, dtype=torch.int32, device=rb.device)
    out.m_IDToSlot[:n].copy_(ar)
    out.m_slotToID[:n].copy_(ar)
    return out


def render_frame_e2e(name, scene, cuda_bvh, ctx, device, steps=3, shard=(0, 1), n_streams=2):
    """The reference's real frame, end to end (Renderer::beginFrame / nextBatch / traceBatch / reconstruct,
    Renderer.cc:112-152,242-300): camera -> primary rays generated on the device -> closest-hit trace -> hit count ->
    diffuse batches of at most 2 Mi rays generated on the device from the primary hits -> closest-hit trace -> reconstruct
    -> the frame's pixels copied to the host.  Nothing but the camera goes up; pixels and one counter come down.
    With shard = (rank, world) this rank renders every world-th 8x8 pixel block.  Wall clock, device idle at both ends.
    The secondary batches depend on the primary batch only (the reference runs them one after the other because it has one
    ray buffer and one work counter, Renderer.cc:242-291, kepler_dynamic_fetch.cu:58): with n_streams = 2 they alternate
    between two ray buffers on two streams, so that one batch's ray generation and the head of its trace run under the tail
    of the batch before it; n_streams = 1 is the reference's order."""
    from .host import reconstruct
    c = CONFIGS[name]
    w, h = c["size"]
    samples = c["samples"]
    device = torch.device(device)
    rank, ws = shard
    tracer = CudaTracer(ctx)
    tracer.setBVH(cuda_bvh)
    raygen = RayGen(1 << 21, ctx)                       # Renderer.cc:46
    origin, m = camera_matrices(scene.camera, w, h)
    far = scene.camera["far"]
    prim = RayBuffer(0, True, device)
    secs = [RayBuffer(0, True, device) for _ in range(max(1, n_streams))]
    streams = [torch.cuda.Stream(device=device) for _ in range(n_streams)] if n_streams > 1 else []
    ntri = scene.getNumTriangles()
    material = torch.full((ntri,), 0xFF808080, dtype=torch.int64, device=device).to(torch.int32)
    shaded = torch.full((ntri,), 0xFFC0C0C0, dtype=torch.int64, device=device).to(torch.int32)
    pixels = torch.zeros(w * h, dtype=torch.int32, device=device)
    host_pixels = torch.empty(w * h, dtype=torch.int32).pin_memory()
    mine = None
    if ws > 1:
        idx = torch.arange(w * h, device=device)
        mine = idx[((idx // 64) % ws) == rank]
    times, counted, launches = [], 0, 0
    for step in range(steps + 1):                        # the first pass warms up
        torch.cuda.synchronize()
        l0 = ctx.launch_count()
        t0 = time.perf_counter()
        raygen.primary(prim, origin, m, w, h, far)
        src = prim
        if mine is not None:
            src = _gather_buffer(prim, mine)
            src.m_slotToID[:src.getSize()].copy_(prim.m_slotToID[mine])
        tracer.traceBatchAsync(src)
        hits = countHits(src.getResultBuffer_dev(), ctx)                       # Renderer::getTotalNumRays (Renderer.cc:221-238)
        newBatch, first, k = True, 0, 0
        cur = torch.cuda.current_stream(device)
        for st in streams:
            st.wait_stream(cur)                                                 # (countHits above has synchronised already)
        while True:
            sec = secs[k % len(secs)]
            with torch.cuda.stream(streams[k % len(streams)] if streams else cur):
                cont, newBatch = raygen.ao(sec, src, scene, samples, far, newBatch)
                if not cont:
                    break
                sec.setNeedClosestHit(True)                                     # Renderer.cc:277
                tracer.traceBatchAsync(sec)
                n_prim = sec.getSize() // samples
                reconstruct("diffuse", samples, first, n_prim, src, sec, material, shaded, pixels, ctx)
            first += n_prim
            k += 1
        for st in streams:
            cur.wait_stream(st)
        host_pixels.copy_(pixels, non_blocking=True)
        torch.cuda.synchronize()
        if step:
            times.append((time.perf_counter() - t0) * 1e3)
            counted = hits * samples
            launches = ctx.launch_count() - l0
    return {"ms_per_step": float(np.median(times)), "counted": int(counted), "h2d": 19 * 4, "d2h": w * h * 4 // ws + 8, "launches": int(launches),
            "api": "RayGen::primary -> traceBatch -> countHits -> {RayGen::ao (diffuse, 2 Mi-ray batches) -> traceBatch -> reconstruct} -> pixels to the host; "
                   "wall clock per frame, median of %d; secondary batches alternate between %d stream(s)" % (len(times), max(1, n_streams)),
            "pixels": host_pixels.clone()}
