"""Times the kernels either side of the trace on C2/C3-sized inputs and sets each against
the HBM roofline with the algorithmic bytes DESIGN.md section 4 lists.  CUDA events on the launching stream, 5 warm-ups,
median of 20, a 256 MiB write between repetitions so that nothing is served from L2.
usage: python tools/bench_aux.py  ->  one line per kernel + gpurun_out/bench_aux.json"""
import importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")


def main():
    dev = torch.device("cuda:0")
    ctx = b2rt.Context.get(0)
    peak = 6450.3
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def timed(fn, reps=20, warm=5):
        out = []
        for i in range(warm + reps):
            flush.fill_(i & 0xff)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record()
            torch.cuda.synchronize()
            if i >= warm:
                out.append(e0.elapsed_time(e1))
        return float(np.median(out))

    scene, bvh = wl.build_scene_and_bvh("c2", log=print)
    arrays = bvh.arrays()
    w, h = 1024, 768
    n = w * h
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rows = []

    def report(name, ms, nbytes, unit_count, unit):
        gbs = nbytes / (ms * 1e-3) * 1e-9
        rows.append({"kernel": name, "ms": ms, "algorithmic_bytes": nbytes, "GB/s": gbs, "frac_of_hbm_peak": gbs / peak,
                     "units": unit_count, "unit": unit, "Munits/s": unit_count / ms * 1e-3})
        print("%-34s %8.3f ms  %8.1f MB  %7.1f GB/s  %5.1f %% of %.0f GB/s   %9.1f M%s/s" % (
            name, ms, nbytes / 1e6, gbs, 100 * gbs / peak, peak, unit_count / ms * 1e-3, unit), flush=True)

    # a11: 4 B LUT read + 32 B ray + 2 x 4 B id maps written per ray
    ms = timed(lambda: rg.primary(prim, origin, m, w, h, scene.camera["far"]))
    report("rayGenPrimaryKernel", ms, n * (4 + 32 + 8), n, "rays")
    tracer.traceBatch(prim)

    # a12: per primary 32 B ray + 16 B result + 12 B normal read, S x (32 B ray + 8 B maps) written
    S = 32
    sec = b2rt.RayBuffer()
    per_batch = (1 << 21) // S
    def gen_ao():
        rg.ao(sec, prim, scene, S, 1.0, True)
    ms = timed(gen_ao)
    report("rayGenAOKernel (32 spp batch)", ms, per_batch * (32 + 16 + 12) + per_batch * S * 40, per_batch * S, "rays")

    # a13: the whole Morton sort of a 2 Mi-ray batch: AABB (32 B read), keys (32 B read, 28 B written), 20 radix passes of
    # (8 B word + 4 B slot) read and written twice per pass (histogram read + scatter read/write), gather + reorder (68 + 40 B)
    nsec = sec.getSize()
    rays0 = sec.m_rays[:nsec].clone(); i2s0 = sec.m_IDToSlot[:nsec].clone(); s2i0 = sec.m_slotToID[:nsec].clone()
    def sort():
        sec.m_rays[:nsec].copy_(rays0); sec.m_IDToSlot[:nsec].copy_(i2s0); sec.m_slotToID[:nsec].copy_(s2i0)
        sec.mortonSort(ctx)
    def restore_only():
        sec.m_rays[:nsec].copy_(rays0); sec.m_IDToSlot[:nsec].copy_(i2s0); sec.m_slotToID[:nsec].copy_(s2i0)
    t_restore = timed(restore_only)
    # default path: 7 prefix passes + tie settling (3 x 8 B gathered per ray in a run); mode 1: the full 19-20-pass LSD sort
    ms = timed(sort) - t_restore
    fell = ctx.sort_fell_back()
    report("mortonSort (AABB+keys+prefix radix+ties+reorder)%s" % (" FELL BACK" if fell else ""), ms,
           nsec * (32 + 32 + 36 + 7 * (12 + 24) + 24 + 12 + 68 + 40), nsec, "rays")
    for shift in (83, 99):
        ctx.set_sort_mode(shift)
        ms = timed(sort) - t_restore
        report("mortonSort, prefix shift %d%s" % (shift, " FELL BACK" if ctx.sort_fell_back() else ""), ms,
               nsec * (32 + 32 + 36 + ((147 - shift + 7) // 8) * (12 + 24) + 24 + 12 + 68 + 40), nsec, "rays")
    ctx.set_sort_mode(1)
    ms = timed(sort) - t_restore
    ctx.set_sort_mode(0)
    report("mortonSort, full 192-bit LSD (mode 1)", ms, nsec * (32 + 32 + 28 + 20 * (12 + 24) + 68 + 40), nsec, "rays")

    # a14 / f1
    ms = timed(lambda: b2rt.countHits(prim.getResultBuffer_dev(), ctx))
    report("countHitsKernel", ms, n * 16, n, "rays")
    ntri = len(scene.tris)
    mat = torch.full((ntri,), 0xFF8080FF, dtype=torch.int64, device=dev).to(torch.int32)
    pix = torch.zeros(n, dtype=torch.int32, device=dev)
    ms = timed(lambda: b2rt.reconstruct("primary", 1, 0, n, prim, prim, mat, mat, pix, ctx))
    report("reconstructKernel (primary)", ms, n * (4 + 16 + 4 + 4 + 4), n, "pixels")

    # a8 / f3: Woop transform on the device: 44 B read (3 indices + 3 x 12 B vertices, gathered) + 48 B written per reference
    rows_h, tris_h = b2rt.leaf_references(arrays["nodes"], arrays["triIndex"])
    verts_d = torch.from_numpy(np.ascontiguousarray(scene.verts, np.float32).reshape(-1).copy()).to(dev)
    index_d = torch.from_numpy(np.ascontiguousarray(scene.tris, np.int32).reshape(-1).copy()).to(dev)
    rows_d, tris_d = torch.from_numpy(rows_h).to(dev), torch.from_numpy(tris_h).to(dev)
    nrows = int(np.asarray(arrays["triIndex"]).size)
    woop = torch.empty(nrows * 4, dtype=torch.float32, device=dev)
    import ctypes as C
    def woopify():
        b2rt._lib.check(b2rt.lib().b2rt_woopify(ctx.handle, C.c_void_p(verts_d.data_ptr()), C.c_void_p(index_d.data_ptr()),
                                                C.c_void_p(rows_d.data_ptr()), C.c_void_p(tris_d.data_ptr()), len(rows_h),
                                                C.c_void_p(woop.data_ptr()), nrows, None), ctx.handle)
    ms = timed(woopify)
    report("woopify (fill + transform)", ms, len(rows_h) * (8 + 12 + 36 + 48) + nrows * 16, len(rows_h), "refs")

    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump({"hbm_peak_gbs": peak, "kernels": rows}, open(os.path.join(ROOT, "gpurun_out", "bench_aux.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
