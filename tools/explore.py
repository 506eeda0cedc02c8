"""Tuning / comparison harness (not the bench): times the traversal kernel's variants and the
reference's CUDA tracer (oracle/_ref/libref_gpu.so) on a config's batches, checks parity between
them.  Usage: python tools/explore.py c2 [variants...]"""
import importlib
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")
import oracle_lib as O


def median_ms(fn, warm=3, reps=7):
    for _ in range(warm):
        fn()
    return float(np.median([fn() for _ in range(reps)]))


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    configs = [a for a in sys.argv[2:] if not a.startswith("--")]
    sort_batches = "--sorted" in sys.argv      # Morton-sort every secondary batch first (RayBuffer::mortonSort), untimed
    ctx = b2rt.Context.get(0)
    scene, bvh = wl.build_scene_and_bvh(name, log=lambda s: print(s, flush=True))
    shard = (0, 1)
    for a in sys.argv:
        if a.startswith("--shard="):           # --shard=r,w : rank r's share of a sharded workload (c5), traced on this one GPU
            shard = tuple(int(x) for x in a.split("=")[1].split(","))
    frame = wl.Frame(name, scene, bvh.arrays(), "cuda:0", ctx, shard=shard)
    counted = frame.counted_rays()
    if sort_batches:
        for rb, _ in frame.batches:
            if rb is not frame.primary:
                rb.mortonSort(ctx)
        print("batches Morton-sorted", flush=True)
    print("batches", [(b.getSize(), c) for b, c in frame.batches], flush=True)
    for rb, _ in frame.batches[:2]:
        work = frame.tracer.countWork(rb)
        n = work["rays"]
        print("work/ray: nodes %.1f tris %.1f leaves %.1f hits %.3f maxstack %d longest ray %d steps" % (
            work["nodes"] / n, work["tris"] / n, work["leaves"] / n, work["hits"] / n, work["max_stack"], work["max_chain"]))
    out = {}
    # reference CUDA tracer
    if O.have_ref_gpu():
        ref = O.RefGpuTracer()
        ref.set_bvh(frame.cuda_bvh.m_nodes, frame.cuda_bvh.m_triWoop, frame.cuda_bvh.m_triIndex)
        sm = ctx.sm_count()
        for label, warps in (("ref720", 720), ("ref_scaled40", sm * 40), ("ref_scaled32", sm * 32), ("ref_scaled48", sm * 48)):
            res = [torch.zeros_like(b.m_results) for b, _ in frame.batches]
            ms = median_ms(lambda: sum(ref.trace(b.getRayBuffer_dev(), r, not b.getNeedClosestHit(), warps)
                                       for (b, _), r in zip(frame.batches, res)))
            out[label] = counted / ms / 1e3
            print("%-16s %8.3f ms  %9.1f Mrays/s" % (label, ms, out[label]), flush=True)
        ref_res = res
    sweeps = configs or ["v1", "v6", "v7", "v100", "v101", "v102", "v103"]
    for s in sweeps:
        kw = {}
        overlap = 0
        for tok in s.split(","):
            if tok.startswith("v"): kw["variant"] = int(tok[1:])
            if tok.startswith("t"): kw["top_nodes"] = int(tok[1:])
            if tok.startswith("f"): kw["fetch_threshold"] = int(tok[1:])
            if tok.startswith("b"): kw["blocks_per_sm"] = int(tok[1:])
            if tok.startswith("x"): kw["tune"] = int(tok[1:])
            if tok.startswith("o"): overlap = int(tok[1:])
        ctx.set_trace_config(**kw)
        if kw.get("tune", 0) & 4 or getattr(main, "_was_greedy", False):
            frame.tracer.setBVH(frame.cuda_bvh)         # the collapse mode takes effect at the next setBVH
            main._was_greedy = bool(kw.get("tune", 0) & 4)
        ms = median_ms((lambda: frame.trace_all_overlapped(overlap)) if overlap else frame.trace_all)
        out[s] = counted / ms / 1e3
        print("%-16s %8.3f ms  %9.1f Mrays/s" % (s, ms, out[s]), flush=True)
        if True:
            # SIMD efficiency and launch timeline, per batch (instrumented kernel, same configuration)
            frame.tracer.collectStats = True
            for rb, _ in frame.batches[:2]:
                frame.tracer.traceBatch(rb)
                st = frame.tracer.getStats()
                if st["warps"] == 0:        # the wide kernel keeps the SIMD counters only
                    print("    stats: TRV %.1f%% IST %.1f%% fetches %d" % (100.0 * st["node_lanes"] / max(1, 32 * st["node_iters"]),
                                                                         100.0 * st["tri_lanes"] / max(1, 32 * st["tri_iters"]), st["fetches"]))
                    continue
                print("    stats: TRV %.1f%% IST %.1f%% fetches %d | warps %d pool empty %.1f us, warp exit mean %.1f p50 %.1f p90 %.1f p99 %.1f last %.1f us"
                      " | last warp: last fetch at %.1f us, then %d node + %d triangle iterations | shares %d" % (
                    100.0 * st["node_lanes"] / max(1, 32 * st["node_iters"]), 100.0 * st["tri_lanes"] / max(1, 32 * st["tri_iters"]), st["fetches"],
                    st["warps"], st["ns_pool_empty"] / 1e3, st["ns_mean_exit"] / 1e3, st["ns_exit_p50"] / 1e3, st["ns_exit_p90"] / 1e3,
                    st["ns_exit_p99"] / 1e3, st["ns_last_exit"] / 1e3, st["ns_straggler_fetch"] / 1e3, st["straggler_node_iters"], st["straggler_tri_iters"], st["shares"]))
            frame.tracer.collectStats = False
    ctx.set_trace_config()
    frame.trace_all()
    if O.have_ref_gpu():
        # parity with the reference CUDA tracer on identical rays + BVH
        for (b, _), r in zip(frame.batches, ref_res):
            mine = b.getResults(); theirs = r[: b.getSize()].cpu().numpy().view(O.RESULT_DTYPE).reshape(-1)
            if b.getNeedClosestHit():
                same = mine["id"] == theirs["id"]
                hit = same & (mine["id"] >= 0)
                rel = np.abs(mine["t"][hit] - theirs["t"][hit]) / np.abs(theirs["t"][hit])
                print("parity vs ref GPU: id agree %.6f, max rel t %.2e" % (same.mean(), rel.max(initial=0)))
            else:
                print("parity vs ref GPU (any-hit): hit/miss agree %.6f" % ((mine["id"] >= 0) == (theirs["id"] >= 0)).mean())
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    with open(os.path.join(ROOT, "gpurun_out", "explore_%s.json" % name), "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
