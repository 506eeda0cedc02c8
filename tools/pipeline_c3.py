"""End-to-end rate of a sorted AO frame ("the sort only pays off end to end"): C3 = 910 348-triangle knot,
1024x768, 32 AO rays per pixel in 2 Mi-ray batches.  Per batch: rayGenAOKernel -> mortonSort -> any-hit trace, all on the
device, CUDA events around each stage; the same frame again without the sort.  Mrays/s = counted rays (Renderer.cc:221-238)
over the summed stage time.  usage: python tools/pipeline_c3.py -> gpurun_out/pipeline_c3.json"""
import importlib, json, os, sys
import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")


def main():
    ctx = b2rt.Context.get(0)
    scene, bvh = wl.build_scene_and_bvh("c3", log=print)
    tracer = b2rt.CudaTracer(ctx)
    tracer.setBVH(b2rt.CudaBVH(bvh))
    w, h, S = 1024, 768, 32
    origin, m = b2rt.camera_matrices(scene.camera, w, h)
    rg = b2rt.RayGen(1 << 21, ctx)
    prim = b2rt.RayBuffer()
    rg.primary(prim, origin, m, w, h, scene.camera["far"])
    tracer.traceBatch(prim)
    hits = b2rt.countHits(prim.getResultBuffer_dev(), ctx)
    counted = hits * S
    lo, hi = scene.verts.min(0), scene.verts.max(0)
    radius = 0.05 * float(np.linalg.norm(hi - lo))
    sec = b2rt.RayBuffer()

    def frame(sort):
        t = {"gen": 0.0, "sort": 0.0, "trace": 0.0}
        nb, batches = True, 0
        while True:
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            e[0].record()
            cont, nb = rg.ao(sec, prim, scene, S, radius, nb)
            if not cont:
                break
            e[1].record()
            if sort:
                sec.mortonSort(ctx)
            e[2].record()
            t["trace"] += tracer.traceBatch(sec)
            torch.cuda.synchronize()
            t["gen"] += e[0].elapsed_time(e[1]); t["sort"] += e[1].elapsed_time(e[2])
            batches += 1
        return t, batches

    out = {"counted_rays": counted, "launched_rays": w * h * S}
    for sort in (True, False):
        for _ in range(2):
            frame(sort)
        runs = [frame(sort) for _ in range(5)]
        med = {k: float(np.median([r[0][k] for r in runs])) for k in ("gen", "sort", "trace")}
        total = sum(med.values())
        key = "sorted" if sort else "unsorted"
        out[key] = {"ms": med, "ms_total": total, "batches": runs[0][1], "Mrays/s_pipeline": counted / total * 1e-3,
                    "Mrays/s_trace_only": counted / med["trace"] * 1e-3}
        print("%-8s gen %.3f ms  sort %.3f ms  trace %.3f ms  | pipeline %.1f Mrays/s, trace only %.1f Mrays/s (%d batches)" % (
            key, med["gen"], med["sort"], med["trace"], out[key]["Mrays/s_pipeline"], out[key]["Mrays/s_trace_only"], runs[0][1]), flush=True)
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "pipeline_c3.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
