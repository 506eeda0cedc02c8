"""Stress test of the any-hit kernels' subtree sharing (trace.cu SHARE): C3's twelve 2 Mi-ray AO batches, many
repetitions per variant, result records poisoned before every launch so that an untraced ray shows.  Found the
non-uniform poolEmpty race of the first version (a crash once in a few thousand launches).
usage: python tools/stress_anyhit.py [reps] [variants, comma separated]"""
import importlib, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
variants = [int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "9,1").split(",")]
ctx = b2rt.Context.get(0)
scene, bvh = wl.build_scene_and_bvh("c3", log=print)
frame = wl.Frame("c3", scene, bvh.arrays(), "cuda:0", ctx)
reference = []
ctx.set_trace_config(variant=1)
for rb, _ in frame.batches:
    frame.tracer.traceBatch(rb)
    reference.append((rb.m_results[:rb.getSize(), 0] >= 0).clone())
bad = 0
for v in variants:
    ctx.set_trace_config(variant=v)
    for rep in range(reps):
        for i, (rb, _) in enumerate(frame.batches):
            rb.m_results.fill_(-2)
            frame.tracer.traceBatch(rb)
            ids = rb.m_results[:rb.getSize(), 0]
            unwritten = int((ids == -2).sum())
            differ = int(((ids >= 0) != reference[i]).sum())
            if unwritten or differ:
                bad += 1
                print("variant %d rep %d batch %d: %d unwritten, %d hit/miss differences" % (v, rep, i, unwritten, differ), flush=True)
    print("variant %d: %d x %d launches done" % (v, reps, len(frame.batches)), flush=True)
print("FAILED" if bad else "OK")
sys.exit(1 if bad else 0)
