"""What one traversal step costs when nothing hides its latency: traces single-warp batches (32 consecutive rays of a
config's batches) with the instrumented kernel and prints nanoseconds per warp-level loop iteration -- the quantity that
sets the length of a launch's tail.  Usage: python tools/lone_warp.py c2 [variant ...]"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "c2"
    variants = [int(v) for v in sys.argv[2:]] or [8, 1]
    ctx = b2rt.Context.get(0)
    scene, bvh = wl.build_scene_and_bvh(name, log=lambda s: print(s, flush=True))
    frame = wl.Frame(name, scene, bvh.arrays(), "cuda:0", ctx)
    frame.trace_all()                                   # BVH into L2
    rng = np.random.RandomState(5)
    for bi, (rb, _) in enumerate(frame.batches[:2]):
        starts = [int(s) & ~31 for s in rng.randint(0, rb.getSize() - 32, 24)]
        for v in variants:
            ctx.set_trace_config(variant=v)
            frame.tracer.collectStats = True
            rows = []
            for s in starts:
                one = wl._slice_buffer(rb, s, s + 32)
                frame.tracer.traceBatch(one)            # warm: this warp's nodes into L2 / this launch's code into the I-cache
                frame.tracer.traceBatch(one)
                st = frame.tracer.getStats()
                iters = st["node_iters"] + st["tri_iters"]
                rows.append((st["ns_last_exit"], st["node_iters"], st["tri_iters"], iters))
            frame.tracer.collectStats = False
            ns = np.array([r[0] for r in rows], float)
            it = np.array([r[3] for r in rows], float)
            nn = np.array([r[1] for r in rows], float)
            nt = np.array([r[2] for r in rows], float)
            # least squares: ns = a + b_node * node_iters + b_tri * tri_iters
            A = np.stack([np.ones_like(nn), nn, nt], 1)
            coef, *_ = np.linalg.lstsq(A, ns, rcond=None)
            print("batch %d variant %3d: %d single-warp launches, mean %.0f node + %.0f triangle iterations, %.1f us;"
                  " %.0f ns per iteration overall; fit: %.1f us + %.0f ns/node step + %.0f ns/triangle step" % (
                      bi, v, len(rows), nn.mean(), nt.mean(), ns.mean() / 1e3, ns.sum() / it.sum(), coef[0] / 1e3, coef[1], coef[2]),
                  flush=True)
    ctx.set_trace_config()


if __name__ == "__main__":
    main()
