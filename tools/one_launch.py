"""Builds a workload's frame and traces every batch a few times with the library defaults (or a variant / tune bits):
the program ncu is pointed at for a full capture of ONE launch of the traversal kernel.
usage: python tools/one_launch.py c5 [variant] [tune]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
b2rt = importlib.import_module("gpu-ray-tracing_b200")
wl = importlib.import_module("gpu-ray-tracing_b200.workloads")

name = sys.argv[1] if len(sys.argv) > 1 else "c5"
ctx = b2rt.Context.get(0)
if len(sys.argv) > 2:
    ctx.set_trace_config(variant=int(sys.argv[2]), tune=int(sys.argv[3]) if len(sys.argv) > 3 else 0)
scene, bvh = wl.build_scene_and_bvh(name)
frame = wl.Frame(name, scene, bvh.arrays(), "cuda:0", ctx)
for _ in range(3):
    ms = frame.trace_all()
print("%s: %d counted rays, %.3f ms per step, %.1f Mrays/s" % (name, frame.counted_rays(), ms, frame.counted_rays() / ms * 1e-3))
