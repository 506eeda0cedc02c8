"""Builds gpu-ray-tracing_b200/libb2rt.so in-tree with nvcc for sm_100a (no JIT cache, no fallback).

    python gpu-ray-tracing_b200/build.py [--force] [--verbose]

nvcc cross-compiles without a GPU, so this runs in the CPU-only container as well as on the
B200 box (where the prebuilt .so that travelled with the snapshot is normally used as is).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb2rt.so")
SCENES_LIB = os.path.join(HERE, "libb2rt_scenes.so")      # benchmark scene generators (include/b2rt_scenes.h): host only, not part of libb2rt
REFABI_LIB = os.path.join(HERE, "libb2rt_refabi.so")      # the reference's launcher ABI on top of libb2rt (include/b2rt_refabi.h)
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_COMMON = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-I", INCLUDE]

# (source, extra flags).  -fmad=false: values that feed parity tests bit-for-bit (ray generation,
# Morton keys, pixel reconstruction) must not depend on FMA contraction; the traversal kernel
# spells out every FMA itself.
CUDA_SOURCES = [
    ("trace.cu", ["-Xptxas", "-v"]),
    ("trace_wide.cu", ["-Xptxas", "-v"]),
    ("api.cu", []),
    ("raygen.cu", ["-fmad=false"]),
    ("raysort.cu", ["-fmad=false"]),
    ("misc.cu", ["-fmad=false"]),
    ("woopify.cu", ["-fmad=false"]),
    ("static_scratch.cu", []),
    ("multi_gpu.cu", []),
]
HOST_SOURCES = ["host/bvh_builder.cpp", "host/host_util.cpp", "host/wide_bvh.cpp", "host/bvhcache_key.cpp"]
HOST_FLAGS = ["-O2", "-ffp-contract=off", "-fPIC", "-std=c++17", "-pthread", "-Wall", "-I", INCLUDE]


def _newer(target, deps):
    if not os.path.exists(target):
        return False
    t = os.path.getmtime(target)
    return all(os.path.getmtime(d) <= t for d in deps)


def _run(cmd, verbose):
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build step failed: " + " ".join(cmd))
    return r.stdout + r.stderr


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "wide_bvh.h"), os.path.join(INCLUDE, "b2rt.h"),
               os.path.join(INCLUDE, "b2rt_tools.h"), os.path.join(INCLUDE, "b2rt_mg.h"), os.path.abspath(__file__)]
    objs = []
    log = []
    for src, extra in CUDA_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace("/", "_") + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            log.append(_run([NVCC] + NVCC_COMMON + extra + ["-c", s, "-o", o], verbose))
    for src in HOST_SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src.replace("/", "_") + ".o")
        objs.append(o)
        if force or not _newer(o, [s] + headers):
            log.append(_run([CXX] + HOST_FLAGS + ["-c", s, "-o", o], verbose))
    if force or not _newer(LIB, objs):
        log.append(_run([NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-lpthread", "-ldl"], verbose))
    scenes_src = os.path.join(CSRC, "host", "scenegen.cpp")
    if force or not _newer(SCENES_LIB, [scenes_src, os.path.join(INCLUDE, "b2rt_scenes.h"), os.path.abspath(__file__)]):
        log.append(_run([CXX] + HOST_FLAGS + ["-shared", scenes_src, "-o", SCENES_LIB], verbose))
    refabi_src = os.path.join(CSRC, "host", "refabi.cpp")
    if force or not _newer(REFABI_LIB, [refabi_src, LIB, os.path.join(INCLUDE, "b2rt_refabi.h")] + headers):
        cuda_home = os.path.dirname(os.path.dirname(NVCC))
        log.append(_run([CXX] + HOST_FLAGS + ["-I", os.path.join(cuda_home, "include"), "-shared", refabi_src, "-o", REFABI_LIB,
                         "-L", HERE, "-lb2rt", "-L", os.path.join(cuda_home, "lib64"), "-lcudart", "-Wl,-rpath,$ORIGIN"], verbose))
    with open(os.path.join(OBJ, "build.log"), "a" if not force else "w") as f:
        f.write("".join(log))
    return LIB


if __name__ == "__main__":
    lib = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(lib)
