"""The C++ host mirror (include/b2rt_fw.hpp) compiles and links against libb2rt.so everywhere;
on a B200 the reference-style call sequence of tests/cpp/test_fw.cpp runs and is checked against the oracle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "gpu-ray-tracing_b200")
EXE = os.path.join(ROOT, "tests", "cpp", "test_fw")


def _build(b2rt, oracle):
    src = os.path.join(ROOT, "tests", "cpp", "test_fw.cpp")
    deps = [src, os.path.join(ROOT, "include", "b2rt_fw.hpp"), os.path.join(ROOT, "include", "b2rt.h")]
    if os.path.exists(EXE) and all(os.path.getmtime(d) <= os.path.getmtime(EXE) for d in deps):
        return EXE
    cuda = "/usr/local/cuda"
    cmd = ["g++", "-O2", "-std=c++17", "-Wall", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "oracle"),
           "-I", cuda + "/include", src, "-o", EXE, "-L", PKG, "-lb2rt", "-lb2rt_scenes", "-L", os.path.join(ROOT, "oracle"), "-lb2rt_oracle",
           "-L", cuda + "/lib64", "-lcudart", "-Wl,-rpath," + PKG, "-Wl,-rpath," + os.path.join(ROOT, "oracle")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    return EXE


def test_cpp_shim_compiles_and_links(b2rt, oracle):
    assert os.path.exists(_build(b2rt, oracle))


@pytest.mark.gpu
def test_cpp_shim_parity(b2rt, oracle, ctx):
    exe = _build(b2rt, oracle)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "test_fw ok" in r.stdout, r.stdout + r.stderr
