"""The C-ABI library loads on a machine without a GPU and exports every symbol include/b2rt.h
declares; the ctypes binding covers each of them; compute entry points fail loudly (no CPU path)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared(header):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(b2rt_[a-z_0-9]+)\s*\(", text)))


def test_header_symbols_exported(b2rt):
    names = _declared("b2rt.h")
    assert len(names) >= 30
    lib = C.CDLL(b2rt.LIB_PATH)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, "declared in include/b2rt.h but not exported: %s" % missing
    assert sorted(b2rt.EXPORTS) == names, "ctypes binding and header disagree"
    # instrumentation entry points live in their own header (include/b2rt_tools.h), same library
    tools = [n for n in _declared("b2rt_tools.h") if n not in names]
    assert sorted(b2rt.TOOLS_EXPORTS) == tools and not [n for n in tools if not hasattr(lib, n)]
    # the multi-GPU entry points (include/b2rt_mg.h) are in the same library; NCCL is bound at run time, so loading needs none
    mg = [n for n in _declared("b2rt_mg.h") if n.startswith("b2rt_mg_")]
    assert sorted(b2rt.MG_EXPORTS) == mg and not [n for n in mg if not hasattr(lib, n)]
    # the benchmark scenes are a separate host-only library (include/b2rt_scenes.h): nothing of it in libb2rt.so
    scenes = _declared("b2rt_scenes.h")
    sl = C.CDLL(b2rt.SCENES_LIB_PATH)
    assert sorted(b2rt.SCENES_EXPORTS) == scenes and not [n for n in scenes if not hasattr(sl, n)]
    assert not [n for n in scenes if hasattr(lib, n)]


def test_reference_kernel_abi_exported(b2rt):
    """libb2rt_refabi.so: the reference's own C-linkage launcher names (include/b2rt_refabi.h)."""
    names = [n for n in _declared_ref()]
    path = os.path.join(os.path.dirname(b2rt.LIB_PATH), "libb2rt_refabi.so")
    assert os.path.exists(path), "libb2rt_refabi.so not built"
    lib = C.CDLL(path)
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def _declared_ref():
    text = open(os.path.join(ROOT, "include", "b2rt_refabi.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b((?:launch_|bind_|unbind_|copy_)[A-Za-z_0-9]+)\s*\(", text)))


def test_no_gpu_means_loud_failure(b2rt):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(b2rt.B2rtError) as e:
        b2rt.Context(0)
    assert "no CUDA device" in str(e.value) or "CUDA" in str(e.value)
    # a NULL context is refused, never silently computed on the host
    assert b2rt.lib().b2rt_trace(None, None, None, 16, 0, None, None) != 0


def test_product_never_imports_oracle():
    """Nothing under gpu-ray-tracing_b200/ may import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "gpu-ray-tracing_b200")
    for dirpath, _, files in os.walk(pkg):
        if os.path.basename(dirpath) == "build":
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                for needle in ("oracle_lib", "b2rt_oracle", "b2o_", "oracle/_ref", "libref_gpu", "ref_cpu"):
                    assert needle not in text, "%s mentions %s" % (os.path.join(dirpath, f), needle)
