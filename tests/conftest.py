import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def b2rt():
    """The product package (directory name has a hyphen, hence importlib)."""
    build = importlib.import_module("gpu-ray-tracing_b200.build")
    build.build()
    return importlib.import_module("gpu-ray-tracing_b200")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib


@pytest.fixture(scope="session")
def ctx(b2rt):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return b2rt.Context.get(0)
