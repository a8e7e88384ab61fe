import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

REFERENCE = "/root/reference"  # present in the build container only; never on the GPU box


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "reference: reads /root/reference (skipped where it is absent)")


@pytest.fixture(scope="session")
def rt():
    """The product package (loads libraytrace_b200.so; fails loudly if it was not built)."""
    pkg = importlib.import_module("rust-raytrace_b200")
    pkg.lib()
    return pkg


@pytest.fixture(scope="session")
def scenes():
    return importlib.import_module("rust-raytrace_b200.scenes")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.lib()
    return oracle


@pytest.fixture(scope="session")
def gpu(rt):
    if rt.device_count() < 1:
        pytest.fail("test marked gpu but no CUDA device is visible (the product has no CPU fallback)")
    return rt


needs_reference = pytest.mark.skipif(not os.path.isdir(REFERENCE), reason="/root/reference not present")
