import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def lib():
    """libc3vq.so, built in-tree if stale (nvcc cross-compiles without a GPU)."""
    from c3dgs_b200 import build, _lib
    if build.stale():
        build.build()
    return _lib.load()


def golden_path(name):
    return os.path.join(ROOT, "tests", "golden", f"vq_{name}.npz")


GOLDEN_CASES = ["sh45", "sh48", "shape7", "cov6"]
