import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")
    config.addinivalue_line("markers", "slow: long-running")


def has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def built():
    """Native libraries are built once per session (no-op when up to date)."""
    from tclb_b200 import build
    build.build_core()
    return True


@pytest.fixture(scope="session")
def tclb_bin(built):
    return os.path.join(ROOT, "bin", "tclb")


@pytest.fixture
def reference_dir():
    if not os.path.isdir(REFERENCE):
        pytest.skip("reference checkout not mounted")
    return REFERENCE
