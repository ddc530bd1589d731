import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_cuda():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture
def arcq_option():
    """Force a kernel / tile choice through arcq_set_option for one test (every choice computes the same result);
    restores the shape-driven defaults afterwards."""
    from arcquant_b200 import agemm
    touched = []

    def set_(name, value):
        agemm.set_option(name, str(value))
        touched.append(name)

    yield set_
    for name in touched:
        agemm.set_option(name, "")
