import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def bgr():
    import bottleneck_guided_rrt_b200 as B
    B.lib()
    return B


def canon_equal(idx_a, d2_a, idx_b, d2_b):
    """bit-exact comparison of kNN results: d2 arrays identical bit for bit, index arrays
    identical (both sides order ties by index)."""
    return (np.array_equal(idx_a, idx_b)
            and np.array_equal(d2_a.view(np.uint32), d2_b.view(np.uint32)))
