import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


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
def golden():
    def load(name):
        return dict(np.load(os.path.join(GOLDEN, name + ".npz")))
    return load


def simple_test_data(N, d):
    """create_simple_test_data, reference src/util/attention_helper.h:151-173."""
    Q = np.zeros((N, d), np.float32)
    for i in range(min(N, d)):
        Q[i, i] = 1.0
    i = np.arange(N, dtype=np.float32)[:, None]
    j = np.arange(d, dtype=np.float32)[None, :]
    V = (i * np.float32(4.0) + j + np.float32(1.0)).astype(np.float32)
    return Q, Q.copy(), V
