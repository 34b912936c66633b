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


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def unpack_masks(bits, shape, values=None):
    """Inverse of np.packbits(masks != 0) used by oracle/gen_golden.py."""
    shape = tuple(int(x) for x in shape)
    m = np.unpackbits(bits)[: int(np.prod(shape))].reshape(shape).astype(np.uint8)
    if values is not None:
        m = m * np.asarray(values, np.uint8).reshape(-1, *([1] * (len(shape) - 1)))
    return m


@pytest.fixture(scope="session")
def has_cuda():
    import torch
    return torch.cuda.is_available()
