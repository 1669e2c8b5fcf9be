import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.dirname(os.path.abspath(__file__))):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    """Load tests/golden/<name>.npz; big tensors were stored as float32 (+ fp64 checksums)."""
    with np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False) as z:
        return {k: z[k] for k in z.files}


def golden_tol(arr):
    """Comparison tolerance implied by how a golden tensor was stored."""
    return 2e-7 if arr.dtype == np.float32 else 1e-12


@pytest.fixture(scope="session")
def golden():
    return load_golden
