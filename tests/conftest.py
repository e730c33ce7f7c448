import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
GOLDEN = os.path.join(ROOT, "tests", "golden")

# Tolerances of BASELINE.json's north_star: relative factor difference <= 1e-11;
# ||A - R^T R|| / ||A|| and ||Q^T Q - I|| <= 1e-13 * n.
TOL_FACTOR = 1e-11
TOL_RESIDUAL_PER_N = 1e-13


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def relerr(a, b):
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / np.linalg.norm(np.asarray(b)))


@pytest.fixture(scope="session")
def engine():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from capital_b200.engine import Engine
    e = Engine(0)
    yield e
    e.close()
