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


@pytest.fixture(scope="session")
def kitti():
    """Bundled KITTI scan as (124668, 3) float32 -- the N x 3 layout benchmark.py:19 documents."""
    return np.fromfile(os.path.join(GOLDEN, "kitti_000000_xyz.f32"), np.float32).reshape(-1, 3)


@pytest.fixture(scope="session")
def knn_golden():
    return np.load(os.path.join(GOLDEN, "knn_golden.npz"))


@pytest.fixture(scope="session")
def icp_golden():
    return np.load(os.path.join(GOLDEN, "icp_golden.npz"))


@pytest.fixture(scope="session")
def front_golden():
    return np.load(os.path.join(GOLDEN, "front_golden.npz"))


@pytest.fixture(scope="session")
def iss_golden():
    return np.load(os.path.join(GOLDEN, "iss_golden.npz"))


@pytest.fixture(scope="session")
def dbscan_golden():
    return np.load(os.path.join(GOLDEN, "dbscan_golden.npz"))


def golden_cases(g):
    return sorted({k.split("/")[0] for k in g.files})


def csr_sets(off, idx):
    return [tuple(sorted(int(x) for x in idx[off[i]:off[i + 1]])) for i in range(len(off) - 1)]
