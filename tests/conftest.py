import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_json():
    with open(os.path.join(GOLDEN_DIR, "reference_golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def golden_arrays():
    return np.load(os.path.join(GOLDEN_DIR, "reference_golden.npz"))


@pytest.fixture
def cpu_backend(monkeypatch):
    """Route the python-level kernel wrappers to the oracle so that HOST logic (tables,
    laziness, scaler trajectories, bucket bookkeeping) can be exercised without a GPU.
    Test double only: the product has no such path."""
    from tests import fake_backend
    fake_backend.install(monkeypatch)
    yield fake_backend
