import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    """Load tests/golden/<name>.npz -> (dict of arrays, list of case-meta dicts)."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    return z, meta


@pytest.fixture(scope="session")
def golden_ising():
    return load_golden("ising")


@pytest.fixture(scope="session")
def golden_ising_de():
    return load_golden("ising_de")


@pytest.fixture(scope="session")
def golden_rx():
    return load_golden("rx")


@pytest.fixture(scope="session")
def golden_latgas():
    return load_golden("latgas")


@pytest.fixture(scope="session")
def golden_latgas_move():
    return load_golden("latgas_move")


@pytest.fixture(scope="session")
def golden_discs():
    return load_golden("discs")
