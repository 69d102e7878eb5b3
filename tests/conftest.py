import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def golden_npz(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def golden_json(name):
    import json

    with open(os.path.join(GOLDEN, name)) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def rooms():
    return golden_npz("sokoban_rooms.npz")


@pytest.fixture(scope="session")
def steps():
    return golden_npz("sokoban_steps.npz")
