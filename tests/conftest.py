import os
import sys

import numpy as np
import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def golden_unet():
    return np.load(os.path.join(GOLDEN, "unet.npz"))


@pytest.fixture(scope="session")
def golden_schedule():
    return np.load(os.path.join(GOLDEN, "schedule.npz"))


@pytest.fixture(scope="session")
def golden_traj():
    return np.load(os.path.join(GOLDEN, "trajectory.npz"))


@pytest.fixture(scope="session")
def golden_unet64():
    return np.load(os.path.join(GOLDEN, "unet64.npz"))


@pytest.fixture(scope="session")
def golden_manip():
    return np.load(os.path.join(GOLDEN, "manipulation.npz"))


def record_parity(key, value):
    """Appends one measured parity figure to gpurun_out/parity_report.jsonl (kept as evidence under profiles/)."""
    import json
    d = os.path.join(ROOT, "gpurun_out")
    os.makedirs(d, exist_ok=True)
    with open(os.path.join(d, "parity_report.jsonl"), "a") as f:
        f.write(json.dumps({"key": key, "value": value}) + "\n")
