import json
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))
GOLD = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def manifest():
    return json.loads((GOLD / "manifest.json").read_text())


@pytest.fixture(scope="session")
def golden_counts():
    return np.load(GOLD / "counts.npz")


@pytest.fixture(scope="session")
def golden_searches():
    return np.load(GOLD / "searches.npz")


def run_klist(run):
    """k lengths of a golden search run (range runs are stored as [min, max])."""
    return list(range(run["k"][0], run["k"][1] + 1)) if run["range"] else list(run["k"])
