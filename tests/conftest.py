"""pytest configuration: the ``gpu`` marker, golden-fixture loader, repo root on sys.path."""
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


def load_golden(name: str) -> dict:
    """``tests/golden/<name>.npz`` as ``{case: {key: array}}``."""
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    out: dict = {}
    for k in z.files:
        case, key = k.split("/", 1)
        out.setdefault(case, {})[key] = z[k]
    return out


@pytest.fixture(scope="session")
def golden():
    return load_golden
