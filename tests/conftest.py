import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
FLAT = os.path.join(ROOT, "scenes_flat")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


def fixture_dir(name):
    """Fixtures live in tests/golden (committed) or scenes_flat (git-ignored, big)."""
    for d in (GOLDEN, FLAT):
        if os.path.exists(os.path.join(d, name + ".ljs")):
            return d
    return None


@pytest.fixture(scope="session")
def manifest():
    return json.load(open(os.path.join(GOLDEN, "manifest.json")))


@pytest.fixture(scope="session")
def gpu():
    from lajolla_public_b200 import LajollaGPU
    g = LajollaGPU(0)   # raises LjError without a device / library: no fallback
    yield g
    g.close()
