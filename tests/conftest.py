import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "kat_hashes.json")) as f:
        data = json.load(f)
    return {c["name"]: c for c in data["cases"]}


@pytest.fixture(scope="session")
def golden_integer():
    """KITTI-size vectors of the integer / DPX mode (tests/golden/make_golden_integer.py: restatement, window 1x1)."""
    import json

    with open(os.path.join(ROOT, "tests", "golden", "kat_hashes_integer.json")) as f:
        data = json.load(f)
    return {c["name"]: c for c in data["cases"]}


@pytest.fixture(scope="session")
def golden_fullsize():
    """Full-size vectors of the configurations the reference cannot run: 1920x1080 D=256 (float and integer mode) and SAD
    block matching at 1242x375 (tests/golden/make_golden_fullsize.py: widened restatement)."""
    import json

    with open(os.path.join(ROOT, "tests", "golden", "kat_hashes_fullsize.json")) as f:
        data = json.load(f)
    return {c["name"]: c for c in data["cases"]}
