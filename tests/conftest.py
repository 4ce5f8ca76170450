import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def manifest():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "manifest.json")))


def fixture_obu(man_entry):
    from tests.golden.gen_fixtures import read_ivf
    return b"".join(read_ivf(os.path.join(ROOT, "tests", "golden", man_entry["file"])))
