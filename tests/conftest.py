import json
import os
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, HERE, os.path.join(ROOT, "tools")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def kat():
    with open(os.path.join(HERE, "golden", "kat.json")) as f:
        return json.load(f)["cases"]


def golden_cases():
    g = os.path.join(HERE, "golden")
    return sorted(d for d in os.listdir(g) if os.path.isdir(os.path.join(g, d)))


@pytest.fixture(scope="session")
def gpu_lib():
    import finito_b200 as fb

    if fb.device_count() < 1:
        pytest.fail("a test marked gpu ran without a CUDA device")
    return fb
