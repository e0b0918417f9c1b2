import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle import pyoracle
    return pyoracle.load("port")


@pytest.fixture(scope="session")
def ref():
    from oracle import pyoracle
    if not pyoracle.available("reference"):
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return pyoracle.load("reference")
