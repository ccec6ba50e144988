import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU checker (test infrastructure): port = our C restatement, ref = compiled reference headers (may be None)."""
    import oracle as O
    return O


@pytest.fixture(scope="session")
def api():
    from libstevi_b200 import api as A
    return A


@pytest.fixture(scope="session")
def ctx(api):
    c = api.Context(0)
    yield c
    c.close()
