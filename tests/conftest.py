import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle import loader

    return loader.port()


@pytest.fixture(scope="session")
def checker():
    """Strongest CPU checker available: compiled reference (oracle/_ref) if present, else the C port."""
    from oracle import loader

    return loader.best()


@pytest.fixture(scope="session")
def reference():
    from oracle import loader

    r = loader.reference()
    if r is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return r
