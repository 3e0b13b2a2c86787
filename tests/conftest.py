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


# Modules whose Eval tests must hold on BOTH device paths: the tiled pack -> eval -> reduce path and the one-launch
# small-problem kernel (sctl_b200/csrc/small_impl.cuh). The library picks by problem size; the tests pin each in turn.
_BOTH_PATHS = {"test_gpu_parity", "test_gpu_edge_cases", "test_gpu_canary", "test_gpu_threads", "test_gpu_fmm"}


def pytest_generate_tests(metafunc):
    if metafunc.module.__name__.split(".")[-1] in _BOTH_PATHS and "eval_path" in metafunc.fixturenames:
        metafunc.parametrize("eval_path", ["tiled", "small"], indirect=True)


@pytest.fixture
def eval_path(request, monkeypatch):
    if request.param == "tiled":
        monkeypatch.setenv("SCTL_B200_SMALL_MAX_PAIRS", "0")
        monkeypatch.setenv("SCTL_B200_STAGE_MAX_BYTES", "0")  # and the general host-pointer path (per-array copies)
    else:  # force the small kernel wherever its size guard allows it (Ns <= 8192)
        monkeypatch.setenv("SCTL_B200_SMALL_MAX_PAIRS", str(2**31 - 1))
    return request.param
