import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def pkg():
    import isogs_b200
    return isogs_b200.pkg


@pytest.fixture(scope="session")
def syn():
    import importlib

    import isogs_b200  # noqa: F401
    return importlib.import_module("isogs_slam_cpp_b200.synthetic")


@pytest.fixture(autouse=True)
def _no_blend_timeouts(request):
    """After every GPU test: no bounded wait of the blend kernels' mbarrier protocol may have timed out."""
    yield
    if request.node.get_closest_marker("gpu") is None:
        return
    import torch
    if not torch.cuda.is_available():
        return
    import isogs_b200
    torch.cuda.synchronize()
    n = isogs_b200.pkg.lib().isogs_debug_blend_timeouts()
    assert n == 0, f"{n} waits of the blend kernels timed out"
