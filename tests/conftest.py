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

