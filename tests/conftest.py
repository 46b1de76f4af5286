import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def wf():
    """The product package (directory name has a hyphen, hence importlib)."""
    return importlib.import_module("cl-waffe2_b200")


@pytest.fixture(scope="session")
def gpu(wf):
    """Initialised device backend; fails loudly (no CPU fallback) when there is no B200."""
    wf.init(0)
    return wf


@pytest.fixture(scope="session")
def L():
    from oracle import loopnest
    return loopnest
