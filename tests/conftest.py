import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """the oracle (test infrastructure): restated iDynTree + controller algebra + the real vendored qpOASES"""
    from oracle import oracle as o
    o.lib()
    return o


@pytest.fixture(scope="session")
def wl():
    from quadrupedwholebodycontroller_b200 import workloads
    return workloads


@pytest.fixture(scope="session")
def wbc():
    import quadrupedwholebodycontroller_b200 as w
    from quadrupedwholebodycontroller_b200.build import build_library
    build_library()
    return w
