import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def O():
    """The CPU oracle (oracle/): the checker, never the thing measured."""
    import __graft_entry__ as g
    return g.load_oracle()


@pytest.fixture(scope="session")
def pkg():
    """The product package (ctypes face of libtrt_b200.so)."""
    import __graft_entry__ as g
    p = g.load_package()
    p.api.lib()
    return p
