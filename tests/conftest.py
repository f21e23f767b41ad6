import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200)")


@pytest.fixture(scope="session")
def pk():
    import __graft_entry__ as g
    return g.package()


@pytest.fixture(scope="session")
def orc(pk):
    import __graft_entry__ as g
    return g.oracle()


@pytest.fixture(scope="session")
def dev(pk):
    """The product path: libpcb200.so through the C ABI. Fails loudly without it."""
    d = pk.GPU()
    d.gpu_init(0)
    yield d
    d.finalize_GPU()
