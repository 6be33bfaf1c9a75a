import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Make sure the product library and the oracle are built (the .so files travel to the GPU box prebuilt)."""
    from lara_b200 import _lib
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    from oracle import binding
    binding.load()
    return True


@pytest.fixture(scope="session")
def gpu_ctx_factory(built):
    import lara_b200 as lb

    def make(**kw):
        p = lb.Parameters()
        p.set(**kw)
        return lb.Context(p), p
    return make
