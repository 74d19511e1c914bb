import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled into oracle/_ref/libffref.so (the checker)."""
    from oracle import ref as r
    if not r.available():
        pytest.skip("oracle/_ref/libffref.so not built (make -C oracle ref)")
    r.lib()
    return r


@pytest.fixture(scope="session")
def engine():
    """The product: libff_b200 bound to cuda:0 through the C ABI.  Fails loudly without it."""
    import libff_b200 as m
    m.init(0)
    yield m
    m.shutdown()
