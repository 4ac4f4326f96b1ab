import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
os.environ.setdefault("OMP_NUM_THREADS", str(min(8, os.cpu_count() or 1)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Make sure the shared libraries exist (the driver's build() normally did this)."""
    from hypre_b200 import LIB_B200, LIB_HYPRE_B200, LIB_ORACLE
    missing = [p for p in (LIB_B200, LIB_HYPRE_B200, LIB_ORACLE) if not os.path.exists(p)]
    if missing:
        import __graft_entry__ as g
        g.build()
    return True


@pytest.fixture(scope="session")
def H(built):
    from hypre_b200 import hypre
    hypre.lib()
    return hypre


@pytest.fixture(scope="session")
def ctx(H):
    """the binding's own device context (GPU tests only)."""
    import ctypes as C
    from hypre_b200 import capi
    c = H.lib().hypre_b200_context()
    if not c:
        raise capi.B200Error("no CUDA device: " + capi.lib().b200_last_error().decode())
    return C.c_void_p(c)
