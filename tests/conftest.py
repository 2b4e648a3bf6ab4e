import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def crt():
    """the product package (ctypes binding of lib/libcrt_b200.so); builds it when the library is missing"""
    import __graft_entry__ as entry

    if not os.path.exists(os.path.join(entry.PKG_DIR, "lib", "libcrt_b200.so")):
        entry.build()
    return entry.load_package()


@pytest.fixture(scope="session")
def ot(crt):
    """oracle tools (test infrastructure); makes sure the checker is built"""
    import oracle_tools

    oracle_tools.ensure_port_build()
    oracle_tools.ensure_reference_build()
    return oracle_tools


@pytest.fixture(scope="session")
def renderer(crt):
    r = crt.Renderer(0)  # raises on a box without a GPU: the gpu tests must never pass on a fallback
    yield r
    r.close()
