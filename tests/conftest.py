import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle_mod():
    """The CPU oracle (test infrastructure).  Built on demand; `ref` only where /root/reference exists."""
    import oracle
    oracle.build(ref=os.path.isdir("/root/reference/include/mitsuba"))
    return oracle


@pytest.fixture(scope="session")
def gpu_lib():
    """The product library on a real GPU.  Fails (does not skip) if it cannot run."""
    from mitsuba_b200 import lib
    assert os.path.exists(lib.LIB_PATH), "libmtsb200.so missing: build it with __graft_entry__.build()"
    lib.init(-1)
    return lib
