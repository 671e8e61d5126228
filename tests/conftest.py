import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """libgrb200.so, the C++ wrapper driver and the CPU checker must exist (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge

    lib = os.path.join(ROOT, "gnuradio_b200", "libgrb200.so")
    drv = os.path.join(ROOT, "gnuradio_b200", "cxx", "libgrb200_blocks.so")
    port = os.path.join(ROOT, "oracle", "libgrb_oracle.so")
    if not (os.path.exists(lib) and os.path.exists(drv) and os.path.exists(port)):
        ge.build()
    yield


@pytest.fixture(scope="session")
def orc():
    """The most authoritative CPU oracle present: the reference's own compiled sources, else the C port."""
    from oracle import pyoracle as po

    return po.best()


def has_gpu():
    import gnuradio_b200 as grb

    try:
        return grb.device_count() > 0
    except Exception:
        return False
