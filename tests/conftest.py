import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure)."""
    import oracle_py
    oracle_py.build()
    oracle_py.load()
    return oracle_py


@pytest.fixture(scope="session")
def hostmath():
    """The product's math header compiled for the host (tests only)."""
    d = os.path.join(ROOT, "tests", "hostmath")
    subprocess.check_call(["make", "-s", "-C", d])
    return C.CDLL(os.path.join(d, "libhostmath.so"))


@pytest.fixture(scope="session")
def sba_lib():
    so = os.path.join(ROOT, "steviapp_b200", "libsba_b200.so")
    if not os.path.exists(so):
        subprocess.check_call([os.path.join(ROOT, "build.sh")])
    from steviapp_b200 import _capi
    return _capi.load()


def relerr(a, b):
    """SURVEY 8(d) gate: max|a-b| / max(max|b|, 1)."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1.0))


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)
