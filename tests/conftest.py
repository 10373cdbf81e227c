import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    """The C part of the CPU oracle (generator + fused moments), built on demand with gcc."""
    import ctypes
    so = os.path.join(ROOT, "oracle", "libppls_oracle.so")
    src = os.path.join(ROOT, "oracle", "ppls_oracle.c")
    if not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    L = ctypes.CDLL(so)
    from oracle import cbind
    cbind.declare(L)
    return L
