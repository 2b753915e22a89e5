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
def product():
    """The ctypes view of libtowr_b200.so, built in-tree if missing."""
    import towr2025_b200 as T
    if not os.path.exists(T.LIB_PATH):
        T.build()
    T.lib()
    return T


@pytest.fixture(scope="session")
def oracle():
    """TEST INFRASTRUCTURE: the CPU restatement of the reference path."""
    from oracle import oracle_py as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def emu():
    """TEST INFRASTRUCTURE: host emulation of the kernel arithmetic (tests/emu)."""
    import ctypes as C
    d = os.path.join(ROOT, "tests", "emu")
    subprocess.check_call(["make", "-C", d], stdout=subprocess.DEVNULL)
    E = C.CDLL(os.path.join(d, "_build", "libemu.so"))
    E.emu_last_error.restype = C.c_char_p
    return E
