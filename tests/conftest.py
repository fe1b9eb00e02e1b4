import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

EMUL_LIB = os.path.join(ROOT, "tests", "emul", "libgpqhe_b200_emul.so")
REAL_LIB = os.path.join(ROOT, "gpqhe_b200", "libgpqhe_b200.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def emul_lib():
    """Library sources compiled against the fibre-based CUDA stand-in (tests/emul):
    test infrastructure for the CPU tier, never loaded by the package itself."""
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "tests", "emul"), "-j8", "all"])
    return EMUL_LIB


@pytest.fixture(scope="session")
def real_lib():
    if not os.path.exists(REAL_LIB):
        import __graft_entry__ as ge
        ge.build()
    return REAL_LIB
