import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port_oracle():
    """oracle/liboracle.so — the plain-C restatement (built on demand; gcc only)."""
    from oracle import pyorc
    if not pyorc.available("port"):
        pyorc.build("port")
    return pyorc.load("port")


@pytest.fixture(scope="session")
def ref_oracle():
    """oracle/_ref/libsphref.so — the reference's own kernel source compiled as C++.
    Prebuilt here (needs /root/reference to build); skipped where it does not exist."""
    from oracle import pyorc
    if not pyorc.available("reference"):
        if os.path.isdir("/root/reference"):
            pyorc.build("reference")
        else:
            pytest.skip("oracle/_ref not built and /root/reference absent")
    return pyorc.load("reference")
