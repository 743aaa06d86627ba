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
    """The compiled libear reference (oracle/_ref); test infrastructure only."""
    from oracle import ref as r
    if not r.available():
        pytest.skip("oracle/_ref/libear_ref.so not built (make -C oracle)")
    return r


@pytest.fixture(scope="session")
def eb():
    import libear_b200
    return libear_b200
