import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def have_reference():
    return os.path.isfile(os.path.join(REFERENCE, "utils.py"))


@pytest.fixture(scope="session")
def lib():
    """libcrnn.so, built if needed (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as g
    g.build()
    from distributed_char_rnn_b200 import _lib
    return _lib.load()
