import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built():
    """builds the oracle checkers (and the CUDA library if it is missing); nvcc cross-compiles without a GPU"""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def restatement(built):
    from oracle import pyoracle
    return pyoracle.Restatement()


@pytest.fixture(scope="session")
def reference(built):
    from oracle import pyoracle
    if not pyoracle.Reference.available():
        pytest.skip("oracle/_ref/libref_hotpath.so not built (needs /root/reference at build time)")
    return pyoracle.Reference()
