import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ctx():
    """One libg16b200 context on cuda:0.  Fails loudly (no skip, no fallback) if the GPU path is unusable."""
    from gnark_icicle_b200 import lib
    c = lib.Context(0)
    yield c
    c.close()
