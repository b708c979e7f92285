import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def ctx():
    """One device context for the GPU tests.  No skip-on-missing: a GPU test without a
    usable libhistorian_b200.so / device must fail loudly (there is no CPU fallback)."""
    from historian_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()
