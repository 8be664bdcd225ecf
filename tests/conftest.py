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
    """A GPU context; gpu-marked tests fail loudly (never skip to a CPU path) when CUDA is absent."""
    import paffy_b200
    c = paffy_b200.Context(0)
    yield c
    c.close()
