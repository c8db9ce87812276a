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
    """The oracle: the reference engine compiled in place (oracle/_ref)."""
    from oracle import ref_engine
    ref_engine.lib()
    return ref_engine


@pytest.fixture(scope="session")
def gpu_ctx():
    from tunebfree_b200 import engine
    ctx = engine.Context(0)
    yield ctx
    ctx.close()
