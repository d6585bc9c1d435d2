import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def gctx():
    """One GPUContext on device 0 for the whole session (GPU tests only)."""
    from systemml_b200 import GPUContext
    ctx = GPUContext(0)
    yield ctx
    ctx.destroy()
