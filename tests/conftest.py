import os
import sys

import pytest

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """libhpcx.so and the oracle are built in-tree once per session (no-op when fresh)."""
    from libhpc_b200 import build as b
    b.build()
    import oracle
    oracle.build()


@pytest.fixture(scope="session")
def ctx():
    """The CUDA context of the library; GPU tests fail loudly when it cannot be created."""
    import libhpc_b200 as hp
    c = hp.Context(0)
    yield c
    c.close()
