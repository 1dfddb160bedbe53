import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PKG_NAME = "locally-consistent-visual-odometry-pipeline_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module(PKG_NAME)


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def ctx(pkg):
    """One shared vo_ctx on cuda:0.  Fails loudly (no CPU fallback) if there is no GPU."""
    c = pkg.Context(device=0, max_w=2048, max_h=1024)
    yield c
    c.close()
