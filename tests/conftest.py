import importlib
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    """The product package; builds the extension when the .so is missing (CPU box)."""
    p = importlib.import_module("semantic-slam_b200")
    if not os.path.exists(p.capi.LIB_PATH) or not os.path.exists(p.capi.DEV_LIB_PATH):
        import __graft_entry__ as g
        g.build()
    p.load()
    return p


@pytest.fixture
def dev_lib(pkg):
    """Tests that force a code path (tsdf_set_variant) run against libtsdf_b200_dev.so: objects created
    while this fixture is active bind to it."""
    pkg.capi.select(dev=True)
    yield pkg
    pkg.capi.select(dev=False)


@pytest.fixture(scope="session")
def synth(pkg):
    return pkg.synth


@pytest.fixture(scope="session")
def oracle():
    from oracle.oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def golden_cu():
    return np.load(os.path.join(GOLDEN, "tsdfcu_small.npz"))


@pytest.fixture(scope="session")
def golden_py():
    return np.load(os.path.join(GOLDEN, "fusionpy_small.npz"))


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)
