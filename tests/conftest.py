import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "open3d-py-extension_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure only)."""
    from oracle import orc as _orc
    _orc.lib()
    return _orc


@pytest.fixture(scope="session")
def bunny():
    """data/bunny.npy of the reference (35 947 points), committed as a float32 fixture."""
    return np.load(os.path.join(GOLDEN, "bunny_xyz_f32.npy")).astype(np.float64)


@pytest.fixture(scope="session")
def ctx():
    from open3dpypro import _native
    c = _native.Context(0)
    yield c
    c.close()
