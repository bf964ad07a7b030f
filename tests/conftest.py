import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import __graft_entry__ as graft  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    lib = os.path.join(graft.PKG_DIR, "lib", "libmslam_b200.so")
    if not os.path.exists(lib):
        graft.build()
    return graft.load_package()


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.lib()
    return O


@pytest.fixture(scope="session")
def bundled():
    return np.load(os.path.join(GOLDEN, "bundled.npz"))


@pytest.fixture(scope="session")
def synthetic():
    return np.load(os.path.join(GOLDEN, "synthetic.npz"))


def canon(kps, desc=None):
    """canonical (octave, y, x) order -- OpenCV's order inside a level is a std::nth_element artefact"""
    key = np.lexsort((kps["x"], kps["y"], kps["octave"]))
    return (kps[key], desc[key]) if desc is not None else kps[key]
