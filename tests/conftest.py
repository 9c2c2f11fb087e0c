import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as g
    return g.load_package()


@pytest.fixture(scope="session")
def tiny():
    import golden_util
    return golden_util.load_scene_golden("skin_tiny")


@pytest.fixture(scope="session")
def dev(pkg):
    d = pkg.SkinDevice(0)   # raises without a GPU / without the built library: no fallback
    yield d
    d.close()
