import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    # Files named test_zz_* are collected last so that an early stop (-x) has already run the rest.
    items.sort(key=lambda it: (os.path.basename(str(it.fspath)).startswith("test_zz_"),))


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    def load(name):
        return np.load(os.path.join(GOLDEN, name + ".npz"))
    return load


def world_from_npz(g, cls):
    return cls(int(g["dim"]), g["env_lo"], g["env_hi"], g["box_lo"], g["box_hi"], g["sph_c"], g["sph_r"],
               float(g["buffer"]))
