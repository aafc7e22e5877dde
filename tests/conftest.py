"""pytest configuration: registers the `gpu` marker and shared fixtures.

`-m "not gpu"` tests pin the oracle against the golden vectors produced by the
reference's own object code and exercise the host-side logic; `-m gpu` tests are
the parity tests proper and call the CUDA path through the C ABI.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


SCENE_NAMES = ["settingData", "cornellBox", "simple1", "simple2"]


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "reference_golden.npz"))


@pytest.fixture(scope="session")
def scene_path():
    return lambda name: os.path.join(ROOT, "scenes", name + ".head.json")


@pytest.fixture(scope="session")
def gpu_lib():
    """The CUDA library, or a loud failure: GPU tests never fall back."""
    import raytracingmin_b200 as rtm
    lib = rtm.load_library()
    return lib
