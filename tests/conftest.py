"""Test configuration.

`-m "not gpu"`: oracle vs golden vectors, host logic, C-ABI symbol checks, the CPU debug harness of the
scalar stage functions (tests/emu) - no CUDA calls.  `-m gpu`: the parity tests proper, through the C-ABI
of libref_cuda.so on a real B200, against the oracle (unmodified ref_soft) and the committed goldens.
Nothing here reads /root/reference at run time; prebuilt libraries travel with the snapshot.
"""
import importlib
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
EMU_PATH = os.path.join(ROOT, "tests", "emu", "libref_cuda_emu.so")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    import __graft_entry__ as g

    g.build()
    return True


@pytest.fixture(scope="session")
def pkg(built):
    return importlib.import_module("xash3d-fwgs_b200")


@pytest.fixture(scope="session")
def scene(built):
    return importlib.import_module("xash3d-fwgs_b200.scene")


@pytest.fixture(scope="session")
def golden_frames():
    meta = json.load(open(os.path.join(GOLDEN, "frames_golden.json")))
    data = np.load(os.path.join(GOLDEN, "frames_golden.npz"))
    return meta, data


@pytest.fixture(scope="session")
def golden_gamma():
    return np.load(os.path.join(GOLDEN, "gamma_golden.npz"))
