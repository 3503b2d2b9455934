import importlib
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
PKG = "rpg_vision-based_slam_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return importlib.import_module(PKG)


@pytest.fixture(scope="session")
def scene():
    return importlib.import_module(PKG + ".scene")


@pytest.fixture(scope="session")
def oracle():
    from oracle import binding
    binding.build()
    return binding


@pytest.fixture(scope="session")
def hostdbg():
    """CPU build of the CTBA_HD evaluators (debug harness, tests only)."""
    import ctypes
    d = os.path.join(ROOT, "tests", "hostdbg")
    so = os.path.join(d, "libhostdbg.so")
    srcs = [os.path.join(d, "hostdbg.cpp"), os.path.join(ROOT, PKG, "csrc", "ctba_terms.h"), os.path.join(ROOT, PKG, "csrc", "ctba_math.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-o", so, srcs[0]])
    return ctypes.CDLL(so)


def relerr(a, b):
    import numpy as np
    if b.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
