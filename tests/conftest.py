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


def elemerr(a, b, rtol=1e-10, floor=1e-12):
    """Element-wise parity gate of SURVEY.md §8(d): |a - b| <= rtol |b| + floor * colscale for EVERY entry, colscale = the largest
    magnitude of the oracle's column (Jacobians: per column, so that a small column next to a large one is held to its own scale;
    residual vectors: the whole vector).  A third term, 64 eps * (largest magnitude of the whole array) ~ 1.4e-14 * max|b|, is the rounding
    floor between two different FP64 evaluation orders (closed-form Jacobians vs Jets): columns that are mathematically zero at the
    test point (e.g. the rotation of the last knot at u = 0) hold 1e-13-sized noise on BOTH sides and have no scale of their own.
    Returns max |a - b| / bound: parity holds iff the result is <= 1."""
    import numpy as np
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.size == 0:
        return 0.0
    colscale = np.abs(b).max(axis=0, keepdims=True) if b.ndim == 2 else np.abs(b).max()
    bound = rtol * np.abs(b) + floor * colscale + 64.0 * np.finfo(np.float64).eps * np.abs(b).max()
    d = np.abs(a - b)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(d == 0.0, 0.0, d / bound)      # an all-zero oracle array admits only exact zeros
    q = np.nan_to_num(q, nan=np.inf)
    if os.environ.get("CTBA_ELEM_DEBUG") and q.max() > 1.0:   # list the worst offenders
        for i in np.argwhere(q > 1.0)[:12]:
            i = tuple(i)
            print("elemerr violation at", i, "got", a[i], "oracle", b[i], "diff", d[i], "bound", np.broadcast_to(bound, b.shape)[i])
    return float(q.max())
