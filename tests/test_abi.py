"""CPU-side checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/ctba.h declares; without a GPU the product path fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def _declared(name="ctba.h"):
    hdr = open(os.path.join(ROOT, "include", name)).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(ctba_[a-z_0-9]+)\s*\(", hdr)))


def test_library_exports_every_declared_symbol(pkg):
    names = _declared()
    assert len(names) >= 25
    L = pkg.lib()
    for n in names:
        assert hasattr(L, n), f"libctba.so does not export {n}"
    assert set(pkg.EXPORTS) == set(names)
    # the discrete-time estimator's boundary (include/ctba_dt.h)
    names_dt = [n for n in _declared("ctba_dt.h") if n.startswith("ctba_dt_")]
    assert len(names_dt) >= 25
    for n in names_dt:
        assert hasattr(L, n), f"libctba.so does not export {n}"
    assert set(pkg.EXPORTS_DT) == set(names_dt)


def test_no_cpu_fallback_without_gpu(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pkg.CtbaError):
        pkg.BSplineSO3R3(6, 100_000_000, 0)
    with pytest.raises(pkg.CtbaError):
        pkg.DiscreteTimeEstimator(0)
    h = ctypes.c_void_p()
    rc = pkg.lib().ctba_create(6, ctypes.c_int64(100_000_000), ctypes.c_int64(0), 0, ctypes.byref(h))
    assert rc == -2 and not h  # CTBA_E_CUDA


def test_product_code_never_touches_the_oracle():
    pk = os.path.join(ROOT, "rpg_vision-based_slam_b200")
    for dp, _, fs in os.walk(pk):
        for f in fs:
            if f.endswith((".py", ".cu", ".h", ".cpp", ".cuh")):
                txt = open(os.path.join(dp, f)).read()
                assert "liboracle" not in txt and "oracle.binding" not in txt and "oracle/" not in txt.replace("the oracle", ""), f
