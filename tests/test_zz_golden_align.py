"""Golden vectors of the frozen-spline world alignment (tests/golden/align_n*.npz, generator tests/golden/make_golden_align.py).
CPU: (1) the oracle still reproduces them; (2) an INDEPENDENT numpy/scipy statement of
GlobalPositionSensorWithAlignmentCorrectionCostFunctor (global_position_cost_functor.h:82-150) — blending matrices from the binomial
formula, scipy rotations, Jacobian by central differences over the 11 tangent directions — agrees with the committed residuals and
Jacobian, which pins the conventions (window arithmetic, sign of the time correction, right perturbation of R_w_g_corr, order of the
11 columns) independently of the oracle's code.  GPU: the CUDA path reproduces the committed vectors through the C ABI.
(The file sorts last on purpose: its GPU part was added after the round's last GPU run.)"""
import os
import sys

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden_align as mg  # noqa: E402
from test_golden import _blend  # noqa: E402


def _load(name):
    return np.load(os.path.join(HERE, "golden", name + ".npz"))


def _check_case(case, g):
    fp = np.array([case["z"].sum(), case["q"].sum(), case["p"].sum(), float(len(case["t_ns"]))])
    np.testing.assert_allclose(fp, g["fingerprint"], rtol=1e-13)


@pytest.mark.parametrize("name", list(mg.CASES))
def test_oracle_reproduces_alignment_golden(name):
    g = _load(name)
    binding, align_case, case = mg.case_of(name)
    _check_case(case, g)
    r, J, _ = mg.evaluate_at_point(binding, align_case, case)
    np.testing.assert_allclose(r, g["r"], rtol=1e-12, atol=1e-12)
    np.testing.assert_allclose(J[:, 6 * case["N"]:], g["J"], rtol=1e-12, atol=1e-12)
    o = binding.OracleSpline(case["N"], case["dt_ns"], case["start_t_ns"])
    align_case.load_alignment(o, case, True, True)
    s = o.optimize(30)
    np.testing.assert_allclose(s["iterations"][:, 0], g["lm_iterations"][:, 0], rtol=1e-9)
    np.testing.assert_allclose(o.getWorldColmapTransformation()[0], g["T_wg"], atol=1e-9)


def _predict(case, x):
    """x = [t_corr, scale_corr, rotvec of the right perturbation of R_w_g_corr (3), p_w_g_corr (3), p_bp (3)] -> stacked residuals"""
    N, dt_ns = case["N"], case["dt_ns"]
    dt_s = dt_ns * 1e-9
    Mc, Mp = _blend(N, True), _blend(N, False)
    R0 = Rot.from_quat(case["T_wg_init"][:4]) * Rot.from_quat(mg.WG_CORR[:4]) * Rot.from_rotvec(x[2:5])
    p_wg = case["T_wg_init"][4:] + x[5:8]
    s_ = case["scale_init"] + x[1]
    out = []
    for t, z in zip(case["t_ns"], case["z"]):
        st_ns = int(t) - case["start_t_ns"] - int(case["t_off_init"] * 1e9)
        s = st_ns // dt_ns
        u = (st_ns * 1e-9 - x[0] - float(s) * dt_s) / dt_s
        pw = u ** np.arange(N)
        lam, c = Mc @ pw, Mp @ pw
        q, p = case["q"], case["p"]
        R = Rot.from_quat(q[s])
        for i in range(1, N):
            R = R * Rot.from_rotvec(lam[i] * (Rot.from_quat(q[s + i - 1]).inv() * Rot.from_quat(q[s + i])).as_rotvec())
        ant = R.apply(x[8:11]) + c @ p[s:s + N]
        out.append((s_ * R0.apply(ant) + p_wg - z) / case["noise_arg"])
    return np.concatenate(out)


@pytest.mark.parametrize("name", list(mg.CASES))
def test_independent_alignment_residuals_match_golden(name):
    g = _load(name)
    _, _, case = mg.case_of(name)
    _check_case(case, g)
    x0 = np.r_[mg.CORR10[9], mg.CORR10[0], np.zeros(3), mg.WG_CORR[4:], mg.P_BP]
    r = _predict(case, x0)
    assert r.shape == g["r"].shape
    np.testing.assert_allclose(r, g["r"], rtol=1e-9, atol=1e-9)
    J = np.zeros((r.size, 11))
    for c in range(11):
        h = 1e-6
        e = np.zeros(11); e[c] = h
        J[:, c] = (_predict(case, x0 + e) - _predict(case, x0 - e)) / (2 * h)
    np.testing.assert_allclose(J, g["J"], rtol=1e-6, atol=1e-6 * np.abs(g["J"]).max())


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(mg.CASES))
def test_cuda_reproduces_alignment_golden(pkg, name):
    import align_case
    g = _load(name)
    _, _, case = mg.case_of(name)
    _check_case(case, g)
    d = pkg.BSplineSO3R3(case["N"], case["dt_ns"], case["start_t_ns"], device=0)
    align_case.load_alignment(d, case, True, True)
    d.setGlobals(case["scale_init"], mg.P_BP, np.array([0.0, 0.0, -9.81]), case["t_off_init"])
    d.setCorrections(mg.CORR10)
    d.setWorldColmapTransformation(case["T_wg_init"], mg.WG_CORR)
    r, J = d.evaluate(7)
    np.testing.assert_allclose(r, g["r"], rtol=1e-10, atol=1e-10 * np.abs(g["r"]).max())
    np.testing.assert_allclose(J, g["J"], rtol=1e-10, atol=1e-10 * np.abs(g["J"]).max())
    d2 = pkg.BSplineSO3R3(case["N"], case["dt_ns"], case["start_t_ns"], device=0)
    align_case.load_alignment(d2, case, True, True)
    s = d2.optimize(30)
    assert len(s["iterations"]) == len(g["lm_iterations"])
    np.testing.assert_allclose(s["iterations"][:, 0], g["lm_iterations"][:, 0], rtol=1e-8)
    np.testing.assert_allclose(d2.getWorldColmapTransformation()[0], g["T_wg"], atol=1e-8)
