"""Golden vectors (tests/golden/*.npz, generator tests/golden/make_golden.py).

CPU part: (1) the oracle still reproduces the committed vectors; (2) an INDEPENDENT numpy/scipy statement of the residuals
(general-order cumulative blending matrix from the binomial formula, scipy rotations, angular velocity by central
differences of the rotation spline, camera models from scene.project) agrees with the committed residuals, which pins the
conventions (time offsets, gravity sign, T_b_c composition, gate, weights) independently of the oracle's code.
GPU part: the CUDA path reproduces the committed vectors through the C ABI."""
import importlib
import os
import sys
from math import comb, factorial

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden  # noqa: E402

CASES = list(make_golden.CASES)


def _load(name):
    return np.load(os.path.join(HERE, "golden", name + ".npz"))


def _rel(a, b):
    return 0.0 if b.size == 0 else float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


@pytest.mark.parametrize("name", CASES)
def test_oracle_reproduces_golden(oracle, pkg, name):
    g = _load(name)
    sc, iters = make_golden.scene_of(name)
    fp = np.array([sc["uv"].sum(), sc["knot_q"].sum(), sc["landmarks"].sum(), float(sc["uv"].shape[0])])
    np.testing.assert_allclose(fp, g["scene_fingerprint"], rtol=1e-13)
    o = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(o, sc)
    for kind in range(6):
        r, J, _ = o.evaluate(kind)
        assert r.shape == g[f"r{kind}"].shape
        assert _rel(r, g[f"r{kind}"]) < 1e-12 and _rel(J, g[f"J{kind}"]) < 1e-12
    np.testing.assert_allclose(o.cost(), g["cost"], rtol=1e-12)
    s = o.optimize(iters, 2)
    np.testing.assert_allclose(s["iterations"][:, 0], g["lm_iterations"][:, 0], rtol=1e-9)
    assert (s["iterations"][:, 6] == g["lm_iterations"][:, 6]).all()


# ---------------- independent statement of the residuals ----------------
def _blend(N, cumulative):
    """General-order uniform B-spline blending matrix (row = knot, column = power of u), Sommer et al. / spline_common.h semantics."""
    M = np.zeros((N, N))
    for s in range(N):
        for n in range(N):
            acc = 0.0
            for l in range(s, N):
                acc += (-1) ** (l - s) * comb(N, l - s) * (N - 1 - l) ** (N - 1 - n)
            M[s, n] = comb(N - 1, n) * acc / factorial(N - 1)
    if cumulative:
        M = np.cumsum(M[::-1], axis=0)[::-1]
    return M


def _pose(sc, q, p, s, u):
    N = sc["order"]
    pw = u ** np.arange(N)
    lam = _blend(N, True) @ pw
    c = _blend(N, False) @ pw
    R = Rot.from_quat(q[s])
    for i in range(1, N):
        d = (Rot.from_quat(q[s + i - 1]).inv() * Rot.from_quat(q[s + i])).as_rotvec()
        R = R * Rot.from_rotvec(lam[i] * d)
    return R, c @ p[s:s + N]


def _window(sc, st_ns):
    return int(st_ns // sc["dt_ns"]), float(st_ns % sc["dt_ns"]) / sc["dt_ns"]


@pytest.mark.parametrize("name", CASES)
def test_independent_residuals_match_golden(scene, name):
    g = _load(name)
    sc, _ = make_golden.scene_of(name)
    N, q, p = sc["order"], sc["knot_q"], sc["knot_p"]
    K = q.shape[0]
    R_bc, p_bc, s_ = Rot.from_quat(sc["q_bc"]), sc["p_bc"], sc["scale"]
    # --- reprojection (reprojection_cost_functor.h:47-109): every accepted frame, gate at 100 px^2, weight 1/sigma
    res = []
    for f, t in enumerate(sc["frame_t_ns"]):
        st = int(t - sc["start_t_ns"])
        b, e = sc["obs_begin"][f], sc["obs_begin"][f + 1]
        if st < 0 or e == b:
            continue
        s, u = _window(sc, st)
        if s + N > K:
            continue
        R, pos = _pose(sc, q, p, s, u)
        R_wc = R * R_bc
        p_wc = s_ * R.apply(p_bc) + pos
        pc = R_wc.inv().apply(sc["landmarks"][sc["lm_index"][b:e]] - p_wc)
        err = scene.project(sc["cam_model"], sc["intr"], sc["dist"], pc) - sc["uv"][b:e]
        err[(err ** 2).sum(axis=1) > 100.0] = 0.0
        res.append(err / sc["sigma_reproj_px"])
    r0 = np.concatenate(res).ravel()
    assert r0.shape == g["r0"].shape
    assert _rel(r0, g["r0"]) < 1e-9
    # --- global position (global_position_cost_functor.h:45-73)
    res = []
    for t, z in zip(sc["gp_t_ns"], sc["gp_pos"]):
        st = int(t - sc["start_t_ns"] - int(sc["t_s_gp"] * 1e9))
        if st < 0:
            continue
        s, u = _window(sc, st)
        if s + N > K:
            continue
        R, pos = _pose(sc, q, p, s, u)
        res.append((s_ * R.apply(sc["p_bp"]) + pos - z) / sc["gp_std"])
    r3 = np.concatenate(res) if res else np.zeros(0)
    assert r3.shape == g["r3"].shape and _rel(r3, g["r3"]) < 1e-9
    # --- IMU (imu_cost_functor.h:48-139): zero initial biases; gravity ADDED to the world acceleration; t_imu = t_cam + offset
    ra, rg = [], []
    dt = sc["dt_ns"] * 1e-9
    for t, am, gm in zip(sc["imu_t_ns"], sc["imu_acc"], sc["imu_gyr"]):
        st = int(t + int(-sc["t_offset_cam_imu"] * 1e9) - sc["start_t_ns"])
        if st < 0:
            continue
        s, u = _window(sc, st)
        if s + N > K:
            continue
        R, _ = _pose(sc, q, p, s, u)
        h = 1e-4
        Rm, _ = _pose(sc, q, p, s, u - h)   # the polynomial piece extends smoothly, so u +- h may leave [0,1)
        Rp, _ = _pose(sc, q, p, s, u + h)
        w = (Rm.inv() * Rp).as_rotvec() / (2 * h * dt)
        c2 = (_blend(N, False) @ np.array([n * (n - 1) * u ** (n - 2) if n >= 2 else 0.0 for n in range(N)])) / dt ** 2
        a_w = c2 @ p[s:s + N]
        ra.append((R.inv().apply(a_w + sc["g"]) - am) / sc["sigma_acc_c"])
        rg.append((w - gm) / sc["sigma_omega_c"])
    if ra:
        assert np.concatenate(ra).shape == g["r1"].shape
        assert _rel(np.concatenate(ra), g["r1"]) < 1e-9
        assert _rel(np.concatenate(rg), g["r2"]) < 1e-6     # central differences of the rotation spline
    else:
        assert g["r1"].size == 0
    # --- gravity norm (gravity_norm_cost_functor.h:26-39)
    if sc["gravity_norm"]:
        assert abs(g["r5"][0] - 1e6 * (sc["g"] @ sc["g"] - 9.80665 ** 2)) <= 1e-3
    # --- cost = 1/2 sum rho: Cauchy(3) per FRAME block on reprojection, trivial loss elsewhere
    c0 = 0.0
    off = 0
    for f, t in enumerate(sc["frame_t_ns"]):
        st = int(t - sc["start_t_ns"])
        n = sc["obs_begin"][f + 1] - sc["obs_begin"][f]
        if st < 0 or n == 0 or _window(sc, st)[0] + N > K:
            continue
        sq = (g["r0"][off:off + 2 * n] ** 2).sum()
        off += 2 * n
        c0 += 0.5 * 9.0 * np.log1p(sq / 9.0)
    assert abs(c0 - g["cost"][0]) <= 1e-10 * max(c0, 1.0)
    for kind in (1, 2, 3, 4, 5):
        assert abs(0.5 * (g[f"r{kind}"] ** 2).sum() - g["cost"][kind]) <= 1e-10 * max(g["cost"][kind], 1.0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_cuda_reproduces_golden(pkg, name):
    g = _load(name)
    sc, iters = make_golden.scene_of(name)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    pkg.load_scene(gpu, sc)
    for kind in range(6):
        r, J = gpu.evaluate(kind)
        assert r.shape == g[f"r{kind}"].shape
        assert _rel(r, g[f"r{kind}"]) < 1e-10 and _rel(J, g[f"J{kind}"]) < 1e-10, (kind, _rel(r, g[f"r{kind}"]), _rel(J, g[f"J{kind}"]))
    np.testing.assert_allclose(gpu.cost(), g["cost"], rtol=1e-8, atol=1e-300)
    s = gpu.optimize(iters)
    assert s["iterations"].shape == g["lm_iterations"].shape
    np.testing.assert_allclose(s["iterations"][:, 0], g["lm_iterations"][:, 0], rtol=1e-8)    # per-iteration cost
    assert (s["iterations"][:, 6] == g["lm_iterations"][:, 6]).all()
    qg, pg = gpu.getKnots()
    assert np.abs(pg - g["knot_p"]).max() < 1e-3
    assert np.rad2deg((Rot.from_quat(qg).inv() * Rot.from_quat(g["knot_q"])).magnitude().max()) < 0.01
    np.testing.assert_allclose(gpu.get3DPoints(), g["landmarks"], atol=1e-3)
