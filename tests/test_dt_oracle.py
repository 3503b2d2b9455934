"""The oracle's restatement of the discrete-time estimator (oracle/dt_problem.h; BASELINE configs[3]) pinned by
independent checks — CPU only:
  * SE3 plus / plus-Jacobian (se3.hpp:135-204, 761-782) against each other by central differences;
  * the Jet Jacobians of the reprojection and global-position functors against central differences of the residuals THROUGH the SE3 plus;
  * the IMU factor's MINIMAL Jacobians (imu_cost_functor.h:667-856, `jacobians_minimal`) against central differences of its residual under
    the perturbation they are derived for (OKVIS: position additive in W, rotation multiplied from the left by deltaQ) — which pins the
    transcription of redoPreintegration / F0 / F1;
  * the re-integration rule (|delta b_g| * dt > 1e-4, :693-704);
  * and the layout hazard of SURVEY.md §8(a): the 15 x 7 blocks the factor hands to Ceres are built by liftJacobian for the layout [t, q]
    while the Sophus block is [q, t], so the LOCAL Jacobian the reference computes is NOT the derivative of its residual.  The oracle keeps it
    (parity is on what the reference computes); this test fails if somebody "repairs" it."""
import importlib

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot


def _setup(oracle, pkg, scene, **kw):
    sc = scene.make_scene(order=6, duration_s=2.5, n_landmarks=90, track_len=8, imu_hz=200, gp_hz=10, cam_hz=10.0, seed=5,
                          flags=(True, True, False, True, True), **kw)
    d = scene.dt_scene_from(sc, perturb=(0.002, 0.05))
    o = oracle.OracleDiscrete()
    pkg.load_dt_scene(o, d)
    o.setCorrections(np.array([0.002, 0.001, -0.002, 0.0015, np.sqrt(1 - 0.001**2 - 0.002**2 - 0.0015**2), 0.002, -0.003, 0.001, 0.0004]), -0.0003)
    return o, d


def test_se3_plus_jacobian_matches_plus(oracle):
    rng = np.random.default_rng(0)
    for _ in range(5):
        T = np.r_[Rot.from_rotvec(rng.normal(size=3)).as_quat(), rng.normal(size=3) * 3]
        J = oracle.se3_plus_jacobian(T)
        Jn = np.zeros((7, 6)); h = 1e-6
        for k in range(6):
            e = np.zeros(6); e[k] = h
            Jn[:, k] = (oracle.se3_plus(T, e) - oracle.se3_plus(T, -e)) / (2 * h)
        np.testing.assert_allclose(J, Jn, atol=2e-9)
    # exp of a finite twist: rotation part = SO3 exp, translation part = V(omega) upsilon (closed form via scipy)
    T = np.r_[0, 0, 0, 1.0, 0, 0, 0]; d = np.array([0.3, -0.2, 0.5, 0.4, 0.1, -0.7])
    out = oracle.se3_plus(T, d)
    th = np.linalg.norm(d[3:]); K = np.array([[0, -d[5], d[4]], [d[5], 0, -d[3]], [-d[4], d[3], 0]])
    V = np.eye(3) + (1 - np.cos(th)) / th**2 * K + (th - np.sin(th)) / th**3 * K @ K
    np.testing.assert_allclose(out[4:], V @ d[:3], atol=1e-14)
    assert (Rot.from_quat(out[:4]).inv() * Rot.from_rotvec(d[3:])).magnitude() < 1e-14


def _numeric_local(o, kind, rows, width_blocks, eval_res):
    """central differences of the residual rows w.r.t. the local coordinates of the listed blocks: ('pose', i) | ('sb', i) | ('glob', name)"""
    cols = []
    P0, sb0 = o.getStates()
    g0 = o.getGlobals()
    h = 1e-6
    for blk in width_blocks:
        n = {"pose": 6, "sb": 9}.get(blk[0], None)
        if blk[0] == "glob":
            n = {"scale": 1, "R_bc": 3, "p_bc": 3, "t_cam_imu": 1, "t_gs": 1, "p_bp": 3}[blk[1]]
        for k in range(n):
            res = []
            for sgn in (+1, -1):
                P, sb, g = P0.copy(), sb0.copy(), g0.copy()
                if blk[0] == "pose":
                    e = np.zeros(6); e[k] = sgn * h
                    P[blk[1]] = oracle_mod.se3_plus(P0[blk[1]], e)
                elif blk[0] == "sb":
                    sb[blk[1], k] += sgn * h
                else:
                    name = blk[1]
                    if name == "scale": g[0] += sgn * h
                    elif name == "R_bc":
                        e = np.zeros(3); e[k] = sgn * h
                        g[1:5] = (Rot.from_quat(g0[1:5]) * Rot.from_rotvec(e)).as_quat()
                    elif name == "p_bc": g[5 + k] += sgn * h
                    elif name == "t_cam_imu": g[8] += sgn * h
                    elif name == "t_gs": g[9] += sgn * h
                    elif name == "p_bp": g[10 + k] += sgn * h
                o.initializeStatesKeep(P, sb)
                o.setCorrections(g[:9], g[9]); o.setPbp(g[10:13])
                res.append(eval_res()[rows])
            cols.append((res[0] - res[1]) / (2 * h))
    o.initializeStatesKeep(P0, sb0); o.setCorrections(g0[:9], g0[9]); o.setPbp(g0[10:13])
    return np.stack(cols, axis=1)


oracle_mod = None


def test_reprojection_and_gp_jacobians_vs_central_differences(oracle, pkg, scene):
    global oracle_mod
    oracle_mod = oracle
    o, d = _setup(oracle, pkg, scene)
    # ---- reprojection: observation q of frame f ----
    r, J = o.evaluate(0)
    ob = d["obs_begin"]
    checked = 0
    for f in (3, 9, 15):
        for q in range(ob[f], min(ob[f] + 3, ob[f + 1])):
            rows = slice(2 * q, 2 * q + 2)
            if np.abs(r[rows]).max() == 0:
                continue
            Jn = _numeric_local(o, 0, rows, [("pose", f), ("glob", "scale"), ("glob", "R_bc"), ("glob", "p_bc"), ("glob", "t_cam_imu")], lambda: o.evaluate(0, jac=False)[0])
            np.testing.assert_allclose(J[rows, :14], Jn, rtol=2e-6, atol=2e-5 * np.abs(J[rows]).max())
            checked += 1
    assert checked >= 4
    # ---- global position ----
    r, J = o.evaluate(2)
    for i in (0, 5, 11):
        f0 = int(d["gp_frame0"][i])
        rows = slice(3 * i, 3 * i + 3)
        Jn = _numeric_local(o, 2, rows, [("pose", f0), ("pose", f0 + 1), ("glob", "scale"), ("glob", "t_gs"), ("glob", "p_bp")], lambda: o.evaluate(2, jac=False)[0])
        np.testing.assert_allclose(J[rows], Jn, rtol=2e-6, atol=2e-6 * np.abs(J[rows]).max())


def _okvis_perturb(T, dp, da):
    """OKVIS minimal perturbation of a pose: p + dp, deltaQ(da) * q"""
    q = (Rot.from_rotvec(da) * Rot.from_quat(T[:4])).as_quat()
    return np.r_[q, T[4:] + dp]


def test_imu_factor_minimal_jacobians_and_reintegration(oracle, pkg, scene):
    o, d = _setup(oracle, pkg, scene)
    k = 4
    f0 = d["imu"][k][0]
    r0, Jmin = o.imuMinimal(k)
    assert o.imuRedoCount(k) == 1 and np.isfinite(r0).all()
    P0, sb0 = o.getStates()
    h = 1e-7
    Jn = np.zeros((15, 30))
    for side, f in ((0, f0), (1, f0 + 1)):
        for c in range(15):
            res = []
            for sgn in (+1, -1):
                P, sb = P0.copy(), sb0.copy()
                if c < 3:
                    e = np.zeros(3); e[c] = sgn * h; P[f] = _okvis_perturb(P0[f], e, np.zeros(3))
                elif c < 6:
                    e = np.zeros(3); e[c - 3] = sgn * h; P[f] = _okvis_perturb(P0[f], np.zeros(3), e)
                else:
                    sb[f, c - 6] += sgn * h
                o.initializeStatesKeep(P, sb)
                res.append(o.imuMinimal(k)[0])
            Jn[:, 15 * side + c] = (res[0] - res[1]) / (2 * h)
    o.initializeStatesKeep(P0, sb0)
    assert o.imuRedoCount(k) == 1                      # |delta b_g| dt stayed far below 1e-4: the cached pre-integration was reused throughout
    scale = np.abs(Jmin).max(axis=0, keepdims=True)
    assert (np.abs(Jmin - Jn) <= 2e-5 * scale + 1e-6 * np.abs(Jmin).max()).all(), np.abs(Jmin - Jn).max()
    # ---- re-integration: a gyro-bias change of more than 1e-4 / dt redoes the pre-integration, a smaller one does not ----
    dt = d["cam_ts"][f0 + 1] - d["cam_ts"][f0]
    sb = sb0.copy(); sb[f0, 3] += 0.5e-4 / dt
    o.initializeStatesKeep(P0, sb); o.imuMinimal(k)
    assert o.imuRedoCount(k) == 1
    sb[f0, 3] += 1.0e-4 / dt
    o.initializeStatesKeep(P0, sb); r_redo, _ = o.imuMinimal(k)
    assert o.imuRedoCount(k) == 2
    # after the redo the residual is the exact (re-integrated) one; the first-order bias correction before it was close
    o2, _ = _setup(oracle, pkg, scene)
    o2.initializeStatesKeep(P0, sb)
    np.testing.assert_allclose(o2.imuMinimal(k)[0], r_redo, rtol=1e-9, atol=1e-9)


def test_imu_factor_local_jacobian_is_what_the_reference_computes_not_the_derivative(oracle, pkg, scene):
    global oracle_mod
    oracle_mod = oracle
    o, d = _setup(oracle, pkg, scene)
    k = 4
    f0 = d["imu"][k][0]
    r, J = o.evaluate(1)
    rows = slice(15 * k, 15 * k + 15)
    Jn = _numeric_local(o, 1, rows, [("pose", f0), ("sb", f0), ("pose", f0 + 1), ("sb", f0 + 1)], lambda: o.evaluate(1, jac=False)[0])
    # speed / bias columns are plain Euclidean blocks: they ARE the derivative
    np.testing.assert_allclose(J[rows, 6:15], Jn[:, 6:15], rtol=1e-4, atol=1e-5 * np.abs(Jn).max())
    np.testing.assert_allclose(J[rows, 21:30], Jn[:, 21:30], rtol=1e-4, atol=1e-5 * np.abs(Jn).max())
    # pose columns: J_minimal * liftJacobian([q, t] read as [t, q]) * Dx_this_mul_exp_x_at_0 — not the derivative (SURVEY §8(a), D-row hazard)
    assert np.abs(J[rows, :6] - Jn[:, :6]).max() > 1e-2 * np.abs(Jn[:, :6]).max()


@pytest.mark.parametrize("shape", ["gvi", "gv"])
def test_sparse_lm_equals_dense_lm(oracle, pkg, scene, shape):
    """the sparse LM that bench.py's CPU legs time on the discrete-time problem (oracle/sparse_lm.h with 15-scalar state nodes) against the
    dense LM of lm.h: same accept / reject sequence, per-iteration cost 1e-9, same final states"""
    res = []
    for sparse in (False, True):
        sc = scene.make_scene(order=6, duration_s=3.0, n_landmarks=120, track_len=8, imu_hz=200, gp_hz=10, cam_hz=10.0, seed=5, flags=(False, True, False, True, True))
        d = scene.dt_scene_from(sc, perturb=(0.005, 0.1) if shape == "gvi" else (0.02, 0.5))
        o = oracle.OracleDiscrete()
        pkg.load_dt_scene(o, d, with_imu=shape == "gvi")
        s = o.optimizeSparse(8, 4) if sparse else o.optimize(8, 4)
        res.append((s, o.getStates(), o.get3DPoints(), o.getGlobals()))
    (sa, (Pa, sba), la, ga), (sb_, (Pb, sbb), lb, gb) = res
    assert sa["iterations"].shape == sb_["iterations"].shape
    assert (sa["iterations"][:, 6] == sb_["iterations"][:, 6]).all() and sa["num_successful_steps"] >= 2
    np.testing.assert_allclose(sb_["iterations"][:, 0], sa["iterations"][:, 0], rtol=1e-9)
    np.testing.assert_allclose(Pb, Pa, atol=1e-8); np.testing.assert_allclose(sbb, sba, atol=1e-7)
    np.testing.assert_allclose(lb, la, atol=1e-7); np.testing.assert_allclose(gb, ga, atol=1e-8)
