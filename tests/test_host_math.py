"""The hand-derived analytic Jacobians (csrc/ctba_terms.h, compiled for the host by the tests-only debug
harness tests/hostdbg) against the Jet-autodiff oracle, element-wise, tolerance 1e-10 relative
(north_star).  Runs on CPU so formula regressions are caught without a GPU."""
import ctypes as C

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

from conftest import elemerr

TOL = 1e-10
P = lambda a: a.ctypes.data_as(C.c_void_p)


def _setup(oracle, scene, N, model, rng, small_rotation=False):
    dt_ns = 100_000_000
    step = 1e-9 if small_rotation else 0.1
    q = [Rot.from_rotvec(rng.normal(size=3) * 0.3).as_quat()]
    for i in range(N - 1):
        q.append((Rot.from_quat(q[-1]) * Rot.from_rotvec(rng.normal(size=3) * step)).as_quat())
    if small_rotation:
        q[2] = q[1].copy()  # identical consecutive knots -> log small-angle branch
    q = np.array(q); p = rng.normal(size=(N, 3))
    cam = {4: "euroc", 5: "fpv", 7: "fov"}[model]
    c = scene.camera_preset(cam)
    corr = np.zeros(10); corr[0] = 0.05; corr[1:5] = Rot.from_rotvec([0.01, -0.02, 0.03]).as_quat(); corr[5:8] = [0.01, 0.02, -0.01]
    corr[8] = 0.003; corr[9] = -0.002
    p_bp = np.array([0.1, -0.05, 0.2]); g = np.array([0.1, -0.2, 9.7]); scale = 1.3
    dt_bias = 0.1 * N / 8.0
    o = oracle.OracleSpline(N, dt_ns, 0)
    o.setKnots(q, p); o.setCamera(c["cam_model"], c["intr"], c["dist"], c["q_bc"], c["p_bc"], 0.0)
    o.setGlobals(scale, p_bp, g, 0.0); o.setCorrections(corr)
    cfg = np.array([model] + list(c["intr"]) + list(c["dist"]) + list(c["q_bc"]) + list(c["p_bc"]) + [0.1, dt_bias, scale], dtype=np.float64)
    glb = np.zeros(16); glb[0] = corr[0]; glb[1:5] = corr[1:5]; glb[5:8] = corr[5:8]; glb[8:11] = p_bp; glb[11:14] = g; glb[14] = corr[8]; glb[15] = corr[9]
    return dict(o=o, q=q, p=p, c=c, corr=corr, p_bp=p_bp, g=g, scale=scale, cfg=cfg, glb=glb, dt_ns=dt_ns, dt_bias=dt_bias)


@pytest.mark.parametrize("N", [4, 6])
@pytest.mark.parametrize("model", [4, 5, 7])
@pytest.mark.parametrize("small", [False, True])
def test_terms_vs_oracle(oracle, scene, hostdbg, N, model, small):
    rng = np.random.default_rng(100 * N + model + (7 if small else 0))
    S = _setup(oracle, scene, N, model, rng, small)
    o, q, p, c = S["o"], S["q"], S["p"], S["c"]
    for u in (0.0, 0.37, 0.999999):
        smp = o.sample(0, u)
        Rwb = Rot.from_quat(smp[:4]); pwb = smp[4:7]
        Rbc = Rot.from_quat(c["q_bc"]) * Rot.from_quat(S["corr"][1:5]); pbc = Rot.from_quat(c["q_bc"]).apply(S["corr"][5:8]) + c["p_bc"]
        s = S["scale"] + S["corr"][0]
        n = 9
        pc = np.c_[rng.uniform(-1, 1, n), rng.uniform(-.6, .6, n), rng.uniform(2, 6, n)]
        pw = np.ascontiguousarray((Rwb * Rbc).apply(pc) + (s * Rwb.apply(pbc) + pwb))
        uv = scene.project(c["cam_model"], c["intr"], c["dist"], pc) + rng.normal(size=(n, 2))
        uv[0] += 20.0   # beyond the 10 px gate -> zero residual AND zero derivative
        uv = np.ascontiguousarray(uv)
        o2 = oracle.OracleSpline(N, S["dt_ns"], 0)
        o2.setKnots(q, p); o2.setCamera(c["cam_model"], c["intr"], c["dist"], c["q_bc"], c["p_bc"], 0.0)
        o2.setGlobals(S["scale"], S["p_bp"], S["g"], 0.0); o2.setCorrections(S["corr"]); o2.set3DPoints(pw)
        nb = o2.initializeBiases(S["dt_bias"], [0, 0, 0], [0, 0, 0])
        t_ns = np.array([int(round(u * S["dt_ns"]))], dtype=np.int64)
        assert o2.addFrames(t_ns, [0, n], uv, np.arange(n), 0.8) == 1
        r_o, J_o, dropped = o2.evaluate(0)
        assert dropped == 0.0 and np.abs(J_o[:2]).max() == 0.0 and np.abs(r_o[:2]).max() == 0.0
        r_d = np.zeros(2 * n); J_d = np.zeros((2 * n, 6 * N + 10))
        uu = float(t_ns[0] % S["dt_ns"]) / float(S["dt_ns"])
        hostdbg.dbg_frame_reproj(N, P(S["cfg"]), P(S["glb"]), P(q), P(p), C.c_double(uu), n, P(pw), P(uv), C.c_double(1 / 0.8), P(r_d), P(J_d))
        assert elemerr(r_d, r_o) <= 1.0 and elemerr(J_d, J_o) <= 1.0
        # IMU + GP
        ba = rng.normal(size=(nb, 3)) * 0.05; bg = rng.normal(size=(nb, 3)) * 0.01
        o2.setBiases(ba, bg)
        t_imu = np.array([int(0.61 * S["dt_ns"])], dtype=np.int64)
        acc = rng.normal(size=(1, 3)) + [0, 0, 9.8]; gyr = rng.normal(size=(1, 3)) * 0.3
        assert o2.addImu(t_imu, acc, gyr, 0.1, 0.05) == 1
        gpz = rng.normal(size=(1, 3))
        assert o2.addGp(np.array([int(0.23 * S["dt_ns"])], dtype=np.int64), gpz, 0.1) == 1
        ra_o, Ja_o, d1 = o2.evaluate(1); rg_o, Jg_o, d2 = o2.evaluate(2); rp_o, Jp_o, _ = o2.evaluate(3)
        assert d1 == 0.0 and d2 == 0.0   # bias blocks 4..N-1 registered by the reference have exactly zero Jacobian
        dt_bias_ns = int(np.uint64(S["dt_bias"] / 1e-9)); st = int(t_imu[0]); sb = st // dt_bias_ns; ub = (st % dt_bias_ns) / dt_bias_ns
        if sb + N > nb:
            te = (nb - N + 1) * dt_bias_ns - 1000000; sb = te // dt_bias_ns; ub = (te % dt_bias_ns) / dt_bias_ns
        ra = np.zeros(3); rg = np.zeros(3); Ja = np.zeros((3, 6 * N + 16)); Jg = np.zeros((3, 3 * N + 13))
        hostdbg.dbg_imu(N, P(S["cfg"]), P(S["glb"]), P(q), P(p), P(np.ascontiguousarray(ba[sb:sb + 4])), P(np.ascontiguousarray(bg[sb:sb + 4])),
                        C.c_double(st * 1e-9), C.c_double(0.0), C.c_double(ub), P(acc), P(gyr), C.c_double(10.0), C.c_double(20.0), P(ra), P(rg), P(Ja), P(Jg))
        assert elemerr(ra, ra_o) <= 1.0 and elemerr(rg, rg_o) <= 1.0
        assert elemerr(Ja, Ja_o) <= 1.0 and elemerr(Jg, Jg_o) <= 1.0
        rp = np.zeros(3); Jp = np.zeros((3, 6 * N + 5))
        hostdbg.dbg_gp(N, P(S["cfg"]), P(S["glb"]), P(q), P(p), C.c_double(int(0.23 * S["dt_ns"]) * 1e-9), C.c_double(0.0), P(gpz), C.c_double(10.0), P(rp), P(Jp))
        assert elemerr(rp, rp_o) <= 1.0 and elemerr(Jp, Jp_o) <= 1.0


@pytest.mark.parametrize("N", [4, 6])
@pytest.mark.parametrize("small", [False, True])
def test_pose_prior_terms_vs_oracle(oracle, scene, hostdbg, N, small):
    """prior_eval (ctba_terms.h; PositionCostFunctor / OrientationCostFunctor) compiled for the host against the oracle's Jet autodiff,
    including a measurement that coincides with the spline (zero residual) and one half a turn away."""
    rng = np.random.default_rng(900 + N + (3 if small else 0))
    S = _setup(oracle, scene, N, 4, rng, small)
    q, p = S["q"], S["p"]
    for u in (0.0, 0.41, 0.999999):
        t_ns = np.array([int(round(u * S["dt_ns"]))], dtype=np.int64)
        uu = float(t_ns[0] % S["dt_ns"]) / float(S["dt_ns"])
        smp = S["o"].sample(0, uu)
        for k, dq in enumerate(([0.0, 0.0, 0.0, 1.0], Rot.from_rotvec(rng.normal(size=3) * 0.3).as_quat(), Rot.from_rotvec([0.0, 3.0, 0.5]).as_quat())):
            meas_q = (Rot.from_quat(smp[:4]) * Rot.from_quat(dq)).as_quat()
            meas_p = smp[4:7] + (0.0 if k == 0 else rng.normal(size=3) * 0.2)
            o2 = oracle.OracleSpline(N, S["dt_ns"], 0)
            o2.setKnots(q, p)
            assert o2.addPositionMeasurements(t_ns, meas_p[None, :], 0.05) == 1
            assert o2.addOrientationMeasurements(t_ns, meas_q[None, :], 0.02) == 1
            r_o, J_o, _ = o2.evaluate(6)
            for which, meas4, inv_std, sl in ((0, np.r_[meas_p, 0.0], 20.0, slice(0, 3)), (1, meas_q, 50.0, slice(3, 6))):
                r_d = np.zeros(3); J_d = np.zeros((3, 3 * N))
                assert hostdbg.dbg_prior(N, P(S["cfg"]), P(q), P(p), which, C.c_double(uu), P(np.ascontiguousarray(meas4, dtype=np.float64)),
                                         C.c_double(inv_std), P(r_d), P(J_d)) == 0
                np.testing.assert_allclose(r_d, r_o[sl], rtol=1e-10, atol=1e-11)
                np.testing.assert_allclose(J_d, J_o[sl], rtol=1e-10, atol=1e-10)


@pytest.mark.parametrize("N", [4, 6])
@pytest.mark.parametrize("small", [False, True])
def test_world_alignment_term_vs_oracle(oracle, scene, hostdbg, N, small):
    """align_eval (ctba_terms.h; GlobalPositionSensorWithAlignmentCorrectionCostFunctor over the 11 free tangent directions of the
    alignment problem) compiled for the host against the oracle's Jet autodiff; the oracle's Jacobian w.r.t. the (constant) knots is ignored."""
    rng = np.random.default_rng(1300 + N + (3 if small else 0))
    S = _setup(oracle, scene, N, 4, rng, small)
    q, p = S["q"], S["p"]
    T_wg = np.r_[Rot.from_rotvec([0.4, -1.1, 2.0]).as_quat(), [12.0, -7.0, 3.5]]
    for t_corr, ident in ((0.0, True), (0.004, False), (-0.003, False)):
        corr = np.r_[[0.0, 0.0, 0.0, 1.0] if ident else Rot.from_rotvec(rng.normal(size=3) * 0.2).as_quat(), rng.normal(size=3) * (0.0 if ident else 0.5)]
        scale_corr = 0.0 if ident else 0.07
        o2 = oracle.OracleSpline(N, S["dt_ns"], 0)
        o2.setKnots(q, p)
        o2.setGlobals(S["scale"], S["p_bp"], S["g"], 0.0)
        c10 = np.zeros(10); c10[0] = scale_corr; c10[4] = 1.0; c10[9] = t_corr
        o2.setCorrections(c10)
        o2.setWorldColmapTransformation(T_wg, corr)
        t_ns = np.array([int(0.45 * S["dt_ns"])], dtype=np.int64)
        meas = rng.normal(size=(1, 3)) * 3.0
        assert o2.addWorldColmapAlignmentMeasurements(t_ns, meas, 0.25) == 1
        r_o, J_o, _ = o2.evaluate(7)
        x12 = np.r_[t_corr, scale_corr, corr[:4], corr[4:], S["p_bp"]]
        e_d = np.zeros(3); J_d = np.zeros((3, 11))
        st_s = float(t_ns[0]) * 1e-9
        assert hostdbg.dbg_align(N, P(S["cfg"]), P(T_wg), P(np.ascontiguousarray(x12)), P(q), P(p), C.c_double(st_s), C.c_double(0.0),
                                 P(np.ascontiguousarray(meas[0])), P(e_d), P(J_d)) == 0
        np.testing.assert_allclose(4.0 * e_d, r_o, rtol=1e-10, atol=1e-11)
        np.testing.assert_allclose(4.0 * J_d, J_o[:, 6 * N:], rtol=1e-10, atol=1e-10)
        assert np.abs(J_o[:, 6 * N:]).min(axis=0).max() > 0      # every one of the 11 directions is exercised
