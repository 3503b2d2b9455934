"""Pose priors of the spline-fitting tools (addPositionMeasurement / addOrientationMeasurement, bspline_so3r3.h:1048-1123;
fit_spline_to_colmap.cpp:139-230): oracle functors against an independent numpy statement on CPU, device residuals / Jacobians / LM trace
against the oracle on the GPU."""
import numpy as np
import pytest


def _fit_problem(scene, order, seed, n_meas=400, noise_p=0.01, noise_r=0.005):
    """a spline initialised off the ground truth + position / orientation samples of the ground-truth trajectory"""
    rng = np.random.default_rng(seed)
    sc = scene.make_scene(order=order, duration_s=5.0, n_landmarks=10, track_len=3, imu_hz=0, gp_hz=0, seed=seed, perturb=(0.05, 2.0, 0.0))
    end = (sc["knot_p"].shape[0] - order + 1) * sc["dt_ns"] + sc["start_t_ns"]
    t = np.sort(rng.integers(sc["start_t_ns"], end, n_meas, dtype=np.int64))
    t = np.concatenate([[sc["start_t_ns"] - 10], t, [end + 5]])          # one before the start, one beyond the end: skipped
    return sc, t, rng


def _quat_mul(a, b):
    ax, ay, az, aw = a; bx, by, bz, bw = b
    return np.array([aw * bx + ax * bw + ay * bz - az * by, aw * by + ay * bw + az * bx - ax * bz, aw * bz + az * bw + ax * by - ay * bx,
                     aw * bw - ax * bx - ay * by - az * bz])


def _measurements(oracle, sc, t, rng, noise_p, noise_r):
    gt = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    gt.setKnots(sc["gt_knot_q"], sc["gt_knot_p"])
    inside = t[1:-1]
    pose = gt.sampleAt(0, inside)
    pos = np.zeros((t.size, 3)); quat = np.tile([0.0, 0.0, 0.0, 1.0], (t.size, 1))
    pos[1:-1] = pose[:, 0:3] + rng.normal(size=(inside.size, 3)) * noise_p
    for i in range(inside.size):
        w = rng.normal(size=3) * noise_r
        th = np.linalg.norm(w)
        dq = np.concatenate([np.sin(th / 2) * w / th, [np.cos(th / 2)]])
        quat[1 + i] = _quat_mul(pose[i, 3:7], dq)
    return pos, quat


@pytest.mark.parametrize("order", [4, 6])
def test_oracle_prior_functors_against_numpy(scene, oracle, order):
    sc, t, rng = _fit_problem(scene, order, 3)
    pos, quat = _measurements(oracle, sc, t, rng, 0.01, 0.005)
    o = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    o.setKnots(sc["knot_q"], sc["knot_p"])
    assert o.addPositionMeasurements(t, pos, 0.02) == t.size - 2          # the two outside the spline are skipped (:1052, :1059)
    assert o.addOrientationMeasurements(t, quat, 0.01) == t.size - 2
    r, J, _ = o.evaluate(6)
    n = t.size - 2
    pose = o.sampleAt(0, t[1:-1])
    np.testing.assert_allclose(r[:3 * n].reshape(n, 3), (pose[:, :3] - pos[1:-1]) / 0.02, rtol=1e-12, atol=1e-12)
    for i in range(0, n, 37):
        mi = quat[1 + i] * np.array([-1, -1, -1, 1.0])
        dr = _quat_mul(pose[i, 3:7], mi)
        dr /= np.linalg.norm(dr)
        np.testing.assert_allclose(r[3 * (n + i):3 * (n + i) + 3], 2 * dr[:3] / 0.01, rtol=1e-11, atol=1e-11)
    # position Jacobian: blending coefficient times identity
    Jp = J[:3 * n].reshape(n, 3, order, 3)
    assert np.abs(Jp[:, 0, :, 1]).max() == 0 and np.allclose(Jp[:, 0, :, 0].sum(axis=1), 1 / 0.02)
    c = o.costTerms()
    assert np.isclose(c[7], 0.5 * np.sum(r * r)) and np.isclose(c[6], c[7])


@pytest.mark.gpu
@pytest.mark.parametrize("order", [4, 6])
def test_prior_residuals_jacobians_and_fit_match_oracle(pkg, scene, oracle, order):
    sc, t, rng = _fit_problem(scene, order, 11 + order)
    pos, quat = _measurements(oracle, sc, t, rng, 0.01, 0.005)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    for h in (gpu, cpu):
        h.setKnots(sc["knot_q"], sc["knot_p"])
        assert h.addPositionMeasurements(t, pos, 1.0) == t.size - 2        # fit_spline_to_colmap passes noise 1.0
        assert h.addOrientationMeasurements(t, quat, 1.0) == t.size - 2
    rg, Jg = gpu.evaluate(pkg.CTBA_PRIOR)
    rc, Jc, _ = cpu.evaluate(6)
    np.testing.assert_allclose(rg, rc, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(Jg, Jc, rtol=1e-10, atol=1e-12)
    np.testing.assert_allclose(gpu.costTerms(), cpu.costTerms(), rtol=1e-12)
    e = gpu.evaluateErrors(pkg.CTBA_PRIOR)
    np.testing.assert_allclose(e, np.linalg.norm(rc.reshape(-1, 3), axis=1), rtol=1e-10, atol=1e-13)
    sg, sc_ = gpu.optimize(8), cpu.optimize(8)
    ig, ic = sg["iterations"], sc_["iterations"]
    assert ig.shape == ic.shape
    np.testing.assert_allclose(ig[:, 0], ic[:, 0], rtol=1e-8)                     # per-iteration cost
    assert (ig[:, 6] == ic[:, 6]).all()                                            # accept / reject sequence
    assert ig[-1, 0] < 0.2 * ig[0, 0]                                              # what is left is the measurement noise
    qg, pg = gpu.getKnots(); qc, pc = cpu.getKnots()
    np.testing.assert_allclose(pg, pc, atol=1e-3)
    assert np.abs(np.abs(np.sum(qg * qc, axis=1)) - 1).max() < 1e-8
    # the fitted spline follows the measurements
    fit = gpu.sampleAt(pkg.CTBA_SAMPLE_POSE, t[1:-1])
    assert np.abs(fit[:, :3] - pos[1:-1]).max() < 0.08
