"""The oracle's sparse LM (oracle/sparse_lm.h: landmark Schur complement + bordered band Cholesky on host threads — the CPU path that
bench.py times as the reference arm and that gives LM traces at sizes the dense solver cannot reach) against the oracle's dense LM
(oracle/lm.h, the full normal equations with landmarks, another elimination order): same accept / reject sequence, per-iteration cost
1e-9, final state far inside the 1 mm / 0.01 deg gate.  CPU only."""
import numpy as np
import pytest

CASES = {
    "gvi_n6": dict(order=6, duration_s=3.0, n_landmarks=120, track_len=10, imu_hz=100, gp_hz=10, dt_s_bias=0.25, seed=3,
                   flags=(True, True, True, True, True), gravity_norm=True, outlier_frac=0.03),
    "gv_n4": dict(order=4, duration_s=3.0, n_landmarks=120, track_len=10, imu_hz=0, gp_hz=10, seed=3, flags=(False, True, False, True, True)),
    "vi_n6_fixed_globals": dict(order=6, duration_s=2.5, n_landmarks=90, track_len=8, imu_hz=100, gp_hz=0, dt_s_bias=0.25, seed=8,
                                flags=(False, False, False, False, False)),
    "gi_n4_no_frames": dict(order=4, duration_s=3.0, n_landmarks=10, track_len=4, imu_hz=100, gp_hz=20, dt_s_bias=0.25, seed=5,
                            flags=(False, True, True, True, True), gravity_norm=True),
}


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("threads", [1, 4])
def test_sparse_lm_equals_dense_lm(oracle, pkg, scene, name, threads):
    kw = dict(CASES[name])
    sc = scene.make_scene(**kw)
    with_frames = name != "gi_n4_no_frames"
    a = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"]); pkg.load_scene(a, sc, with_frames=with_frames)
    b = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"]); pkg.load_scene(b, sc, with_frames=with_frames)
    sa = a.optimize(6, 2)
    sb = b.optimizeSparse(6, threads)
    ia, ib = sa["iterations"], sb["iterations"]
    assert ia.shape == ib.shape and sa["termination_type"] == sb["termination_type"]
    assert (ia[:, 6] == ib[:, 6]).all()                                  # same accept / reject decisions
    np.testing.assert_allclose(ib[:, 0], ia[:, 0], rtol=1e-9)            # per-iteration cost
    np.testing.assert_allclose(ib[1:, 3], ia[1:, 3], rtol=1e-6, atol=1e-12)   # step norms
    np.testing.assert_allclose(ib[:, 5], ia[:, 5], rtol=1e-6)            # trust-region radii
    (qa, pa), (qb, pb) = a.getKnots(), b.getKnots()
    assert np.abs(pa - pb).max() < 1e-5 and np.abs(qa - qb).max() < 1e-5
    assert np.abs(a.get3DPoints() - b.get3DPoints()).max() < 1e-5
    np.testing.assert_allclose(b.getGlobals(), a.getGlobals(), atol=1e-6)
    t = sb["timings"]
    assert t["solves"] >= 1 and t["n_band"] % 6 == 0 and t["half_bandwidth"] > 0


def test_sparse_lm_runs_where_dense_cannot(oracle, pkg, scene):
    """20 s, 400 landmarks: the cost must decrease monotonically over accepted steps and the thread count must not change the trace."""
    sc = scene.make_scene(order=6, duration_s=20.0, n_landmarks=400, track_len=12, imu_hz=100, gp_hz=10, dt_s_bias=0.5, seed=21,
                          flags=(False, True, True, True, True), gravity_norm=True, outlier_frac=0.02)
    logs = []
    for threads in (1, 6):
        o = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"]); pkg.load_scene(o, sc)
        s = o.optimizeSparse(4, threads)
        logs.append(s["iterations"])
        acc = s["iterations"][s["iterations"][:, 6] == 1, 0]
        assert acc.size >= 2 and (np.diff(acc) < 0).all()
    assert (logs[0][:, 6] == logs[1][:, 6]).all()
    np.testing.assert_allclose(logs[1][:, 0], logs[0][:, 0], rtol=1e-10)
