"""Spline sampling getters (getPose / getAccel / getGyro / get*Bias, bspline_so3r3.h:122-327; traceSpline, save.h:99-121):
the oracle's restatement against its own fixed-(s, u) evaluation on CPU, and the device kernel against the oracle on the GPU."""
import numpy as np
import pytest

KINDS = range(7)


def _times(sc, rng, n, bias=False, n_knots=None, dt_ns_bias=None):
    N = sc["order"]
    if bias:
        end = (n_knots - N + 1) * dt_ns_bias + sc["start_t_ns"]
    else:
        end = (sc["knot_p"].shape[0] - N + 1) * sc["dt_ns"] + sc["start_t_ns"]
    t = rng.integers(sc["start_t_ns"], end, n, dtype=np.int64)
    t[0] = sc["start_t_ns"]; t[1] = end - 1                      # both ends of the validity range
    t[2] = sc["start_t_ns"] + 3 * sc["dt_ns"]                    # exactly on a knot boundary (u = 0)
    return t


@pytest.mark.parametrize("order", [4, 6])
def test_oracle_getters_consistent_with_fixed_u_evaluation(pkg, scene, oracle, order):
    rng = np.random.default_rng(order)
    sc = scene.make_scene(order=order, duration_s=3.0, n_landmarks=30, track_len=5, imu_hz=100, gp_hz=0, dt_s_bias=0.25, seed=2)
    o = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(o, sc)
    t = _times(sc, rng, 40)
    pose, acc_b, acc_w, gyr_b, gyr_w = (o.sampleAt(k, t) for k in range(5))
    g = np.asarray(sc["g"])
    for i, ti in enumerate(t):
        st = int(ti - sc["start_t_ns"]); s, u = st // sc["dt_ns"], (st % sc["dt_ns"]) / sc["dt_ns"]
        ref = o.sample(s, u)      # q(4) p(3) w(3) wd(3) vel(3) acc(3)
        np.testing.assert_allclose(pose[i], np.concatenate([ref[4:7], ref[0:4]]), rtol=0, atol=1e-14)
        np.testing.assert_allclose(gyr_b[i], ref[7:10], rtol=0, atol=1e-13)
        np.testing.assert_allclose(acc_w[i], ref[16:19] + g, rtol=0, atol=1e-12)
        q = ref[0:4]; R = np.array(oracle.so3_matrix(q)).reshape(3, 3)
        np.testing.assert_allclose(acc_b[i], R.T @ (ref[16:19] + g), rtol=0, atol=1e-11)
        np.testing.assert_allclose(gyr_w[i], R.T @ ref[7:10], rtol=0, atol=1e-12)
    with pytest.raises(ValueError):
        o.sampleAt(0, np.array([sc["start_t_ns"] - 1]))
    with pytest.raises(ValueError):
        o.sampleAt(0, np.array([t[1] + 1]))


@pytest.mark.gpu
@pytest.mark.parametrize("order", [4, 6])
def test_device_sampling_matches_oracle(pkg, scene, oracle, order):
    rng = np.random.default_rng(10 + order)
    sc = scene.make_scene(order=order, duration_s=4.0, n_landmarks=60, track_len=6, imu_hz=100, gp_hz=10, dt_s_bias=0.25, seed=5)
    gpu = pkg.BSplineSO3R3(sc["order"], sc["dt_ns"], sc["start_t_ns"], device=0)
    cpu = oracle.OracleSpline(sc["order"], sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(gpu, sc); pkg.load_scene(cpu, sc)
    # non-trivial biases
    nb = gpu.numBiasKnots()
    ba, bg = rng.normal(size=(nb, 3)) * 0.1, rng.normal(size=(nb, 3)) * 0.01
    gpu.setBiases(ba, bg); cpu.setBiases(ba, bg)
    t = _times(sc, rng, 500)
    dt_ns_bias = int(0.25 / 1e-9)
    tb = _times(sc, rng, 200, bias=True, n_knots=nb, dt_ns_bias=dt_ns_bias)
    for kind in KINDS:
        tt = tb if kind >= pkg.CTBA_SAMPLE_ACCEL_BIAS else t
        a, b = gpu.sampleAt(kind, tt), cpu.sampleAt(kind, tt)
        assert a.shape == b.shape
        np.testing.assert_allclose(a, b, rtol=1e-10, atol=1e-10, err_msg=f"kind {kind}")
    # after an optimisation the samples follow the device state without an explicit read-back
    gpu.optimize(3); cpu.setKnots(*gpu.getKnots())
    np.testing.assert_allclose(gpu.sampleAt(pkg.CTBA_SAMPLE_POSE, t), cpu.sampleAt(0, t), rtol=1e-10, atol=1e-10)
    tr = gpu.traceSpline(2_000_000)         # 500 Hz, like the drivers (optimize_continuous_time.cpp)
    assert tr.shape[1] == 8 and tr[0, 0] == sc["start_t_ns"] * 1e-9 and tr[-1, 0] < gpu.getEndTns() * 1e-9
    np.testing.assert_allclose(np.linalg.norm(tr[:, 4:8], axis=1), 1.0, atol=1e-12)
    with pytest.raises(pkg.CtbaError):
        gpu.sampleAt(pkg.CTBA_SAMPLE_POSE, np.array([sc["start_t_ns"] - 1]))
    with pytest.raises(pkg.CtbaError):
        gpu.sampleAt(pkg.CTBA_SAMPLE_GYRO_BODY, np.array([gpu.getEndTns() + 1_000_000]))
