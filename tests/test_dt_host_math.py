"""The closed-form discrete-time terms of the product (csrc/ctba_dt_math.h: reprojection and global-position residuals with their analytic
local Jacobians, the IMU pre-integration factor with its cache and re-integration rule, SE3 plus), compiled for the host by the tests-only
harness tests/hostdbg/hostdbg_dt.cpp, against the Jet-autodiff oracle (oracle/dt_problem.h) — element-wise, 1e-10 relative (north_star).
Runs on the CPU so that formula regressions are caught without a GPU."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

from conftest import elemerr, ROOT, PKG

P = lambda a: a.ctypes.data_as(C.c_void_p)


@pytest.fixture(scope="module")
def hdt():
    d = os.path.join(ROOT, "tests", "hostdbg")
    so = os.path.join(d, "libhostdbg_dt.so")
    srcs = [os.path.join(d, "hostdbg_dt.cpp"), os.path.join(ROOT, PKG, "csrc", "ctba_dt_math.h"), os.path.join(ROOT, PKG, "csrc", "ctba_math.h")]
    if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["g++", "-O1", "-std=c++17", "-fPIC", "-shared", "-o", so, srcs[0]])
    L = C.CDLL(so)
    L.hdt_imu_cache_size.restype = C.c_int
    return L


def dt_cfg(d, cauchy_a=3.0, g_max=7.8, a_max=150.0):
    return np.array([d["cam_model"]] + list(d["intr"]) + list(d["dist"]) + list(d["q_bc"]) + list(d["p_bc"]) +
                    [d["width"], d["height"], d["t_offset_cam_imu"], d["scale"], d["t_offset_gs_imu"], d["gyro_noise_density"], d["acc_noise_density"],
                     d["gyro_random_walk"], d["acc_random_walk"], g_max, a_max] + list(d["g_W"]) + [cauchy_a], dtype=np.float64)


def _setup(oracle, pkg, scene, cam="euroc", seed=5):
    sc = scene.make_scene(order=6, duration_s=2.5, n_landmarks=90, track_len=8, imu_hz=200, gp_hz=10, cam_hz=10.0, seed=seed, cam=cam,
                          flags=(True, True, False, True, True), outlier_frac=0.05)
    d = scene.dt_scene_from(sc, perturb=(0.002, 0.05))
    o = oracle.OracleDiscrete()
    pkg.load_dt_scene(o, d)
    o.setCorrections(np.array([0.002, 0.001, -0.002, 0.0015, np.sqrt(1 - 0.001**2 - 0.002**2 - 0.0015**2), 0.002, -0.003, 0.001, 0.0004]), -0.0003)
    return o, d


def test_se3_plus(oracle, hdt):
    rng = np.random.default_rng(1)
    for k in range(6):
        T = np.r_[Rot.from_rotvec(rng.normal(size=3)).as_quat(), rng.normal(size=3) * 3]
        d = rng.normal(size=6) * (1e-12 if k == 0 else 0.3)
        out = np.zeros(7); J = np.zeros((7, 6))
        hdt.hdt_se3_plus(P(T), P(d), P(out)); hdt.hdt_se3_plus_jacobian(P(T), P(J))
        np.testing.assert_allclose(out, oracle.se3_plus(T, d), rtol=0, atol=1e-15 * 8)
        np.testing.assert_allclose(J, oracle.se3_plus_jacobian(T), rtol=0, atol=1e-15)


@pytest.mark.parametrize("cam", ["euroc", "fpv"])
def test_reprojection_vs_oracle(oracle, pkg, scene, hdt, cam):
    o, d = _setup(oracle, pkg, scene, cam)
    r_o, J_o = o.evaluate(0)
    Pz, _ = o.getStates(); glb = o.getGlobals(); lm = o.get3DPoints()
    cfg = dt_cfg(d); ob = d["obs_begin"]
    r = np.zeros_like(r_o); J = np.zeros_like(J_o)
    for f in range(len(ob) - 1):
        a, b = int(ob[f]), int(ob[f + 1])
        if a == b:
            continue
        pw = np.ascontiguousarray(lm[d["lm_index"][a:b]]); uv = np.ascontiguousarray(d["uv"][a:b]); vel = np.ascontiguousarray(d["vel"][a:b])
        rr = np.zeros(2 * (b - a)); JJ = np.zeros((2 * (b - a), 17))
        hdt.hdt_reproj(P(cfg), P(np.ascontiguousarray(Pz[d["state_ids"][f]])), P(glb), b - a, P(pw), P(uv), P(vel), C.c_double(1.0 / d["sigma_reproj_px"]), P(rr), P(JJ))
        r[2 * a:2 * b] = rr; J[2 * a:2 * b] = JJ
    assert (r_o == 0).any() and (r_o != 0).sum() > 100           # gated and live reprojections both present
    assert ((r == 0) == (r_o == 0)).all()
    assert elemerr(r, r_o) <= 1.0
    assert elemerr(J, J_o) <= 1.0


def test_global_position_vs_oracle(oracle, pkg, scene, hdt):
    o, d = _setup(oracle, pkg, scene)
    r_o, J_o = o.evaluate(2)
    Pz, _ = o.getStates(); glb = o.getGlobals(); cfg = dt_cfg(d)
    ts_imu = d["cam_ts"] + d["t_offset_cam_imu"]
    r = np.zeros_like(r_o); J = np.zeros_like(J_o)
    for i in range(d["gp_t"].size):
        f0 = int(d["gp_frame0"][i]); rr = np.zeros(3); JJ = np.zeros((3, 17))
        hdt.hdt_gp(P(cfg), P(np.ascontiguousarray(Pz[f0])), P(np.ascontiguousarray(Pz[f0 + 1])), P(glb), C.c_double(d["gp_t"][i]), P(np.ascontiguousarray(d["gp_pos"][i])),
                   C.c_double(ts_imu[f0]), C.c_double(ts_imu[f0 + 1]), C.c_double(1.0 / d["gp_std"]), P(rr), P(JJ))
        r[3 * i:3 * i + 3] = rr; J[3 * i:3 * i + 3] = JJ
    assert r.size >= 30
    assert elemerr(r, r_o) <= 1.0
    assert elemerr(J, J_o) <= 1.0


def test_imu_factor_vs_oracle_with_reintegration(oracle, pkg, scene, hdt):
    """Same sequence of Evaluate calls on both sides: first call (integrates), a small bias move (cache reused, first-order correction), a
    large gyro-bias move (re-integration), then back to the start (re-integration again: the linearisation point moved)."""
    o, d = _setup(oracle, pkg, scene)
    cfg = dt_cfg(d)
    ts_imu = d["cam_ts"] + d["t_offset_cam_imu"]
    nf = len(d["imu"])
    caches = [np.zeros(hdt.hdt_imu_cache_size() // 8 + 1) for _ in range(nf)]
    for c in caches:
        hdt.hdt_imu_cache_init(P(c))
    P0, sb0 = o.getStates()
    rng = np.random.default_rng(3)
    sb0 = sb0 + rng.normal(size=sb0.shape) * np.r_[np.full(3, 0.01), np.full(3, 1e-3), np.full(3, 1e-2)]
    worst_r, worst_J = 0.0, 0.0
    redo_expected = None
    for step, dbg in enumerate((0.0, 1e-5, 5e-3, 0.0)):
        sb = sb0.copy(); sb[:, 3:6] += dbg
        o.initializeStatesKeep(P0, sb)
        r_o, J_o = o.evaluate(1)
        for k, (f0, t, gyr, acc) in enumerate(d["imu"]):
            r = np.zeros(15); J = np.zeros((15, 30))
            rc = hdt.hdt_imu(P(cfg), P(caches[k]), len(t), P(np.ascontiguousarray(t)), P(np.ascontiguousarray(gyr)), P(np.ascontiguousarray(acc)), C.c_double(ts_imu[f0]),
                             C.c_double(ts_imu[f0 + 1]), P(np.ascontiguousarray(P0[f0])), P(np.ascontiguousarray(sb[f0])), P(np.ascontiguousarray(P0[f0 + 1])),
                             P(np.ascontiguousarray(sb[f0 + 1])), P(r), P(J))
            assert rc >= 0
            assert hdt.hdt_imu_redo_count(P(caches[k])) == o.imuRedoCount(k) == (1, 1, 2, 3)[step]
            worst_r = max(worst_r, elemerr(r, r_o[15 * k:15 * k + 15])); worst_J = max(worst_J, elemerr(J, J_o[15 * k:15 * k + 15]))
    assert nf >= 10
    assert worst_r <= 1.0, worst_r
    assert worst_J <= 1.0, worst_J
