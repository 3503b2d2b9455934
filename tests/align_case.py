"""Synthetic align_spline_to_global_frame problem (src/align_spline_to_global_frame.cpp:139-226): a spline expressed in a gauge frame g
(COLMAP's, arbitrary scale) and global position measurements of the antenna in the world frame w,
    z = s R_wg (R_gb(t) p_bp + p_gb(t)) + p_wg + noise,
with a perturbed initial T_wg / scale, a few gross outliers (the reason for CauchyLoss(0.10), bspline_so3r3.h:1189) and a time offset."""
import numpy as np
from scipy.spatial.transform import Rotation as Rot


def make_alignment_case(scene, oracle, N, seed=7, duration_s=20.0, gp_hz=10.0, noise=0.02, outlier_frac=0.05, start_t_ns=5_000_000_000):
    rng = np.random.default_rng(seed)
    dt_ns = 100_000_000
    dt_s = dt_ns * 1e-9
    K = int(duration_s / dt_s) + N + 1
    traj = scene.Trajectory(rng)
    tk = (np.arange(K) - (N - 2) / 2.0) * dt_s
    q = traj.rot(tk).as_quat()
    p = traj.pos(tk)
    s_true, t_off_true = 2.3, 0.012
    R_wg = Rot.from_rotvec([0.3, -0.8, 1.9]); p_wg = np.array([25.0, -13.0, 4.0]); p_bp = np.array([0.08, -0.03, 0.15])
    # ground-truth antenna track, sampled through the oracle's own getPose
    o = oracle.OracleSpline(N, dt_ns, start_t_ns)
    o.setKnots(q, p)
    n = int((duration_s - 1.0) * gp_hz)
    t_meas_ns = start_t_ns + (0.3e9 + np.arange(n) * (1e9 / gp_hz)).astype(np.int64)
    poses = o.sampleAt(0, t_meas_ns - int(t_off_true * 1e9))      # tx ty tz qx qy qz qw at the spline time of each measurement
    ant_g = Rot.from_quat(poses[:, 3:7]).apply(p_bp) + poses[:, 0:3]
    z = s_true * R_wg.apply(ant_g) + p_wg + rng.normal(size=(n, 3)) * noise
    out = rng.random(n) < outlier_frac
    z[out] += rng.normal(size=(int(out.sum()), 3)) * 3.0
    T_wg_init = np.r_[(R_wg * Rot.from_rotvec(rng.normal(size=3) * 0.05)).as_quat(), p_wg + rng.normal(size=3) * 0.5]
    return dict(N=N, dt_ns=dt_ns, start_t_ns=start_t_ns, q=q, p=p, t_ns=t_meas_ns, z=z, T_wg_init=T_wg_init, scale_init=s_true * 1.04,
                t_off_init=0.0, truth=dict(scale=s_true, R_wg=R_wg.as_quat(), p_wg=p_wg, p_bp=p_bp, t_off=t_off_true), noise_arg=1.0)


def load_alignment(prob, case, optimize_pbp, optimize_t):
    """The call sequence of align_spline_to_global_frame.cpp:153-203 on a BSplineSO3R3-shaped object (oracle or device mirror)."""
    prob.setKnots(case["q"], case["p"])
    prob.setWorldColmapTransformation(case["T_wg_init"])
    prob.setGlobals(case["scale_init"], np.zeros(3), np.array([0.0, 0.0, -9.81]), case["t_off_init"])
    prob.setFlags(False, optimize_pbp, False, False, optimize_t)
    return prob.addWorldColmapAlignmentMeasurements(case["t_ns"], case["z"], case["noise_arg"])
