"""drivers/: the file formats either side of the solve and the align_spline_to_global_frame tool re-hosted on the C ABI.
CPU: loaders / writers / flat YAML of drivers/ctba_io.hpp against files written here (including the reference's "one record too many, then
pop_back" idiom and the re-orthonormalisation of pose files); the tool builds with -Wall -Werror and stops loudly without a GPU.
GPU: the tool end to end on a synthetic dataset, result files against the oracle's solve of the same problem."""
import os
import subprocess

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

from align_case import load_alignment, make_alignment_case

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "rpg_vision-based_slam_b200", "csrc")


def _gxx(src, exe, pkg):
    pkg.lib()
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), src, "-L", LIBDIR, "-lctba",
                           "-Wl,-rpath," + LIBDIR, "-o", exe])
    return exe


def _write_pose(fn, q_xyzw, p, digits=9):
    R = Rot.from_quat(q_xyzw).as_matrix()
    with open(fn, "w") as f:
        for r in range(3):
            f.write(" ".join(f"{v:.{digits}g}" for v in list(R[r]) + [p[r]]) + "\n")
        f.write("0 0 0 1\n")


def _write_dataset(d, case, order, optimize_pbp, optimize_t):
    dt_ms = int(case["dt_ns"] * 1e-9 * 1000)
    sub = f"order_{order}/dt_{dt_ms}_ms"
    os.makedirs(os.path.join(d, "fitted", sub)); os.makedirs(os.path.join(d, "align", sub))
    with open(os.path.join(d, "fitted", sub, "spline_knots.txt"), "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for i, (q, p) in enumerate(zip(case["q"], case["p"])):
            t = (case["start_t_ns"] + i * case["dt_ns"]) * 1e-9
            f.write(" ".join(f"{v:.17g}" for v in [t, *p, *q]) + "\n")
    with open(os.path.join(d, "gp.txt"), "w") as f:
        f.write("# timestamp x y z\n")
        for t, z in zip(case["t_ns"], case["z"]):
            f.write(" ".join(f"{v:.17g}" for v in [t * 1e-9, *z]) + "\n")
    _write_pose(os.path.join(d, "align", sub, "T_wg_init.txt"), case["T_wg_init"][:4], case["T_wg_init"][4:], digits=17)
    with open(os.path.join(d, "align", sub, "scale_init.txt"), "w") as f:
        f.write(f"{case['scale_init']:.17g}\n")
    yml = os.path.join(d, "cfg.yaml")
    with open(yml, "w") as f:
        f.write(f"""dataset: Custom

## Files
gp_fn: "{d}/gp.txt"   # measurements
colmap_spline_dir: "{d}/fitted"
spline_global_ref_alignment_dir: '{d}/align'

## Parameters
spline_order: {order}
spline_control_nodes_dt_s: {case['dt_ns'] * 1e-9} # [s]
init_t_offset_cam_gp: {case['t_off_init']}
optimize_pBP: {'true' if optimize_pbp else 'false'} # pBP = body - antenna
optimize_spline_globsens_time_offset: {'True' if optimize_t else 'False'}
n_max_iters: 30
n_threads: 4
""")
    return yml, os.path.join(d, "align", sub)


def _colmap_track(rng, duration_s, hz, t0=1000.0):
    """COLMAP body estimates: t tx ty tz qx qy qz qw at the camera rate, a smooth track with a little noise"""
    t = t0 + np.arange(int(duration_s * hz) + 1) / hz
    p = np.c_[2.0 * np.cos(0.5 * t), 2.0 * np.sin(0.5 * t), 0.3 * np.sin(0.9 * t)] + rng.normal(size=(len(t), 3)) * 0.002
    q = (Rot.from_rotvec(np.c_[0.2 * np.sin(0.7 * t), 0.1 * np.cos(0.4 * t), 0.5 * t]) * Rot.from_rotvec(rng.normal(size=(len(t), 3)) * 0.001)).as_quat()
    return t, q, p


def _write_traj(fn, t, q, p):
    with open(fn, "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for i in range(len(t)):
            f.write(" ".join(f"{v:.17g}" for v in [t[i], *p[i], *q[i]]) + "\n")


def _interp(t, q, p, its):
    """linearInterpolation (util/trajectory_utils.h:96-143) + Sophus::interpolate (interpolate.hpp:27-34), restated with scipy"""
    op, oq = [], []
    for x in its:
        i = int(np.searchsorted(t, x, side="right")) - 1
        a = (x - t[i]) / (t[i + 1] - t[i])
        op.append(p[i] + a * (p[i + 1] - p[i]))
        r1, r2 = Rot.from_quat(q[i]), Rot.from_quat(q[i + 1])
        oq.append((r1 * Rot.from_rotvec(a * (r1.inv() * r2).as_rotvec())).as_quat())
    return np.array(op), np.array(oq)


def test_io_formats_roundtrip(pkg, tmp_path):
    d = str(tmp_path)
    rng = np.random.default_rng(5)
    gp = np.c_[100.0 + 0.1 * np.arange(7), rng.normal(size=(7, 3))]
    with open(os.path.join(d, "gp.txt"), "w") as f:
        f.write("# t x y z\n")
        for r in gp: f.write(" ".join(f"{v:.17g}" for v in r) + "\n")
    q = Rot.from_rotvec(rng.normal(size=(5, 3))).as_quat(); p = rng.normal(size=(5, 3))
    with open(os.path.join(d, "spline_knots.txt"), "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for i in range(5): f.write(" ".join(f"{v:.17g}" for v in [50.0 + 0.1 * i, *p[i], *(2.0 * q[i])]) + "\n")     # un-normalised on purpose
    qT = Rot.from_rotvec([2.9, 0.4, -0.6]).as_quat()       # trace < 0: the other branch of the matrix -> quaternion conversion
    _write_pose(os.path.join(d, "T.txt"), qT, [1.5, -2.5, 3.5], digits=6)   # 6 digits: visibly non-orthogonal before the fix
    open(os.path.join(d, "scalar.txt"), "w").write("# scale\n2.25\n")
    open(os.path.join(d, "vec.txt"), "w").write("0.5 -0.25 0.125\n")
    open(os.path.join(d, "cfg.yaml"), "w").write('gp_fn: "/a b/c#d.txt"  # comment\nspline_order: 6\nspline_control_nodes_dt_s: 0.1 # [s]\n'
                                                 "optimize_pBP: false\noptimize_spline_globsens_time_offset: True\n")
    traj_t, traj_q, traj_p = _colmap_track(rng, 2.0, 10.0)
    _write_traj(os.path.join(d, "traj.txt"), traj_t, traj_q, traj_p)
    its = np.array([traj_t[0], traj_t[0] + 0.03, traj_t[3], 0.5 * (traj_t[7] + traj_t[8]), traj_t[-1] - 1e-9])
    np.savetxt(os.path.join(d, "interp_ts.txt"), its, fmt="%.17g")
    # calibration files in the two layouts of the reference's datasets (datasets/UZH_FPV/calib/*/camchain-*.yaml + imu.yaml, datasets/EuRoC/calib/*/calib.yaml)
    q_ci = [Rot.from_rotvec(rng.normal(size=3)) for _ in range(2)]; p_ci = rng.normal(size=(2, 3)) * 0.05
    def _rows(R, t): return [list(R.as_matrix()[r]) + [t[r]] for r in range(3)] + [[0.0, 0.0, 0.0, 1.0]]
    def _flow(rows, sep): return sep.join(", ".join(f"{v:.17g}" for v in r) for r in rows)
    with open(os.path.join(d, "camchain.yaml"), "w") as f:
        f.write("cam0:\n  T_cam_imu: [%s]\n  camera_model: pinhole\n  distortion_coeffs: [-0.0137, 0.0207, -0.0127,\n    0.0025]\n  distortion_model: equidistant\n"
                "  intrinsics: [278.66, 278.48, 319.75, 241.96]\n  resolution: [640, 480]\n  timeshift_cam_imu: -0.016684572091862235\n" % _flow(_rows(q_ci[0], p_ci[0]), ",\n  "))
        f.write("cam1:\n  T_cam_imu:\n  [%s]\n  camera_model: pinhole\n  distortion_coeffs: [-0.0084, 0.0114, -0.0069,\n    0.0015]\n  distortion_model: equidistant\n"
                "  intrinsics: [277.61, 277.63, 314.89, 236.04]\n  resolution: [640, 480]\n  timeshift_cam_imu: -0.0165\n" % _flow(_rows(q_ci[1], p_ci[1]), ",\n  "))
    open(os.path.join(d, "imu.yaml"), "w").write("accelerometer_noise_density: 0.1\naccelerometer_random_walk: 0.002\ngyroscope_noise_density: 0.05\n"
                                                 "gyroscope_random_walk: 4.0e-05\nrostopic: /snappy_imu\nupdate_rate: 500.0\n")
    R_bc1 = Rot.from_rotvec([1.2, -1.2, 1.2]); p_bc1 = np.array([-0.0198, 0.0454, 0.0079])
    with open(os.path.join(d, "euroc.yaml"), "w") as f:
        f.write('label: "Euroc"\ncameras:\n')
        for k, (R, t, fx) in enumerate(((Rot.from_rotvec([0.1, 0.2, 0.3]), np.zeros(3), 458.65), (R_bc1, p_bc1, 457.58))):
            f.write(f"- camera:\n    label: cam{k}\n    image_height: 480\n    image_width: 752\n    type: pinhole\n    intrinsics:\n      cols: 1\n      rows: 4\n"
                    f"      data: [{fx}, 456.13, 379.99, 255.23]\n    distortion:\n      type: radial-tangential\n      parameters:\n        cols: 1\n        rows: 4\n"
                    f"        data: [-0.2836, 0.0745, -0.000104,\n    -3.55e-05]\n  T_B_C:\n    cols: 4\n    rows: 4\n    data: [{_flow(_rows(R, t), ',' + chr(10) + '           ')}]\n")
        f.write("\nimu_params:\n  delay_imu_cam: 0.004\n  sigma_omega_c: 0.05\n  sigma_acc_c: 0.1\n  sigma_omega_bias_c: 4.0e-05\n  sigma_acc_bias_c: 0.002\n  imu_rate: 200\n")
    exe = _gxx(os.path.join(ROOT, "tests", "cpp", "io_roundtrip.cpp"), str(tmp_path / "io_roundtrip"), pkg)
    out = subprocess.run([exe, d], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK"), out.stdout + out.stderr
    kv = {l.split(" ", 1)[0]: l.split(" ", 1)[1] for l in out.stdout.splitlines() if " " in l and not l.startswith("Created")}
    assert int(kv["gp_count"]) == 7 and int(kv["knot_count"]) == 5
    np.testing.assert_array_equal(np.array(kv["gp_last"].split(), dtype=float), gp[-1])
    k1 = np.array(kv["knot1"].split(), dtype=float)
    np.testing.assert_array_equal(k1[:4], [50.1, *p[1]])
    np.testing.assert_allclose(k1[4:], q[1], atol=1e-15)                       # normalised like Sophus::SO3d(Quaterniond)
    pose = np.array(kv["pose"].split(), dtype=float)
    assert abs(np.linalg.norm(pose[:4]) - 1.0) < 1e-15
    assert (Rot.from_quat(pose[:4]).inv() * Rot.from_quat(qT)).magnitude() < 1e-5
    # the re-orthonormalised rotation is the polar factor U V^T of what was in the file (fixRotationMatrix, utils.h:34-50)
    M = np.loadtxt(os.path.join(d, "T.txt"))[:3, :3]
    U, _, Vt = np.linalg.svd(M)
    np.testing.assert_allclose(Rot.from_quat(pose[:4]).as_matrix(), U @ Vt, atol=1e-14)
    np.testing.assert_array_equal(pose[4:], [1.5, -2.5, 3.5])
    assert float(kv["scalar"]) == 2.25 and kv["vec"].split() == ["0.5", "-0.25", "0.125"]
    assert kv["yaml"] == "/a b/c#d.txt|6|0.10000000000000001|0|1|0"
    ip, iq = _interp(traj_t, traj_q, traj_p, its)
    for i in range(len(its)):
        v = np.array(kv[f"interp{i}"].split(), dtype=float)
        np.testing.assert_allclose(v[:3], ip[i], atol=1e-14)
        assert min(np.abs(v[3:] - iq[i]).max(), np.abs(v[3:] + iq[i]).max()) < 1e-13
    for cam in range(2):
        head, *vals = kv[f"kalibr{cam}"].split(" ")
        v = np.array(vals, dtype=float)
        assert head == "pinhole|equidistant"
        np.testing.assert_array_equal(v[:4], [[278.66, 278.48, 319.75, 241.96], [277.61, 277.63, 314.89, 236.04]][cam])
        np.testing.assert_array_equal(v[4:8], [[-0.0137, 0.0207, -0.0127, 0.0025], [-0.0084, 0.0114, -0.0069, 0.0015]][cam])
        np.testing.assert_array_equal(v[8:16], [640, 480, [-0.016684572091862235, -0.0165][cam], 500.0, 0.1, 0.002, 0.05, 4.0e-05])
        T_bc = q_ci[cam].inv()                                   # the drivers use T_b_c = T_cam_imu^-1
        assert (Rot.from_quat(v[16:20]).inv() * T_bc).magnitude() < 1e-14
        np.testing.assert_allclose(v[20:23], -T_bc.apply(p_ci[cam]), atol=1e-15)
    head, *vals = kv["euroc1"].split(" ")
    v = np.array(vals, dtype=float)
    assert head == "pinhole|radial-tangential"
    np.testing.assert_array_equal(v[:6], [457.58, 255.23, -0.2836, -3.55e-05, -0.004, 200.0])      # timeshift = -delay_imu_cam
    assert (Rot.from_quat(v[6:10]).inv() * R_bc1).magnitude() < 1e-14
    np.testing.assert_allclose(v[10:13], p_bc1, atol=1e-15)
    # writers: 3x4 + "0 0 0 1" with 9 significant digits, "# x y z" header, 16-digit scalars (save.h:156-178)
    T_out = np.loadtxt(os.path.join(d, "T_out.txt"))
    assert T_out.shape == (4, 4) and list(T_out[3]) == [0, 0, 0, 1]
    np.testing.assert_allclose(T_out[:3, :3], Rot.from_quat(pose[:4]).as_matrix(), atol=5e-9)
    assert open(os.path.join(d, "p_out.txt")).read() == "# x y z\n0.5 -0.25 0.125\n"
    assert open(os.path.join(d, "s_out.txt")).read() == "2.25\n"


def test_align_tool_builds_and_fails_loudly_without_gpu(pkg, oracle, scene, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "align_spline_to_global_frame.cpp"), str(tmp_path / "align_tool"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    case = make_alignment_case(scene, oracle, 6, seed=11, duration_s=5.0)
    yml, _ = _write_dataset(str(tmp_path), case, 6, True, True)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    assert "Loaded %d leica measurements" % len(case["t_ns"]) in out.stdout and "Loaded %d spline knots" % len(case["q"]) in out.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("order,optimize_pbp,optimize_t", [(6, True, True), (4, False, False)])
def test_align_tool_end_to_end_on_gpu(pkg, oracle, scene, tmp_path, order, optimize_pbp, optimize_t):
    exe = _gxx(os.path.join(ROOT, "drivers", "align_spline_to_global_frame.cpp"), str(tmp_path / "align_tool"), pkg)
    case = make_alignment_case(scene, oracle, order, seed=21 + order, duration_s=30.0)
    yml, res = _write_dataset(str(tmp_path), case, order, optimize_pbp, optimize_t)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    # the oracle on the same problem, loaded the way the tool loads it (pose file with 17 digits: the same T_wg_init up to rounding)
    o = oracle.OracleSpline(order, case["dt_ns"], case["start_t_ns"])
    load_alignment(o, case, optimize_pbp, optimize_t)
    so = o.optimize(30)
    assert so["termination_type"] == 0
    T_o, _ = o.getWorldColmapTransformation()
    o.relinearize()
    g = o.getGlobals()
    T_wg = np.loadtxt(os.path.join(res, "T_wg.txt"))
    np.testing.assert_allclose(T_wg[:3, :3], Rot.from_quat(T_o[:4]).as_matrix(), atol=2e-8)      # 9 significant digits in the file
    np.testing.assert_allclose(T_wg[:3, 3], T_o[4:], rtol=2e-8, atol=1e-7)
    assert abs(float(open(os.path.join(res, "scale.txt")).read()) - (g[18] + g[0])) < 1e-8
    assert abs(float(open(os.path.join(res, "spline_globalsensor_time_offset.txt")).read()) - g[17]) < 1e-9
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "p_bp.txt")), g[10:13], atol=1e-7)
    # spline.txt: the frozen spline sampled at 100 Hz like traceSpline (save.h:99-117)
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = (len(case["q"]) - order + 1) * case["dt_ns"] + case["start_t_ns"] - 1_000_000
    t_ns = np.arange(case["start_t_ns"], end_ns, 10_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:], ref, atol=1e-12)
    assert f"Number of errors: {len(case['t_ns'])}" in out.stdout


def _write_fit_dataset(d, rng, order, duration_s=8.0, num_errors=150):
    t, q, p = _colmap_track(rng, duration_s, 20.0)
    _write_traj(os.path.join(d, "colmap_body_estimates.txt"), t, q, p)
    yml = os.path.join(d, "fit.yaml")
    with open(yml, "w") as f:
        f.write(f"""## Files
colmap_fn: "{d}/colmap_body_estimates.txt"
colmap_spline_dir: "{d}/results"
## Parameters
start_t: {t[2] - 1e-6:.12f} # time of first image
end_t: {t[-3] + 1e-6:.12f}
spline_control_nodes_dt_s: 0.1
num_errors: {num_errors}
n_max_iters: 20
n_threads: 4
""")
    return yml, os.path.join(d, "results", "colmap_fitted_spline", f"order_{order}", "dt_100_ms"), (t[2:-2], q[2:-2], p[2:-2])


def _fit_pipeline_python(oracle, order, t, q, p, num_errors):
    """src/fit_spline_to_colmap.cpp:126-190 restated: knots and measurements by linear interpolation, then the oracle's priors"""
    dt_s = 0.1
    knots_ts = []
    x = t[0]
    while x <= t[-1]:
        knots_ts.append(x); x += dt_s
    kp, kq = _interp(t, q, p, knots_ts)
    dt_meas = (t[-1] - t[0]) / float(num_errors)
    meas_ts = []
    x = t[0]
    while x <= t[-1]:
        meas_ts.append(x); x += dt_meas
    mp, mq = _interp(t, q, p, meas_ts)
    o = oracle.OracleSpline(order, int(dt_s * 1e9), int(t[0] * 1e9))
    o.setKnots(kq, kp)
    t_ns = np.array([int(v * 1e9) for v in meas_ts], dtype=np.int64)
    n_pos = o.addPositionMeasurements(t_ns, mp, 1.0)
    n_ori = o.addOrientationMeasurements(t_ns, mq, 1.0)
    return o, np.array(knots_ts), n_pos, n_ori


def test_fit_tool_builds_and_fails_loudly_without_gpu(pkg, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "fit_spline_to_colmap.cpp"), str(tmp_path / "fit_tool"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    yml, _, _ = _write_fit_dataset(str(tmp_path), np.random.default_rng(2), 6)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    assert "Loaded 161 colmap estimates" in out.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("order", [6, 4])
def test_fit_tool_end_to_end_on_gpu(pkg, oracle, tmp_path, order):
    exe = _gxx(os.path.join(ROOT, "drivers", "fit_spline_to_colmap.cpp"), str(tmp_path / "fit_tool"), pkg)
    num_errors = 150
    yml, res, (t, q, p) = _write_fit_dataset(str(tmp_path), np.random.default_rng(30 + order), order, num_errors=num_errors)
    out = subprocess.run([exe, yml, "0", str(order)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    o, knots_ts, n_pos, n_ori = _fit_pipeline_python(oracle, order, t, q, p, num_errors)
    assert f"Number of position errors: {n_pos}\n" in out.stdout and f"Number of orientation errors: {n_ori}\n" in out.stdout
    so = o.optimize(20)
    kq_o, kp_o = o.getKnots()
    kn = np.loadtxt(os.path.join(res, "spline_knots.txt"))
    assert kn.shape == (len(knots_ts), 8)
    np.testing.assert_allclose(kn[:, 0], knots_ts, rtol=1e-14)
    # weakly constrained end knots amplify the 1e-8 differences of the LM traces (measured: < 1e-6); the sampled trajectory (what the next
    # tools consume) is checked far inside the 1 mm / 0.01 deg of north_star
    np.testing.assert_allclose(kn[:, 1:4], kp_o, atol=1e-5)
    sgn = np.sign(np.sum(kn[:, 4:] * kq_o, axis=1))[:, None]
    np.testing.assert_allclose(kn[:, 4:], sgn * kq_o, atol=1e-5)
    line = [l for l in out.stdout.splitlines() if l.startswith("iterations ")][0]
    final_cost = float(line.split("->")[1].split(",")[0])
    assert abs(final_cost - so["final_cost"]) <= 1e-6 * so["final_cost"]
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = (len(knots_ts) - order + 1) * 100_000_000 + int(t[0] * 1e9) - 1_000_000
    t_ns = np.arange(int(t[0] * 1e9), end_ns, 10_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:4], ref[:, :3], atol=1e-5)
    ang = (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(ref[:, 3:])).magnitude()
    assert ang.max() < np.deg2rad(0.001)


def _write_gpfit_dataset(d, rng, order, num_errors=120):
    t = 2000.0 + np.arange(160) / 20.0                        # last stamp t0 + 7.95: strictly after the last 0.1 s knot (see the tool's comment)
    z = np.c_[3.0 * np.cos(0.4 * t), 2.0 * np.sin(0.6 * t), 0.5 * np.sin(0.9 * t)] + rng.normal(size=(len(t), 3)) * 0.01
    with open(os.path.join(d, "gp.txt"), "w") as f:
        f.write("# timestamp x y z\n")
        for i in range(len(t)):
            f.write(" ".join(f"{v:.17g}" for v in [t[i], *z[i]]) + "\n")
    yml = os.path.join(d, "gpfit.yaml")
    with open(yml, "w") as f:
        f.write(f'gp_fn: "{d}/gp.txt"\nspline_dir: "{d}/results"\nstart_t: {t[0] - 1.0:.6f}\nend_t: {t[-1] + 1.0:.6f}\n'
                f"spline_control_nodes_dt_s: 0.1\nnum_errors: {num_errors}\nn_max_iters: 20\nn_threads: 4\n")
    return yml, os.path.join(d, "results", "fit_spline_on_gp_meas", f"order_{order}", "dt_100_ms"), (t, z)


@pytest.mark.gpu
@pytest.mark.parametrize("order", [6, 4])
def test_gpfit_tool_end_to_end_on_gpu(pkg, oracle, tmp_path, order):
    """fit_spline_to_gp_measurements: position priors only; the rotation knots are not part of the problem and stay at identity."""
    exe = _gxx(os.path.join(ROOT, "drivers", "fit_spline_to_gp_measurements.cpp"), str(tmp_path / "gpfit_tool"), pkg)
    num_errors = 120
    yml, res, (t, z) = _write_gpfit_dataset(str(tmp_path), np.random.default_rng(70 + order), order, num_errors)
    out = subprocess.run([exe, yml, "0", str(order)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    ident = np.tile([0.0, 0.0, 0.0, 1.0], (len(t), 1))
    knots_ts, x = [], t[0]
    while x <= t[-1]:
        knots_ts.append(x); x += 0.1
    kp, kq = _interp(t, ident, z, knots_ts)
    dt_meas = (t[-1] - t[0]) / float(num_errors)
    meas_ts, x = [], t[0]
    while x <= t[-1]:
        meas_ts.append(x); x += dt_meas
    mp, _ = _interp(t, ident, z, meas_ts)
    o = oracle.OracleSpline(order, 100_000_000, int(t[0] * 1e9))
    o.setKnots(kq, kp)
    n_pos = o.addPositionMeasurements(np.array([int(v * 1e9) for v in meas_ts], dtype=np.int64), mp, 1.0)
    assert f"Number of position errors: {n_pos}\n" in out.stdout
    so = o.optimize(20)
    kq_o, kp_o = o.getKnots()
    kn = np.loadtxt(os.path.join(res, "spline_knots.txt"))
    assert kn.shape == (len(knots_ts), 8)
    np.testing.assert_allclose(kn[:, 1:4], kp_o, atol=1e-5)
    np.testing.assert_array_equal(kn[:, 4:], ident[:len(kn)])
    line = [l for l in out.stdout.splitlines() if l.startswith("iterations ")][0]
    assert int(line.split()[1].rstrip(",")) == len(so["iterations"])
    final_cost = float(line.split("->")[1].split(",")[0])
    assert abs(final_cost - so["final_cost"]) <= 1e-6 * so["final_cost"]


def _write_colmap_bin(d, cams, images, points):
    """COLMAP's binary export (reconstruction.cc WriteCamerasBinary / WriteImagesBinary / WritePoints3DBinary)"""
    import struct
    with open(os.path.join(d, "cameras.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(cams)))
        for c in cams:
            f.write(struct.pack("<IiQQ", c["id"], c["model"], c["w"], c["h"]))
            f.write(struct.pack(f"<{len(c['params'])}d", *c["params"]))
    with open(os.path.join(d, "images.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(images)))
        for im in images:
            f.write(struct.pack("<I4d3dI", im["id"], *im["qwxyz"], *im["t"], im["cam"]))
            f.write(im["name"].encode() + b"\0")
            f.write(struct.pack("<Q", len(im["pts"])))
            for (x, y, pid) in im["pts"]:
                f.write(struct.pack("<ddQ", x, y, pid))
    with open(os.path.join(d, "points3D.bin"), "wb") as f:
        f.write(struct.pack("<Q", len(points)))
        for pt in points:
            f.write(struct.pack("<Q3d3Bd", pt["id"], *pt["xyz"], 10, 20, 30, pt["err"]))
            f.write(struct.pack("<Q", len(pt["track"])))
            for (im_id, p2) in pt["track"]:
                f.write(struct.pack("<II", im_id, p2))


def _kept_points(pts):
    """the selection rules of readImagesAndPoints3DBinary (sensor/camera.h:340-385), restated"""
    INVALID = 2 ** 64 - 1
    xs, ys, seen, kept = {}, {}, [], []
    for j, (x, y, pid) in enumerate(pts):
        if pid == INVALID: continue
        if x in xs and y in ys and xs[x] == ys[y]: continue
        if pid in seen: continue
        kept.append((j, x, y, pid)); seen.append(pid)
        xs[x] = j; ys[y] = j
    return kept


def test_colmap_binary_ingest(tmp_path):
    d = str(tmp_path)
    rng = np.random.default_rng(9)
    INVALID = 2 ** 64 - 1
    # OPENCV_FISHEYE (8 parameters) is read exactly; OPENCV (model 4, 8 parameters) is read as fx fy cx cy + 8 doubles, i.e. four past its
    # record (the reference's quirk) — harmless only for the last camera of the file, which is where it is put here
    cams = [dict(id=1, model=5, w=640, h=480, params=[278.6, 278.5, 319.5, 241.2, -0.013, 0.02, -0.01, 0.003]),
            dict(id=2, model=4, w=752, h=480, params=[458.6, 457.3, 367.2, 248.4, -0.28, 0.07, 1.9e-4, 1.8e-5])]
    points = [dict(id=int(i), xyz=rng.normal(size=3) * 4, err=float(rng.random()), track=[(1, 0), (2, 3)]) for i in (7, 3, 900000000000, 12, 5)]
    points.append(dict(id=3, xyz=[9.0, 9.0, 9.0], err=1.0, track=[(3, 1)]))          # repeated id: the first record wins
    images = []
    for k in range(3):
        pts = [(float(x), float(y), int(pid)) for x, y, pid in zip(rng.uniform(0, 700, 9), rng.uniform(0, 470, 9), [7, INVALID, 3, 12, INVALID, 5, 900000000000, 44, 7])]
        pts.insert(4, (pts[0][0], pts[0][1], 3 if k else 99))    # exact duplicate of point 0 (note 1), whatever its 3-D id
        pts.insert(6, (pts[2][0], 123.25, 12))                   # same x as an earlier point, different y: kept unless its 3-D id was seen
        q = Rot.from_rotvec(rng.normal(size=3)).as_quat() * 1.7  # un-normalised on disk
        images.append(dict(id=k + 1, qwxyz=[q[3], q[0], q[1], q[2]], t=list(rng.normal(size=3)), cam=1 + k % 2, name=f"cam0/{1000 + k}.png", pts=pts))
    _write_colmap_bin(d, cams, images, points)
    exe = str(tmp_path / "colmap_ingest")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-Wall", "-Werror", os.path.join(ROOT, "tests", "cpp", "colmap_ingest.cpp"), "-o", exe])
    out = subprocess.run([exe, d], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0 and out.stdout.strip().endswith("OK"), out.stdout + out.stderr
    L = [l.split() for l in out.stdout.splitlines()]
    cam_l = [l for l in L if l[0] == "cam"]
    assert [int(cam_l[0][1]), int(cam_l[0][2]), int(cam_l[0][9])] == [1, 5, 4] and [int(cam_l[1][1]), int(cam_l[1][2]), int(cam_l[1][9])] == [2, 4, 8]
    for c in range(2):
        assert [int(cam_l[c][3]), int(cam_l[c][4])] == [cams[c]["w"], cams[c]["h"]]
        np.testing.assert_array_equal(np.array(cam_l[c][5:9], dtype=float), cams[c]["params"][:4])
        np.testing.assert_array_equal(np.array(cam_l[c][10:14], dtype=float), cams[c]["params"][4:8])
    fr_l = [l for l in L if l[0] == "frame"]; pt_l = [l for l in L if l[0] == "pt"]; lm_l = [l for l in L if l[0] == "lm"]
    assert len(fr_l) == 3
    ids = [p["id"] for p in points[:5]]
    for k, im in enumerate(images):
        assert int(fr_l[k][1]) == k and fr_l[k][3] == im["name"]
        q = np.array(fr_l[k][4:8], dtype=float)
        qn = np.array(im["qwxyz"])[[1, 2, 3, 0]]; qn = qn / np.linalg.norm(qn)
        np.testing.assert_allclose(q, qn, atol=1e-16)
        np.testing.assert_array_equal(np.array(fr_l[k][8:11], dtype=float), im["t"])
        kept = _kept_points(im["pts"])
        got = [(int(l[2]), float(l[3]), float(l[4]), int(l[5]), int(l[6])) for l in pt_l if int(l[1]) == k]
        assert int(fr_l[k][11]) == len(kept) == len(got)
        for (j, x, y, pid), g in zip(kept, got):
            assert g[:4] == (j, x, y, pid) and g[4] == (ids.index(pid) if pid in ids else -1)
    assert [int(l[2]) for l in lm_l] == ids
    np.testing.assert_array_equal(np.array(lm_l[1][3:6], dtype=float), points[1]["xyz"])     # id 3: the first record
    assert all(int(l[7]) == 2 for l in lm_l)


def _t_of(t_ns):
    """what static_cast<int64_t>(t * 1e9) gives for a stamp written with 17 digits"""
    return int(float(f"{t_ns * 1e-9:.17g}") * 1e9)


def _write_gv_dataset(d, sc, order, T_wg_q, T_wg_p, flags):
    """A scene of scene.make_scene as the files optimize_gv_continuous_time reads."""
    R_wg = Rot.from_quat(T_wg_q)
    dt_ms = int(sc["dt_ns"] * 1e-9 * 1000)
    sub = f"order_{order}/dt_{dt_ms}_ms"
    for k in ("fitted", "align"): os.makedirs(os.path.join(d, k, sub))
    os.makedirs(os.path.join(d, "colmap"))
    kq_g = (R_wg.inv() * Rot.from_quat(sc["knot_q"])).as_quat(); kp_g = R_wg.inv().apply(sc["knot_p"] - T_wg_p) / sc["scale"]
    lm_g = R_wg.inv().apply(sc["landmarks"] - T_wg_p) / sc["scale"]
    with open(os.path.join(d, "fitted", sub, "spline_knots.txt"), "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for i in range(len(kq_g)):
            f.write(" ".join(f"{v:.17g}" for v in [(sc["start_t_ns"] + i * sc["dt_ns"]) * 1e-9, *kp_g[i], *kq_g[i]]) + "\n")
    _write_pose(os.path.join(d, "align", sub, "T_wg_init.txt"), T_wg_q, T_wg_p, digits=17)
    open(os.path.join(d, "align", sub, "scale_init.txt"), "w").write(f"{sc['scale']:.17g}\n")
    open(os.path.join(d, "align", sub, "p_bp.txt"), "w").write("# x y z\n" + " ".join(f"{v:.17g}" for v in sc["p_bp"]) + "\n")
    with open(os.path.join(d, "gp.txt"), "w") as f:
        f.write("# timestamp x y z\n")
        for t, z in zip(sc["gp_t_ns"], sc["gp_pos"]):
            f.write(" ".join(f"{v:.17g}" for v in [t * 1e-9, *z]) + "\n")
    F = len(sc["frame_t_ns"])
    with open(os.path.join(d, "colmap", "colmap_cam_estimates.txt"), "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for t in sc["frame_t_ns"]:
            f.write(f"{t * 1e-9:.17g} 0 0 0 0 0 0 1\n")
    ids = [10 + 3 * i for i in range(len(lm_g))]
    cams = [dict(id=1, model=sc["cam_model"], w=sc["width"], h=sc["height"], params=list(sc["intr"]) + list(sc["dist"]))]
    images = []
    for f_ in range(F):
        b, e = sc["obs_begin"][f_], sc["obs_begin"][f_ + 1]
        pts = [(float(sc["uv"][q, 0]), float(sc["uv"][q, 1]), ids[sc["lm_index"][q]]) for q in range(b, e)]
        pts.insert(0, (5.0, 5.0, 2 ** 64 - 1))                      # a 2-D point without a 3-D point
        images.append(dict(id=f_ + 1, qwxyz=[1.0, 0, 0, 0], t=[0.0, 0, 0], cam=1, name=f"{f_}.png", pts=pts))
    points = [dict(id=ids[i], xyz=list(lm_g[i]), err=0.5, track=[]) for i in range(len(lm_g))]
    _write_colmap_bin(os.path.join(d, "colmap"), cams, images, points)
    R_bc = Rot.from_quat(sc["q_bc"]).as_matrix()
    rows = [list(R_bc[r]) + [sc["p_bc"][r]] for r in range(3)] + [[0.0, 0.0, 0.0, 1.0]]
    with open(os.path.join(d, "calib.yaml"), "w") as f:
        f.write("label: \"synthetic\"\ncameras:\n- camera:\n    label: cam0\n    image_height: %d\n    image_width: %d\n    type: pinhole\n" % (sc["height"], sc["width"]))
        f.write("    intrinsics:\n      cols: 1\n      rows: 4\n      data: [%s]\n" % ", ".join(f"{v:.17g}" for v in sc["intr"]))
        f.write("    distortion:\n      type: radial-tangential\n      parameters:\n        cols: 1\n        rows: 4\n        data: [%s,\n    %s]\n"
                % (", ".join(f"{v:.17g}" for v in sc["dist"][:3]), f"{sc['dist'][3]:.17g}"))
        f.write("  T_B_C:\n    cols: 4\n    rows: 4\n    data: [%s]\n" % ",\n           ".join(", ".join(f"{v:.17g}" for v in r) for r in rows))
        f.write("\nimu_params:\n  delay_imu_cam: %.17g\n  sigma_omega_c: 0.05\n  sigma_acc_c: 0.1\n  sigma_omega_bias_c: 4.0e-05\n  sigma_acc_bias_c: 0.002\n  imu_rate: 200\n"
                % (-sc["t_offset_cam_imu"]))
    yml = os.path.join(d, "gv.yaml")
    with open(yml, "w") as f:
        f.write(f"""dataset: Custom
cam_id: 0
camimu_calib_fn: "{d}/calib.yaml"
gp_fn: "{d}/gp.txt"
colmap_dir: "{d}/colmap/"
colmap_spline_dir: "{d}/fitted"
spline_global_ref_alignment_dir: "{d}/align"
full_batch_optimization_dir: "{d}/full_batch"
spline_control_nodes_dt_s: {sc['dt_ns'] * 1e-9}
sampling_spline_rate: 50.0
grid_cell_size: 30
gp_std: {sc['gp_std']}
sigma_reproj_px: {sc['sigma_reproj_px']}
optimize_scale: {'true' if flags[0] else 'false'}
optimize_pBP: {'true' if flags[1] else 'false'}
optimize_spline_globsens_time_offset: {'true' if flags[4] else 'false'}
n_max_outer_iters: 3
n_max_inner_iters: 4
n_threads: 4
""")
    return yml, os.path.join(d, "full_batch", "continuous_time", sub), (kq_g, kp_g, lm_g)


def _gv_scene(scene, order):
    flags = (True, True, False, True, True)     # the driver leaves optimize_camimu_extrinsics at the constructor default (true)
    return scene.make_scene(order=order, duration_s=3.0, n_landmarks=160, track_len=9, imu_hz=0, gp_hz=10, flags=flags, seed=77 + order), flags


def test_gv_driver_builds_and_fails_loudly_without_gpu(pkg, scene, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_gv_continuous_time.cpp"), str(tmp_path / "gv"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc, flags = _gv_scene(scene, 6)
    yml, _, _ = _write_gv_dataset(str(tmp_path), sc, 6, Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0]), flags)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    # everything up to the construction of the spline ran: calibration, measurements, COLMAP export, knots
    assert f"Loaded {len(sc['gp_t_ns'])} global measurements" in out.stdout and f"Loaded: {len(sc['frame_t_ns'])} images" in out.stdout
    assert f"Loaded: {len(sc['landmarks'])} 3D points" in out.stdout and f"Loaded {len(sc['knot_q'])} initial spline knots" in out.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("order", [6, 4])
def test_gv_driver_end_to_end_on_gpu(pkg, oracle, scene, tmp_path, order):
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_gv_continuous_time.cpp"), str(tmp_path / "gv"), pkg)
    sc, flags = _gv_scene(scene, order)
    T_q, T_p = Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0])
    yml, res, (kq_g, kp_g, lm_g) = _write_gv_dataset(str(tmp_path), sc, order, T_q, T_p, flags)
    out = subprocess.run([exe, yml, "0", str(order)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    # the oracle through the same steps (src/optimize_gv_continuous_time.cpp:320-511)
    R_wg = Rot.from_quat(T_q)
    o = oracle.OracleSpline(order, sc["dt_ns"], sc["start_t_ns"])
    o.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"])
    o.setKnots((R_wg * Rot.from_quat(kq_g)).as_quat(), sc["scale"] * R_wg.apply(kp_g) + T_p)
    o.setGlobals(sc["scale"], sc["p_bp"], np.zeros(3), 0.0)
    o.setFlags(*flags)
    o.set3DPoints(sc["scale"] * R_wg.apply(lm_g) + T_p)
    t_fr = np.array([_t_of(t) for t in sc["frame_t_ns"]], dtype=np.int64)
    valid, _ = o.computeValidReprojections(t_fr, sc["obs_begin"], sc["uv"], sc["lm_index"], sc["width"], sc["height"], 30)
    o.addGp(np.array([_t_of(t) for t in sc["gp_t_ns"]], dtype=np.int64), sc["gp_pos"], sc["gp_std"])
    ob = np.r_[0, np.cumsum([valid[sc["obs_begin"][f]:sc["obs_begin"][f + 1]].sum() for f in range(len(t_fr))])].astype(np.int32)
    o.addFrames(t_fr, ob, sc["uv"][valid], sc["lm_index"][valid], sc["sigma_reproj_px"])
    assert f"valid reprojections: {int(valid.sum())}\n" in out.stdout
    s = o.optimize(1); o.relinearize()
    costs = [s["final_cost"]]
    for _ in range(3):
        s = o.optimize(4); o.relinearize(); costs.append(s["final_cost"])
        if s["termination_type"] == 0: break
    got = [float(l.split("->")[1].split(",")[0]) for l in out.stdout.splitlines() if l.startswith("solve ")]
    assert len(got) == len(costs)
    np.testing.assert_allclose(got, costs, rtol=1e-6)
    g = o.getGlobals()
    assert abs(float(open(os.path.join(res, "scale.txt")).read()) - (g[18] + g[0])) < 1e-7
    assert abs(float(open(os.path.join(res, "spline_globalsensor_time_offset.txt")).read()) - g[17]) < 1e-7
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "p_bp.txt")), g[10:13], atol=1e-6)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "points3D.txt")), o.get3DPoints(), atol=1e-5)
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = (len(sc["knot_q"]) - order + 1) * sc["dt_ns"] + sc["start_t_ns"] - 1_000_000
    t_ns = np.arange(sc["start_t_ns"], end_ns, 20_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:4], ref[:, :3], atol=1e-5)                     # north_star: 1 mm / 0.01 deg
    assert (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(ref[:, 3:])).magnitude().max() < np.deg2rad(0.001)
    for fn in ("initial_spline.txt", "initial_points3D.txt", "initial_scale.txt", "initial_p_bp.txt", "gravity.txt", "initial_reprojection_errors.txt"):
        assert os.path.exists(os.path.join(res, fn))


def _write_vi_dataset(d, sc, order, T_wg_q, T_wg_p, flags):
    """the gv dataset plus imu.txt and the keys of full_batch_optimization_vi/continuous_time/*.yaml (T_wg inline over several lines)"""
    _, _, g = _write_gv_dataset(d, sc, order, T_wg_q, T_wg_p, flags)
    with open(os.path.join(d, "imu.txt"), "w") as f:
        f.write("# id timestamp wx wy wz ax ay az\n")
        for i, (t, w, a) in enumerate(zip(sc["imu_t_ns"], sc["imu_gyr"], sc["imu_acc"])):
            f.write(f"{i} " + " ".join(f"{v:.17g}" for v in [t * 1e-9, *w, *a]) + "\n")
    R = Rot.from_quat(T_wg_q).as_matrix()
    rows = [", ".join(f"{v:.17g}" for v in list(R[r]) + [T_wg_p[r]]) for r in range(3)] + ["0.0, 0.0, 0.0, 1.0"]
    dt_ms = int(sc["dt_ns"] * 1e-9 * 1000)
    yml = os.path.join(d, "vi.yaml")
    with open(yml, "w") as f:
        f.write(f"""dataset: Custom
cam_id: 0
camimu_calib_fn: "{d}/calib.yaml"
imu_fn: "{d}/imu.txt"
colmap_dir: "{d}/colmap/"
colmap_spline_dir: "{d}/fitted"
full_batch_optimization_dir: "{d}/full_batch_vi"
spline_control_nodes_dt_s: {sc['dt_ns'] * 1e-9}
bias_spline_control_nodes_dt_s: {sc['dt_s_bias']}
init_acc_bias: [0.0, 0.0, 0.0]
init_gyro_bias: [0.0, 0.0, 0.0]
sampling_spline_rate: 50.0
grid_cell_size: 30
sigma_reproj_px: {sc['sigma_reproj_px']}
sigma_acc_c: {sc['sigma_acc_c']}
sigma_acc_bias_c: {sc['sigma_acc_bias_c']}
sigma_omega_c: {sc['sigma_omega_c']}
sigma_omega_bias_c: {sc['sigma_omega_bias_c']}
bias_err_freq_hz: {sc['bias_err_freq_hz']}
bias_err_dt_s: {sc['bias_err_dt_s']}
initial_scale: {sc['scale']}
T_wg: [{rows[0]},
       {rows[1]},
       {rows[2]},
       {rows[3]}]
optimize_scale: {'true' if flags[0] else 'false'}
optimize_gravity: {'true' if flags[2] else 'false'}
optimize_camimu_extrinsics: {'true' if flags[3] else 'false'}
n_max_outer_iters: 2
n_max_inner_iters: 4
n_threads: 4
""")
    return yml, os.path.join(d, "full_batch_vi", "continuous_time", f"order_{order}", f"dt_{dt_ms}_ms"), g


def _vi_scene(scene, order):
    flags = (False, True, True, True, True)
    sc = scene.make_scene(order=order, duration_s=3.0, n_landmarks=160, track_len=9, imu_hz=200, gp_hz=0, dt_s_bias=0.25, flags=flags, seed=99 + order)
    return sc, flags


def test_vi_driver_builds_and_fails_loudly_without_gpu(pkg, scene, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_vi_continuous_time.cpp"), str(tmp_path / "vi"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc, flags = _vi_scene(scene, 6)
    yml, _, _ = _write_vi_dataset(str(tmp_path), sc, 6, Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0]), flags)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    assert f"Loaded {len(sc['imu_t_ns'])} imu measurements" in out.stdout and f"Loaded {len(sc['knot_q'])} initial spline knots" in out.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("order", [6, 4])
def test_vi_driver_end_to_end_on_gpu(pkg, oracle, scene, tmp_path, order):
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_vi_continuous_time.cpp"), str(tmp_path / "vi"), pkg)
    sc, flags = _vi_scene(scene, order)
    T_q, T_p = Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0])
    yml, res, (kq_g, kp_g, lm_g) = _write_vi_dataset(str(tmp_path), sc, order, T_q, T_p, flags)
    out = subprocess.run([exe, yml, "0", str(order)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    R_wg = Rot.from_quat(T_q)
    o = oracle.OracleSpline(order, sc["dt_ns"], sc["start_t_ns"])
    o.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"])
    o.setKnots((R_wg * Rot.from_quat(kq_g)).as_quat(), sc["scale"] * R_wg.apply(kp_g) + T_p)
    o.setGlobals(sc["scale"], np.zeros(3), np.array([0.0, 0.0, 9.80665]), 0.0)
    o.setFlags(*flags)
    o.set3DPoints(sc["scale"] * R_wg.apply(lm_g) + T_p)
    o.addGravityNormError()
    o.initializeBiases(sc["dt_s_bias"], np.zeros(3), np.zeros(3))
    t_fr = np.array([_t_of(t) for t in sc["frame_t_ns"]], dtype=np.int64)
    valid, _ = o.computeValidReprojections(t_fr, sc["obs_begin"], sc["uv"], sc["lm_index"], sc["width"], sc["height"], 30)
    n_imu = o.addImu(np.array([_t_of(t) for t in sc["imu_t_ns"]], dtype=np.int64), sc["imu_acc"], sc["imu_gyr"], sc["sigma_acc_c"], sc["sigma_omega_c"])
    ob = np.r_[0, np.cumsum([valid[sc["obs_begin"][f]:sc["obs_begin"][f + 1]].sum() for f in range(len(t_fr))])].astype(np.int32)
    o.addFrames(t_fr, ob, sc["uv"][valid], sc["lm_index"][valid], sc["sigma_reproj_px"])
    o.addImuBiasCost(sc["bias_err_freq_hz"], sc["bias_err_dt_s"], sc["sigma_acc_bias_c"], sc["sigma_omega_bias_c"])
    assert f"IMU samples in the problem: {n_imu} of" in out.stdout and f"valid reprojections: {int(valid.sum())}\n" in out.stdout
    s = o.optimize(1); o.relinearize()
    costs = [s["final_cost"]]
    for _ in range(2):
        s = o.optimize(4); o.relinearize(); costs.append(s["final_cost"])
        if s["termination_type"] == 0: break
    got = [float(l.split("->")[1].split(",")[0]) for l in out.stdout.splitlines() if l.startswith("solve ")]
    assert len(got) == len(costs)
    np.testing.assert_allclose(got, costs, rtol=1e-6)
    g = o.getGlobals()
    assert abs(float(open(os.path.join(res, "cam_imu_time_offset.txt")).read()) - g[16]) < 1e-7
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "gravity.txt")), g[13:16], atol=1e-6)
    T_bc = np.loadtxt(os.path.join(res, "T_bc.txt"))
    R_o = (Rot.from_quat(sc["q_bc"]) * Rot.from_quat(g[1:5])).as_matrix(); p_o = sc["p_bc"] + Rot.from_quat(sc["q_bc"]).apply(g[5:8])
    np.testing.assert_allclose(T_bc[:3, :3], R_o, atol=1e-6); np.testing.assert_allclose(T_bc[:3, 3], p_o, atol=1e-6)
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = (len(sc["knot_q"]) - order + 1) * sc["dt_ns"] + sc["start_t_ns"] - 1_000_000
    t_ns = np.arange(sc["start_t_ns"], end_ns, 20_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:4], ref[:, :3], atol=1e-5)
    assert (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(ref[:, 3:])).magnitude().max() < np.deg2rad(0.001)
    ab = np.loadtxt(os.path.join(res, "acc_bias_spline.txt"))
    np.testing.assert_allclose(ab[:, 1:], o.sampleAt(5, (ab[:, 0] * 1e9).round().astype(np.int64)), atol=1e-6)
    pim = np.loadtxt(os.path.join(res, "predicted_imu.txt"))
    assert pim.shape[1] == 7 and len(pim) > 0.8 * n_imu
    for fn in ("initial_spline.txt", "initial_predicted_imu.txt", "gyro_bias_spline.txt", "acc_errors.txt", "gyro_errors.txt", "reprojection_errors.txt", "points3D.txt", "scale.txt"):
        assert os.path.exists(os.path.join(res, fn))


def _write_ct_dataset(d, sc, order, T_wg_q, T_wg_p, flags, t_gp_init):
    """gv + vi datasets and the alignment results optimize_continuous_time starts from (T_wg.txt, scale.txt, spline_globalsensor_time_offset.txt)"""
    _, _, g = _write_vi_dataset(d, sc, order, T_wg_q, T_wg_p, flags)
    dt_ms = int(sc["dt_ns"] * 1e-9 * 1000)
    al = os.path.join(d, "align", f"order_{order}", f"dt_{dt_ms}_ms")
    _write_pose(os.path.join(al, "T_wg.txt"), T_wg_q, T_wg_p, digits=17)
    open(os.path.join(al, "scale.txt"), "w").write(f"{sc['scale']:.17g}\n")
    open(os.path.join(al, "spline_globalsensor_time_offset.txt"), "w").write(f"{t_gp_init:.17g}\n")
    vi = open(os.path.join(d, "vi.yaml")).read().replace("full_batch_vi", "full_batch_ct")
    yml = os.path.join(d, "ct.yaml")
    with open(yml, "w") as f:
        f.write(vi + f'gp_fn: "{d}/gp.txt"\nspline_global_ref_alignment_dir: "{d}/align"\ngp_std: {sc["gp_std"]}\n'
                f"optimize_pBP: {'true' if flags[1] else 'false'}\noptimize_spline_globsens_time_offset: {'true' if flags[4] else 'false'}\n")
    return yml, os.path.join(d, "full_batch_ct", "continuous_time", f"order_{order}", f"dt_{dt_ms}_ms"), g


def _ct_scene(scene, order):
    # gravity fixed: with GP + IMU on this short synthetic scene the 1e6-weighted gravity-norm residual makes the restated LM reject every
    # step of the driver's short solves (in the oracle as well), which would leave nothing to compare
    flags = (False, True, False, True, True)
    return scene.make_scene(order=order, duration_s=3.0, n_landmarks=160, track_len=9, imu_hz=200, gp_hz=10, dt_s_bias=0.25, flags=flags, seed=55 + order), flags


def test_ct_driver_builds_and_fails_loudly_without_gpu(pkg, scene, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_continuous_time.cpp"), str(tmp_path / "ct"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc, flags = _ct_scene(scene, 6)
    yml, _, _ = _write_ct_dataset(str(tmp_path), sc, 6, Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0]), flags, 0.002)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    assert f"Loaded {len(sc['imu_t_ns'])} imu measurements" in out.stdout and f"Loaded {len(sc['gp_t_ns'])} global measurements" in out.stdout
    assert f"Loaded {len(sc['knot_q'])} initial spline knots" in out.stdout and "Loaded scale: 1" in out.stdout


@pytest.mark.gpu
def test_ct_driver_end_to_end_on_gpu(pkg, oracle, scene, tmp_path):
    order = 6
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_continuous_time.cpp"), str(tmp_path / "ct"), pkg)
    sc, flags = _ct_scene(scene, order)
    T_q, T_p = Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0])
    t_gp_init = 0.002
    yml, res, (kq_g, kp_g, lm_g) = _write_ct_dataset(str(tmp_path), sc, order, T_q, T_p, flags, t_gp_init)
    out = subprocess.run([exe, yml, "0", str(order)], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    o, costs = _ct_oracle(oracle, sc, flags, order, T_q, T_p, kq_g, kp_g, lm_g, t_gp_init)
    got = [float(l.split("->")[1].split(",")[0]) for l in out.stdout.splitlines() if l.startswith("solve ")]
    assert len(got) == len(costs)
    np.testing.assert_allclose(got, costs, rtol=1e-6)
    g = o.getGlobals()
    assert abs(float(open(os.path.join(res, "cam_imu_time_offset.txt")).read()) - g[16]) < 1e-7
    assert abs(float(open(os.path.join(res, "spline_globalsensor_time_offset.txt")).read()) - g[17]) < 1e-7
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "p_bp.txt")), g[10:13], atol=1e-6)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "gravity.txt")), g[13:16], atol=1e-6)
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = (len(sc["knot_q"]) - order + 1) * sc["dt_ns"] + sc["start_t_ns"] - 1_000_000
    t_ns = np.arange(sc["start_t_ns"], end_ns, 20_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:4], ref[:, :3], atol=1e-5)
    assert (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(ref[:, 3:])).magnitude().max() < np.deg2rad(0.001)
    for fn in ("initial_T_bc.txt", "initial_gravity.txt", "global_position_errors.txt", "predicted_imu.txt", "acc_bias_spline.txt", "points3D.txt", "T_bc.txt"):
        assert os.path.exists(os.path.join(res, fn))


def _ct_oracle(oracle, sc, flags, order, T_q, T_p, kq_g, kp_g, lm_g, t_gp_init):
    """src/optimize_continuous_time.cpp:405-692 on the oracle"""
    R_wg = Rot.from_quat(T_q)
    o = oracle.OracleSpline(order, sc["dt_ns"], sc["start_t_ns"])
    o.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"])
    o.setKnots((R_wg * Rot.from_quat(kq_g)).as_quat(), sc["scale"] * R_wg.apply(kp_g) + T_p)
    o.setGlobals(sc["scale"], sc["p_bp"], np.array([0.0, 0.0, 9.80665]), t_gp_init)
    o.setFlags(*flags)
    o.set3DPoints(sc["scale"] * R_wg.apply(lm_g) + T_p)
    if flags[2]: o.addGravityNormError()
    o.initializeBiases(sc["dt_s_bias"], np.zeros(3), np.zeros(3))
    t_fr = np.array([_t_of(t) for t in sc["frame_t_ns"]], dtype=np.int64)
    valid, _ = o.computeValidReprojections(t_fr, sc["obs_begin"], sc["uv"], sc["lm_index"], sc["width"], sc["height"], 30)
    o.addGp(np.array([_t_of(t) for t in sc["gp_t_ns"]], dtype=np.int64), sc["gp_pos"], sc["gp_std"])
    o.addImu(np.array([_t_of(t) for t in sc["imu_t_ns"]], dtype=np.int64), sc["imu_acc"], sc["imu_gyr"], sc["sigma_acc_c"], sc["sigma_omega_c"])
    ob = np.r_[0, np.cumsum([valid[sc["obs_begin"][f]:sc["obs_begin"][f + 1]].sum() for f in range(len(t_fr))])].astype(np.int32)
    o.addFrames(t_fr, ob, sc["uv"][valid], sc["lm_index"][valid], sc["sigma_reproj_px"])
    o.addImuBiasCost(sc["bias_err_freq_hz"], sc["bias_err_dt_s"], sc["sigma_acc_bias_c"], sc["sigma_omega_bias_c"])
    s = o.optimize(1); o.relinearize()
    costs = [s["final_cost"]]
    for _ in range(2):
        s = o.optimize(4); o.relinearize(); costs.append(s["final_cost"])
        if s["termination_type"] == 0: break
    return o, costs


def test_gi_driver_builds_and_fails_loudly_without_gpu(pkg, scene, tmp_path):
    """optimize_gi_continuous_time without a GPU: the build, the parsing and the loud failure (the end-to-end run is the -m gpu twin below)."""
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_gi_continuous_time.cpp"), str(tmp_path / "gi"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc, flags = _ct_scene(scene, 6)
    d = str(tmp_path)
    _write_ct_dataset(d, sc, 6, Rot.from_quat([0, 0, 0, 1.0]).as_quat(), np.zeros(3), flags, 0.0)
    os.makedirs(os.path.join(d, "gpfit", "order_6", "dt_100_ms"))
    os.replace(os.path.join(d, "fitted", "order_6", "dt_100_ms", "spline_knots.txt"), os.path.join(d, "gpfit", "order_6", "dt_100_ms", "spline_knots.txt"))
    yml = os.path.join(d, "gi.yaml")
    with open(yml, "w") as f:
        f.write(f'dataset: EuRoC\nimu_fn: "{d}/imu.txt"\ngp_fn: "{d}/gp.txt"\nspline_dir: "{d}/gpfit"\nfull_batch_optimization_dir: "{d}/full_batch_gi"\n'
                f"spline_control_nodes_dt_s: 0.1\nsampling_spline_rate: 50.0\nbias_spline_control_nodes_dt_s: {sc['dt_s_bias']}\ninit_acc_bias: [0.0, 0.0, 0.0]\n"
                f"init_gyro_bias: [0.0, 0.0, 0.0]\nsigma_acc_c: 0.1\nsigma_acc_bias_c: 0.002\nsigma_omega_c: 0.05\nsigma_omega_bias_c: 4.0e-05\n"
                f"bias_err_freq_hz: {sc['bias_err_freq_hz']}\nbias_err_dt_s: {sc['bias_err_dt_s']}\ngp_std: 0.1\noptimize_gravity: false\noptimize_pBP: true\n"
                f"optimize_spline_globsens_time_offset: true\nn_max_outer_iters: 2\nn_max_inner_iters: 4\nn_threads: 4\n")
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    assert f"Loaded {len(sc['imu_t_ns'])} imu measurements" in out.stdout and f"Loaded {len(sc['gp_t_ns'])} global measurements" in out.stdout
    assert f"Loaded {len(sc['knot_q'])} initial spline knots" in out.stdout


def _write_gi_dataset(d, sc):
    flags = (False, True, False, True, True)
    _write_ct_dataset(d, sc, 6, Rot.from_quat([0, 0, 0, 1.0]).as_quat(), np.zeros(3), flags, 0.0)
    sub = os.path.join(d, "gpfit", "order_6", "dt_100_ms")
    os.makedirs(sub)
    _write_traj(os.path.join(sub, "spline_knots.txt"), [(sc["start_t_ns"] + i * sc["dt_ns"]) * 1e-9 for i in range(len(sc["knot_q"]))], sc["knot_q"], sc["knot_p"])
    yml = os.path.join(d, "gi.yaml")
    with open(yml, "w") as f:
        f.write(f'dataset: EuRoC\nimu_fn: "{d}/imu.txt"\ngp_fn: "{d}/gp.txt"\nspline_dir: "{d}/gpfit"\nfull_batch_optimization_dir: "{d}/full_batch_gi"\n'
                f"spline_control_nodes_dt_s: 0.1\nsampling_spline_rate: 50.0\nbias_spline_control_nodes_dt_s: {sc['dt_s_bias']}\ninit_acc_bias: [0.0, 0.0, 0.0]\n"
                f"init_gyro_bias: [0.0, 0.0, 0.0]\nsigma_acc_c: 0.1\nsigma_acc_bias_c: 0.002\nsigma_omega_c: 0.05\nsigma_omega_bias_c: 4.0e-05\n"
                f"bias_err_freq_hz: {sc['bias_err_freq_hz']}\nbias_err_dt_s: {sc['bias_err_dt_s']}\ngp_std: 0.1\noptimize_gravity: false\noptimize_pBP: true\n"
                f"optimize_spline_globsens_time_offset: true\nn_max_outer_iters: 2\nn_max_inner_iters: 4\nn_threads: 4\n")
    return yml, os.path.join(d, "full_batch_gi", "continuous_time", "order_6", "dt_100_ms")


@pytest.mark.gpu
def test_gi_driver_end_to_end_on_gpu(pkg, oracle, scene, tmp_path):
    """optimize_gi_continuous_time (src/optimize_gi_continuous_time.cpp:303-431): GP + IMU + bias cost, no vision, scale left at the constructor's
    0 (so p_bp has no effect — reproduced), only the spline-imu offset re-linearised between the solves.  The tool's solve costs and result
    files against the oracle taken through the same steps."""
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_gi_continuous_time.cpp"), str(tmp_path / "gi"), pkg)
    sc, _ = _ct_scene(scene, 6)
    yml, res = _write_gi_dataset(str(tmp_path), sc)
    out = subprocess.run([exe, yml, "0", "6"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    o = oracle.OracleSpline(6, sc["dt_ns"], sc["start_t_ns"])
    o.setKnots(sc["knot_q"], sc["knot_p"])
    o.setGlobals(0.0, np.zeros(3), np.array([0.0, 0.0, 9.80665]), 0.0)
    o.setFlags(False, True, False, True, True)
    o.set3DPoints(np.zeros((0, 3)))
    o.initializeBiases(sc["dt_s_bias"], np.zeros(3), np.zeros(3))
    o.addGp(np.array([_t_of(t) for t in sc["gp_t_ns"]], dtype=np.int64), sc["gp_pos"], sc["gp_std"])
    o.addImu(np.array([_t_of(t) for t in sc["imu_t_ns"]], dtype=np.int64), sc["imu_acc"], sc["imu_gyr"], sc["sigma_acc_c"], sc["sigma_omega_c"])
    o.addImuBiasCost(sc["bias_err_freq_hz"], sc["bias_err_dt_s"], sc["sigma_acc_bias_c"], sc["sigma_omega_bias_c"])
    s = o.optimize(1); o.relinearizeMask(1)
    costs = [s["final_cost"]]
    for _ in range(2):
        s = o.optimize(4); o.relinearizeMask(1); costs.append(s["final_cost"])
        if s["termination_type"] == 0: break
    got = [float(l.split("->")[1].split(",")[0]) for l in out.stdout.splitlines() if l.startswith("solve ")]
    assert len(got) == len(costs)
    np.testing.assert_allclose(got, costs, rtol=1e-6)
    g = o.getGlobals()
    assert abs(float(open(os.path.join(res, "globalsensor_imu_time_offset.txt")).read()) - g[16]) < 1e-7
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "gravity.txt")), g[13:16], atol=1e-6)
    tr = np.loadtxt(os.path.join(res, "spline.txt"))
    end_ns = int(float((len(sc["knot_q"]) - 6 + 1) * sc["dt_ns"] + sc["start_t_ns"]) - 1e6)
    t_ns = np.arange(sc["start_t_ns"], end_ns, 20_000_000, dtype=np.int64)
    assert tr.shape == (len(t_ns), 8)
    ref = o.sampleAt(0, t_ns)
    np.testing.assert_allclose(tr[:, 1:4], ref[:, :3], atol=1e-5)
    assert (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(ref[:, 3:])).magnitude().max() < np.deg2rad(0.001)
    for fn in ("acc_errors.txt", "gyro_errors.txt", "global_position_errors.txt", "predicted_imu.txt", "acc_bias_spline.txt", "gyro_bias_spline.txt", "p_bp.txt"):
        assert os.path.exists(os.path.join(res, fn))


def test_align_colmap_to_imu_tool(pkg, scene, tmp_path):
    """align_colmap_to_imu is host-only (no solver): the whole tool is checked on the CPU against a numpy restatement of
    src/align_colmap_to_imu.cpp:51-82, 158-190."""
    d = str(tmp_path)
    rng = np.random.default_rng(17)
    sc, flags = _vi_scene(scene, 6)
    _write_vi_dataset(d, sc, 6, Rot.from_quat([0, 0, 0, 1.0]).as_quat(), np.zeros(3), flags)        # calib.yaml
    t_imu = 500.0 + np.arange(60) / 200.0
    R_true = Rot.from_rotvec([0.3, -0.5, 0.2])
    acc = R_true.inv().apply(np.array([0.0, 0.0, 9.81])) + rng.normal(size=(60, 3)) * 0.05
    with open(os.path.join(d, "imu_a.txt"), "w") as f:
        f.write("# id t wx wy wz ax ay az\n")
        for i in range(60):
            f.write(f"{i} {t_imu[i]:.17g} 0 0 0 " + " ".join(f"{v:.17g}" for v in acc[i]) + "\n")
    q_gc = Rot.from_rotvec(rng.normal(size=(3, 3))).as_quat(); p_gc = rng.normal(size=(3, 3))
    t_cam = [t_imu[7] + 1e-4, t_imu[20], t_imu[40]]
    os.makedirs(os.path.join(d, "colmap_a"))
    _write_traj(os.path.join(d, "colmap_a", "colmap_cam_estimates.txt"), t_cam, q_gc, p_gc)
    yml = os.path.join(d, "aci.yaml")
    open(yml, "w").write(f'dataset: Custom\ncam_id: 0\ncamimu_calib_fn: "{d}/calib.yaml"\nimu_fn: "{d}/imu_a.txt"\ncolmap_dir: "{d}/colmap_a/"\nresult_dir: "{d}/aci_out"\n')
    exe = _gxx(os.path.join(ROOT, "drivers", "align_colmap_to_imu.cpp"), str(tmp_path / "aci"), pkg)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 0, out.stdout + out.stderr
    # restatement: samples with t >= first image (index 8 on), edges [m0.t, m10.t] -> samples 0..11 of those
    sel = acc[8:][:12]
    e = sel.mean(axis=0); e /= np.linalg.norm(e)
    ax = np.cross([0.0, 0.0, 1.0], e); ax /= np.linalg.norm(ax)
    R_IB = Rot.from_rotvec(-ax * np.arccos(e[2]))
    T_IB = np.loadtxt(os.path.join(d, "aci_out", "T_IB.txt"))
    np.testing.assert_allclose(T_IB[:3, :3], R_IB.as_matrix(), atol=2e-9); np.testing.assert_array_equal(T_IB[:3, 3], 0.0)
    np.testing.assert_allclose(R_IB.apply(e), [0, 0, 1], atol=1e-12)          # gravity direction ends up on +z
    R_bc, p_bc = Rot.from_quat(sc["q_bc"]), sc["p_bc"]                         # T_cam_imu = T_b_c^-1
    R_ci, p_ci = R_bc.inv(), -R_bc.inv().apply(p_bc)
    R_gb = Rot.from_quat(q_gc[0]) * R_ci; p_gb = Rot.from_quat(q_gc[0]).apply(p_ci) + p_gc[0]
    R_ig = R_IB * R_gb.inv(); p_ig = -R_ig.apply(p_gb)
    T_IG = np.loadtxt(os.path.join(d, "aci_out", "T_IG.txt"))
    np.testing.assert_allclose(T_IG[:3, :3], R_ig.as_matrix(), atol=2e-9); np.testing.assert_allclose(T_IG[:3, 3], p_ig, rtol=2e-8, atol=2e-9)


# ---- optimize_discrete_time (BASELINE configs[3]): src/optimize_discrete_time.cpp:195-726 ---------------------------------------------
def _write_dt_dataset(d, sc, dd, T_wg_q, T_wg_p, flags):
    """A scene of scene.make_scene + its discrete-time view (scene.dt_scene_from: the perturbed frame poses) as the files
    optimize_discrete_time reads: EuRoC-format calibration, imu.txt, gp.txt, the COLMAP export (images.bin written in REVERSED order, so
    that the driver's sort by image name matters), the COLMAP camera trajectory T_GC, T_wg.txt / scale.txt of the alignment step."""
    _write_vi_dataset(d, sc, 6, T_wg_q, T_wg_p, (flags[0], flags[1], False, True, flags[3]))       # calib.yaml, imu.txt, gp.txt (+ files this tool ignores)
    R_wg = Rot.from_quat(T_wg_q); R_bc = Rot.from_quat(sc["q_bc"])
    F = len(sc["frame_t_ns"])
    R_wb = Rot.from_quat(dd["poses"][:, :4]); p_wb = dd["poses"][:, 4:]
    R_gb = R_wg.inv() * R_wb; p_gb = R_wg.inv().apply(p_wb - T_wg_p) / sc["scale"]
    R_gc = R_gb * R_bc; p_gc = R_gb.apply(sc["p_bc"]) + p_gb                                          # T_GC = T_GB T_BC
    with open(os.path.join(d, "colmap_traj.txt"), "w") as f:
        f.write("# timestamp tx ty tz qx qy qz qw\n")
        for i in range(F):
            f.write(" ".join(f"{v:.17g}" for v in [sc["frame_t_ns"][i] * 1e-9, *p_gc[i], *R_gc[i].as_quat()]) + "\n")
    lm_g = R_wg.inv().apply(sc["landmarks"] - T_wg_p) / sc["scale"]
    ids = [10 + 3 * i for i in range(len(lm_g))]
    cams = [dict(id=1, model=sc["cam_model"], w=sc["width"], h=sc["height"], params=list(sc["intr"]) + list(sc["dist"]))]
    images = []
    for f_ in reversed(range(F)):
        b, e = sc["obs_begin"][f_], sc["obs_begin"][f_ + 1]
        pts = [(float(sc["uv"][q, 0]), float(sc["uv"][q, 1]), ids[sc["lm_index"][q]]) for q in range(b, e)]
        images.append(dict(id=F - f_, qwxyz=[1.0, 0, 0, 0], t=[0.0, 0, 0], cam=1, name=f"img_{f_:06d}.png", pts=pts))
    points = [dict(id=ids[i], xyz=list(lm_g[i]), err=0.5, track=[]) for i in range(len(lm_g))]
    os.makedirs(os.path.join(d, "colmap_dt"))
    _write_colmap_bin(os.path.join(d, "colmap_dt"), cams, images, points)
    os.makedirs(os.path.join(d, "align_dt"))
    _write_pose(os.path.join(d, "align_dt", "T_wg.txt"), T_wg_q, T_wg_p, digits=17)
    open(os.path.join(d, "align_dt", "scale.txt"), "w").write(f"{sc['scale']:.17g}\n")
    cam_ts = sc["frame_t_ns"] * 1e-9
    start_t, end_t = cam_ts[1] - 1e-4, cam_ts[-2] + 1e-4                                               # the first and the last frame stay out
    yml = os.path.join(d, "dt.yaml")
    with open(yml, "w") as f:
        f.write(f"""dataset: Custom
cam_id: 0
camimu_calib_fn: "{d}/calib.yaml"
imu_fn: "{d}/imu.txt"
gp_fn: "{d}/gp.txt"
colmap_dir: "{d}/colmap_dt/"
colmap_fn: "{d}/colmap_traj.txt"
alignment_dir: "{d}/align_dt"
full_batch_optimization_dir: "{d}/full_batch_dt"
start_t: {start_t:.17g}
end_t: {end_t:.17g}
grid_cell_size: 30
gp_std: {sc['gp_std']}
sigma_reproj_px: {sc['sigma_reproj_px']}
sigma_acc_c: {sc['sigma_acc_c']}
sigma_acc_bias_c: {sc['sigma_acc_bias_c']}
sigma_omega_c: {sc['sigma_omega_c']}
sigma_omega_bias_c: {sc['sigma_omega_bias_c']}
t_offset_cam_gp_init: {sc['t_s_gp']:.17g}
p_BP: [{sc['p_bp'][0]:.17g}, {sc['p_bp'][1]:.17g}, {sc['p_bp'][2]:.17g}]
optimize_scale: {'true' if flags[0] else 'false'}
optimize_pBP: {'true' if flags[1] else 'false'}
optimize_t_offset_globsens_imu: {'true' if flags[3] else 'false'}
n_max_iters: 8
n_threads: 4
""")
    return yml, os.path.join(d, "full_batch_dt", "discrete_time"), (R_gc, p_gc, lm_g)


def _dt_scene(scene):
    sc = scene.make_scene(order=6, duration_s=3.0, n_landmarks=140, track_len=8, imu_hz=200, gp_hz=10, cam_hz=10.0, seed=23, flags=(False, True, True, True, True),
                          outlier_frac=0.03)
    dd = scene.dt_scene_from(sc, perturb=(0.002, 0.05))
    return sc, dd, (False, True, True, True)       # scale, p_BP, extrinsics (the driver never sets it: constructor default true), time offset


def test_dt_driver_builds_and_fails_loudly_without_gpu(pkg, scene, tmp_path):
    import torch
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_discrete_time.cpp"), str(tmp_path / "dt"), pkg)
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    sc, dd, flags = _dt_scene(scene)
    yml, _, _ = _write_dt_dataset(str(tmp_path), sc, dd, Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0]), flags)
    out = subprocess.run([exe, yml], capture_output=True, text=True, timeout=60)
    assert out.returncode == 3 and "no CPU fallback" in out.stderr, out.stdout + out.stderr
    # everything up to the construction of the estimator ran: calibration, measurements, COLMAP export, trajectory, alignment
    assert f"Loaded {len(sc['imu_t_ns'])} imu measurements" in out.stdout and f"Loaded {len(sc['gp_t_ns'])} global measurements" in out.stdout
    assert f"Full trajectory contains: {len(sc['frame_t_ns'])} images" in out.stdout and f"Full 3D map contains: {len(sc['landmarks'])} 3D points" in out.stdout


@pytest.mark.gpu
def test_dt_driver_end_to_end_on_gpu(pkg, oracle, scene, tmp_path):
    exe = _gxx(os.path.join(ROOT, "drivers", "optimize_discrete_time.cpp"), str(tmp_path / "dt"), pkg)
    sc, dd, flags = _dt_scene(scene)
    T_q, T_p = Rot.from_rotvec([0.2, -0.4, 1.0]).as_quat(), np.array([3.0, -2.0, 1.0])
    yml, res, (R_gc, p_gc, lm_g) = _write_dt_dataset(str(tmp_path), sc, dd, T_q, T_p, flags)
    out = subprocess.run([exe, yml, "0"], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    # the oracle through the same steps (src/optimize_discrete_time.cpp:339-570)
    R_wg = Rot.from_quat(T_q); R_bc = Rot.from_quat(sc["q_bc"])
    sl = slice(1, len(sc["frame_t_ns"]) - 1)
    cam_ts = (sc["frame_t_ns"] * 1e-9)[sl]
    R_gb = R_gc[sl] * R_bc.inv(); p_gb = p_gc[sl] - R_gb.apply(sc["p_bc"])                             # T_GB = T_GC T_CB
    poses = np.concatenate([(R_wg * R_gb).as_quat(), sc["scale"] * R_wg.apply(p_gb) + T_p], axis=1)
    o = oracle.OracleDiscrete()
    o.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"], sc["width"], sc["height"])
    o.setImu(sc["sigma_omega_c"], sc["sigma_acc_c"], sc["sigma_omega_bias_c"], sc["sigma_acc_bias_c"], np.array([0.0, 0.0, 9.80665]))
    o.initializeStates(poses, cam_ts)
    o.initializeLinearVelocity()
    t_gp_imu_init = sc["t_offset_cam_imu"] - sc["t_s_gp"]
    o.setGlobals(sc["scale"], sc["p_bp"], t_gp_imu_init)
    o.setFlags(*flags)
    o.set3DPoints(sc["scale"] * R_wg.apply(lm_g) + T_p)
    b0, b1 = sc["obs_begin"][1], sc["obs_begin"][-2]
    pt_begin = (sc["obs_begin"][1:-1] - b0).astype(np.int32); uv = sc["uv"][b0:b1]; li = sc["lm_index"][b0:b1]
    valid, nv, vel = o.computeValidReprojections(pt_begin, uv, li, sc["width"], sc["height"], 30, cam_ts)
    ts_imu = cam_ts + sc["t_offset_cam_imu"]
    ti = sc["imu_t_ns"] * 1e-9
    n_fac = 0
    for f in range(len(cam_ts) - 1):
        t0, t1 = ts_imu[f], ts_imu[f + 1]
        if ti[0] > t0 or ti[-1] < t1: continue
        k0 = int(np.searchsorted(ti, t0, side="right")) - 1; k1 = min(int(np.searchsorted(ti, t1, side="right")), ti.size - 1)
        o.addImuMeasurements(f, ti[k0:k1 + 1], sc["imu_gyr"][k0:k1 + 1], sc["imu_acc"][k0:k1 + 1]); n_fac += 1
    ob = np.r_[0, np.cumsum([valid[pt_begin[f]:pt_begin[f + 1]].sum() for f in range(len(cam_ts))])].astype(np.int32)
    o.addFrames(np.arange(len(cam_ts), dtype=np.int32), ob, uv[valid], vel[valid], li[valid], sc["sigma_reproj_px"])
    t_gp = sc["gp_t_ns"] * 1e-9
    f0 = np.array([next((i - 1 for i in range(len(ts_imu)) if ts_imu[i] > t + t_gp_imu_init), -1) for t in t_gp])
    ok = f0 >= 0
    o.addGlobalSensMeasurements(t_gp[ok], sc["gp_pos"][ok], f0[ok].astype(np.int32), sc["gp_std"])
    assert f"valid reprojections: {int(valid.sum())}, IMU factors: {n_fac}\n" in out.stdout
    assert f"Global positions in the problem: {int(ok.sum())} of {len(t_gp)}\n" in out.stdout
    so = o.optimize(8, 8)
    line = [l for l in out.stdout.splitlines() if l.startswith("solve: ")][0]
    got0, got1 = [float(v) for v in line.split("cost ")[1].split(",")[0].split(" -> ")]
    assert abs(got0 - so["initial_cost"]) <= 1e-8 * so["initial_cost"] and abs(got1 - so["final_cost"]) <= 1e-6 * so["final_cost"]
    assert f"iterations {len(so['iterations'])} (successful {so['num_successful_steps']}, unsuccessful {so['num_unsuccessful_steps']})" in line
    P, sb = o.getStates()
    tr = np.loadtxt(os.path.join(res, "trajectory.txt"))
    assert tr.shape == (len(cam_ts), 8)
    np.testing.assert_allclose(tr[:, 0], ts_imu, atol=1e-9)
    np.testing.assert_allclose(tr[:, 1:4], P[:, 4:], atol=1e-5)                                        # north_star: 1 mm / 0.01 deg
    assert (Rot.from_quat(tr[:, 4:]).inv() * Rot.from_quat(P[:, :4])).magnitude().max() < np.deg2rad(0.001)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "velocity.txt"))[:, 1:], sb[:, 0:3], atol=1e-5)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "gyro_bias.txt"))[:, 1:], sb[:, 3:6], atol=1e-6)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "acc_bias.txt"))[:, 1:], sb[:, 6:9], atol=1e-5)
    g = o.getGlobals()
    assert abs(float(open(os.path.join(res, "scale.txt")).read()) - (sc["scale"] + g[0])) < 1e-7
    assert abs(float(open(os.path.join(res, "cam_imu_time_offset.txt")).read()) - (sc["t_offset_cam_imu"] + g[8])) < 1e-7
    assert abs(float(open(os.path.join(res, "globalsensor_imu_time_offset.txt")).read()) - (t_gp_imu_init + g[9])) < 1e-7
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "p_BP.txt")), g[10:13], atol=1e-6)
    T_bc = np.loadtxt(os.path.join(res, "T_BC.txt"))
    R_o = (R_bc * Rot.from_quat(g[1:5])).as_matrix(); p_o = sc["p_bc"] + R_bc.apply(g[5:8])
    np.testing.assert_allclose(T_bc[:3, :3], R_o, atol=1e-6); np.testing.assert_allclose(T_bc[:3, 3], p_o, atol=1e-6)
    np.testing.assert_allclose(np.loadtxt(os.path.join(res, "points3D.txt")), o.get3DPoints(), atol=1e-5)
    init = np.loadtxt(os.path.join(res, "initial_trajectory.txt"))
    np.testing.assert_allclose(init[:, 1:4], poses[:, 4:], atol=1e-9)
    assert len(np.loadtxt(os.path.join(res, "reprojection_errors.txt"))) == int(valid.sum())
    assert len(np.loadtxt(os.path.join(res, "global_position_errors.txt"))) == int(ok.sum())
    for fn in ("initial_velocity.txt", "initial_acc_bias.txt", "initial_gyro_bias.txt", "initial_scale.txt", "initial_p_BP.txt", "initial_T_BC.txt", "gravity.txt",
               "initial_points3D.txt", "initial_reprojection_errors.txt", "initial_global_position_errors.txt"):
        assert os.path.exists(os.path.join(res, fn))
