"""Deterministic synthetic scenes of the shapes named in BASELINE.json (datasets are not available offline).

Follows SURVEY.md §8(d): smooth sinusoidal ground-truth motion in the style of the reference's
RandomMeasurements (include/sensor/imu.h:139-348), landmarks with bounded track length, COLMAP camera
models 4 (rad-tan, EuRoC cam values) / 5 (equidistant, UZH-FPV snapdragon values), 1 px pixel noise,
IMU = spline-consistent body rates / specific force + bias + noise, GP = antenna position + noise,
initial state = ground truth perturbed.  Pure numpy/scipy; used by tests, bench.py and the drivers'
scene exporter.  Nothing here touches the oracle.
"""
import numpy as np
from scipy.spatial.transform import Rotation as Rot

G_W = np.array([0.0, 0.0, 9.80665])  # imu_calib.h:35


def camera_preset(name):
    if name == "euroc":  # datasets/EuRoC/calib/Vicon_room/calib.yaml (cam values rounded)
        return dict(cam_model=4, intr=np.array([457.587, 456.134, 379.999, 255.238]), dist=np.array([-0.28368365, 0.07451284, -0.00010473, -3.55590700e-05]),
                    width=752, height=480,
                    q_bc=Rot.from_matrix(np.array([[0.0125552670891, -0.999755099723, 0.0182237714554],
                                                   [0.999598781151, 0.0130119051815, 0.0251588363115],
                                                   [-0.0253898008918, 0.0179005838253, 0.999517347078]])).as_quat(),
                    p_bc=np.array([-0.0198435579556, 0.0453689425024, 0.00786212447038]), t_offset_cam_imu=0.0)
    if name == "fpv":  # UZH-FPV snapdragon cam0, equidistant
        return dict(cam_model=5, intr=np.array([278.66723066, 278.48991409, 319.75221200, 241.96858910]),
                    dist=np.array([-0.013721808247486035, 0.020727425669427896, -0.012786476702685545, 0.0025242267320687625]),
                    width=640, height=480,
                    q_bc=Rot.from_rotvec([1.2, -1.2, 1.2]).as_quat(), p_bc=np.array([0.02, -0.03, 0.01]), t_offset_cam_imu=-0.0167)
    if name == "fov":
        return dict(cam_model=7, intr=np.array([300.0, 300.0, 320.0, 240.0]), dist=np.array([0.9, 0.0, 0.0, 0.0]), width=640, height=480,
                    q_bc=Rot.from_rotvec([0.0, 0.0, 0.3]).as_quat(), p_bc=np.array([0.01, 0.02, 0.03]), t_offset_cam_imu=0.0)
    raise ValueError(name)


def project(cam_model, intr, dist, pc):
    """numpy restatement of Camera::project for data generation (camera.h:79-90, camera_calib.h)."""
    x = pc[:, 0] / pc[:, 2]
    y = pc[:, 1] / pc[:, 2]
    if cam_model == 4:
        k1, k2, p1, p2 = dist
        r2 = x * x + y * y
        rad = k1 * r2 + k2 * r2 * r2
        xd = x + x * rad + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
        yd = y + y * rad + 2 * p2 * x * y + p1 * (r2 + 2 * y * y)
    elif cam_model == 5:
        r = np.sqrt(x * x + y * y)
        th = np.arctan(r)
        thd = th * (1 + dist[0] * th**2 + dist[1] * th**4 + dist[2] * th**6 + dist[3] * th**8)
        sc = thd / np.maximum(r, 1e-12)
        xd, yd = x * sc, y * sc
    elif cam_model == 7:
        s = dist[0]
        r = np.sqrt(x * x + y * y)
        f = np.where(r < 0.001, 1.0, np.arctan(r * 2 * np.tan(s / 2)) / (s * np.maximum(r, 1e-12)))
        xd, yd = x * f, y * f
    else:
        raise ValueError(cam_model)
    return np.stack([intr[0] * xd + intr[2], intr[1] * yd + intr[3]], axis=1)


class Trajectory:
    """Analytic smooth trajectory: p(t) sum of sinusoids, R(t) = Rz(yaw) Ry(pitch) Rx(roll)."""

    def __init__(self, rng, amp_scale=1.0):
        self.A = np.array([3.0, 2.0, 0.6]) * amp_scale
        self.W = rng.uniform(0.25, 0.6, 3)
        self.P = rng.uniform(0, 2 * np.pi, 3)
        self.yaw_rate = rng.uniform(0.1, 0.25)
        self.rw = rng.uniform(0.4, 0.9, 3)
        self.rp = rng.uniform(0, 2 * np.pi, 3)

    def pos(self, t):
        return self.A * np.sin(np.outer(t, self.W) + self.P)

    def acc(self, t):
        return -self.A * self.W**2 * np.sin(np.outer(t, self.W) + self.P)

    def rot(self, t):
        yaw = self.yaw_rate * t + 0.3 * np.sin(self.rw[0] * t + self.rp[0])
        pitch = 0.15 * np.sin(self.rw[1] * t + self.rp[1])
        roll = 0.10 * np.sin(self.rw[2] * t + self.rp[2])
        return Rot.from_euler("ZYX", np.stack([yaw, pitch, roll], axis=1))

    def omega_body(self, t, h=1e-4):
        return (self.rot(t - h).inv() * self.rot(t + h)).as_rotvec() / (2 * h)


def make_scene(order=6, dt_s=0.1, duration_s=6.0, cam_hz=20.0, imu_hz=200.0, gp_hz=10.0, n_landmarks=300, track_len=15, cam="euroc",
               seed=20240101, pixel_noise=1.0, outlier_frac=0.0, perturb=(0.02, 0.5, 0.03), dt_s_bias=0.5, bias_err_freq_hz=None,
               bias_err_dt_s=None, flags=(False, True, False, True, True), start_t_ns=1_000_000_000_000, gravity_norm=False, depth=(2.0, 8.0)):
    """flags = (optimize_scale, optimize_pbp, optimize_gravity, optimize_camimu_extrinsics, optimize_t_offset_s_gs)."""
    rng = np.random.default_rng(seed)
    N = order
    dt_ns = int(round(dt_s * 1e9))
    c = camera_preset(cam)
    traj = Trajectory(rng)
    K = int(np.floor(duration_s / dt_s)) + N + 1
    tk = (np.arange(K) - (N - 2) / 2.0) * dt_s
    gt_q = traj.rot(tk).as_quat()
    gt_p = traj.pos(tk)
    # make consecutive quaternions sign-consistent
    for i in range(1, K):
        if np.dot(gt_q[i], gt_q[i - 1]) < 0:
            gt_q[i] = -gt_q[i]

    R_bc = Rot.from_quat(c["q_bc"])
    p_bc = c["p_bc"]
    scale = 1.0
    p_bp = np.array([0.05, -0.02, 0.10])
    t_s_gp = 0.0

    # ---- frames ----
    F = int(np.floor((duration_s - 0.05) * cam_hz))
    tf = np.arange(F) / cam_hz + 0.0123
    tf = tf[tf < duration_s - 0.02]
    F = tf.size
    frame_t_ns = start_t_ns + np.round(tf * 1e9).astype(np.int64)
    Rwb_f = traj.rot(tf)
    pwb_f = traj.pos(tf)
    Rwc_f = Rwb_f * R_bc
    pwc_f = scale * Rwb_f.apply(p_bc) + pwb_f

    # ---- landmarks: born at a frame, seen for track_len frames ----
    L = n_landmarks
    f0 = np.sort(rng.integers(-track_len // 2, max(F - track_len // 2, 1), L))
    fm = np.clip(f0 + track_len // 2, 0, F - 1)
    W_, H_ = c["width"], c["height"]
    px = np.stack([rng.uniform(0.08 * W_, 0.92 * W_, L), rng.uniform(0.08 * H_, 0.92 * H_, L)], axis=1)
    d = rng.uniform(depth[0], depth[1], L)
    xn = (px[:, 0] - c["intr"][2]) / c["intr"][0]
    yn = (px[:, 1] - c["intr"][3]) / c["intr"][1]
    pc = np.stack([xn * d, yn * d, d], axis=1)
    lm_gt = Rwc_f[fm].apply(pc) + pwc_f[fm]

    # candidate (landmark, frame) pairs
    li = np.repeat(np.arange(L), track_len)
    fi = (f0[:, None] + np.arange(track_len)[None, :]).ravel()
    ok = (fi >= 0) & (fi < F)
    li, fi = li[ok], fi[ok]
    pcs = Rwc_f[fi].inv().apply(lm_gt[li] - pwc_f[fi])
    front = pcs[:, 2] > 0.3
    li, fi, pcs = li[front], fi[front], pcs[front]
    uv = project(c["cam_model"], c["intr"], c["dist"], pcs)
    rn = np.hypot(pcs[:, 0] / pcs[:, 2], pcs[:, 1] / pcs[:, 2])
    inside = (uv[:, 0] > 2) & (uv[:, 0] < W_ - 2) & (uv[:, 1] > 2) & (uv[:, 1] < H_ - 2) & (rn > 2e-3) & (rn < 1.5)
    li, fi, uv = li[inside], fi[inside], uv[inside]
    uv = uv + rng.normal(size=uv.shape) * pixel_noise
    if outlier_frac > 0:
        m = rng.random(uv.shape[0]) < outlier_frac
        ang = rng.uniform(0, 2 * np.pi, m.sum())
        mag = rng.uniform(14.0, 40.0, m.sum())
        uv[m] += np.stack([mag * np.cos(ang), mag * np.sin(ang)], axis=1)
    order_ = np.lexsort((li, fi))
    li, fi, uv = li[order_], fi[order_], uv[order_]
    obs_begin = np.zeros(F + 1, dtype=np.int32)
    np.add.at(obs_begin, fi + 1, 1)
    obs_begin = np.cumsum(obs_begin).astype(np.int32)

    # ---- IMU ----
    if imu_hz and imu_hz > 0:
        ti = np.arange(int(np.floor((duration_s - 0.01) * imu_hz))) / imu_hz + 0.0007   # camera-clock time of each sample
        Ri = traj.rot(ti)
        ba = 0.04 * np.sin(np.outer(ti, [0.05, 0.07, 0.03]) + [0.1, 0.5, 0.9])
        bg = 0.01 * np.sin(np.outer(ti, [0.04, 0.06, 0.02]) + [0.3, 0.2, 0.7])
        imu_acc = Ri.inv().apply(traj.acc(ti) + G_W) + ba + rng.normal(size=(ti.size, 3)) * 0.02
        imu_gyr = traj.omega_body(ti) + bg + rng.normal(size=(ti.size, 3)) * 0.002
        imu_t_ns = start_t_ns + np.round((ti + c["t_offset_cam_imu"]) * 1e9).astype(np.int64)  # t_imu = t_cam + offset (bspline_so3r3.h:33-37)
    else:
        imu_t_ns = np.zeros(0, dtype=np.int64); imu_acc = np.zeros((0, 3)); imu_gyr = np.zeros((0, 3))

    # ---- GP ----
    if gp_hz and gp_hz > 0:
        tg = np.arange(int(np.floor((duration_s - 0.01) * gp_hz))) / gp_hz + 0.0031
        gp_pos = scale * traj.rot(tg).apply(p_bp) + traj.pos(tg) + rng.normal(size=(tg.size, 3)) * 0.01
        gp_t_ns = start_t_ns + np.round((tg + t_s_gp) * 1e9).astype(np.int64)
    else:
        gp_t_ns = np.zeros(0, dtype=np.int64); gp_pos = np.zeros((0, 3))

    # ---- initial state = perturbed ground truth ----
    sp, sr_deg, sl = perturb
    knot_p = gt_p + rng.normal(size=gt_p.shape) * sp
    knot_q = (Rot.from_quat(gt_q) * Rot.from_rotvec(rng.normal(size=(K, 3)) * np.deg2rad(sr_deg))).as_quat()
    for i in range(K):
        if np.dot(knot_q[i], gt_q[i]) < 0:
            knot_q[i] = -knot_q[i]
    landmarks = lm_gt + rng.normal(size=lm_gt.shape) * sl

    if bias_err_dt_s is None:
        bias_err_dt_s = 4.0 * dt_s_bias * 2.0   # reference yaml: 4.0 s with 0.5 s bias knots
    if bias_err_freq_hz is None:
        bias_err_freq_hz = 1.0 / bias_err_dt_s

    return dict(
        order=N, dt_ns=dt_ns, start_t_ns=int(start_t_ns), cam_model=c["cam_model"], intr=c["intr"], dist=c["dist"], q_bc=c["q_bc"], p_bc=p_bc,
        t_offset_cam_imu=c["t_offset_cam_imu"], width=W_, height=H_, scale=scale, p_bp=p_bp, g=G_W.copy(), t_s_gp=t_s_gp, flags=tuple(flags),
        knot_q=np.ascontiguousarray(knot_q), knot_p=np.ascontiguousarray(knot_p), landmarks=np.ascontiguousarray(landmarks),
        gt_knot_q=gt_q, gt_knot_p=gt_p, gt_landmarks=lm_gt, gt_frame_q=Rwb_f.as_quat(), gt_frame_p=pwb_f,
        frame_t_ns=frame_t_ns, obs_begin=obs_begin, uv=np.ascontiguousarray(uv), lm_index=li.astype(np.int32), sigma_reproj_px=1.0,
        imu_t_ns=imu_t_ns, imu_acc=np.ascontiguousarray(imu_acc), imu_gyr=np.ascontiguousarray(imu_gyr), sigma_acc_c=0.1, sigma_omega_c=0.05,
        sigma_acc_bias_c=0.002, sigma_omega_bias_c=4.0e-05, dt_s_bias=dt_s_bias, init_acc_bias=np.zeros(3), init_gyro_bias=np.zeros(3),
        bias_err_freq_hz=bias_err_freq_hz, bias_err_dt_s=bias_err_dt_s,
        gp_t_ns=gp_t_ns, gp_pos=np.ascontiguousarray(gp_pos), gp_std=0.1, gravity_norm=gravity_norm,
    )


# Shapes of BASELINE.json configs.  seeds: 20240101 + cfg.
def config_scene(cfg, **over):
    if cfg == 1:   # optimize_continuous_time, indoor_forward_3_snapdragon-shaped: ~1.5k frames, ~20k landmarks, ~250k obs, IMU 500 Hz, GP 20 Hz
        kw = dict(order=6, dt_s=0.1, duration_s=50.0, cam_hz=30.0, imu_hz=500.0, gp_hz=20.0, n_landmarks=20000, track_len=14, cam="fpv",
                  seed=20240102, flags=(False, True, True, True, True), gravity_norm=True)
    elif cfg == 2:  # optimize_vi_continuous_time, EuRoC MH_01-shaped: 180 s, 20 Hz camera, 200 Hz IMU, ~40k landmarks, ~600k obs
        kw = dict(order=6, dt_s=0.1, duration_s=180.0, cam_hz=20.0, imu_hz=200.0, gp_hz=0.0, n_landmarks=40000, track_len=17, cam="euroc",
                  seed=20240103, flags=(False, True, True, True, True), gravity_norm=True)
    elif cfg == 3:  # optimize_gv_continuous_time: 1 h, 10k frames, 5M obs (500 per frame), GP 10 Hz; 270k landmarks of 28 frames each, of which ~18.5 project inside the image
        kw = dict(order=6, dt_s=0.1, duration_s=3600.0, cam_hz=10000.0 / 3600.0, imu_hz=0.0, gp_hz=10.0, n_landmarks=270000, track_len=28,
                  cam="euroc", seed=20240104, flags=(False, True, False, True, True))
    elif cfg == 4:  # optimize_discrete_time on the configs[2] scene: the same frames / landmarks / observations plus a 100 Hz IMU for the pre-integrated factors
        kw = dict(order=6, dt_s=0.1, duration_s=3600.0, cam_hz=10000.0 / 3600.0, imu_hz=100.0, gp_hz=10.0, n_landmarks=270000, track_len=28,
                  cam="euroc", seed=20240104, flags=(False, True, False, True, True))
    elif cfg == 5:  # knot-spacing / spline-order sweep at a fixed 2 M observations: the configs[2] recipe over 1440 s (4000 frames x 500); pass dt_s, order
        kw = dict(order=6, dt_s=0.1, duration_s=1440.0, cam_hz=10000.0 / 3600.0, imu_hz=0.0, gp_hz=10.0, n_landmarks=108000, track_len=28,
                  cam="euroc", seed=20240106, flags=(False, True, False, True, True))
    else:
        raise ValueError(cfg)
    kw.update(over)
    return make_scene(**kw)


def dt_scene_from(sc, perturb=(0.02, 0.5), seed=7, gyro_random_walk=4.0e-5, acc_random_walk=2.0e-3):
    """The discrete-time problem (BASELINE configs[3]: one SE(3) state + [v, b_g, b_a] per frame) over the data of a scene of make_scene,
    prepared the way src/optimize_discrete_time.cpp:436-570 prepares it: states at the frames (ground truth perturbed), IMU samples per
    consecutive pair with the edge samples (Imu::getMeasurementsContainingEdges, sensor/imu.h:86-135), feature velocities from the next
    frame (DiscreteTimeEstimator::computeFeatureVelocity, estimator.h:293-325), global positions with their left-closest state
    (getLeftClosestStateId, sensor/global_position_sensor.h:82-96).  Every observation of the scene is taken as a valid reprojection."""
    rng = np.random.default_rng(seed)
    F = sc["frame_t_ns"].size
    cam_ts = sc["frame_t_ns"] * 1e-9
    ts_imu = cam_ts + sc["t_offset_cam_imu"]
    q = (Rot.from_quat(sc["gt_frame_q"]) * Rot.from_rotvec(rng.normal(size=(F, 3)) * np.deg2rad(perturb[1]))).as_quat()
    t = sc["gt_frame_p"] + rng.normal(size=(F, 3)) * perturb[0]
    poses = np.ascontiguousarray(np.concatenate([q, t], axis=1))
    # feature velocities
    ob = sc["obs_begin"]
    vel = np.zeros_like(sc["uv"])
    for f in range(F - 1):
        a0, a1, b0, b1 = ob[f], ob[f + 1], ob[f + 1], ob[f + 2]
        if a1 == a0 or b1 == b0:
            continue
        la, lb = sc["lm_index"][a0:a1], sc["lm_index"][b0:b1]
        order = np.argsort(lb, kind="stable")
        pos = np.searchsorted(lb[order], la)
        pos_c = np.minimum(pos, lb.size - 1)
        hit = lb[order][pos_c] == la
        j = b0 + order[pos_c[hit]]
        vel[a0:a1][hit] = (sc["uv"][j] - sc["uv"][a0:a1][hit]) / (cam_ts[f + 1] - cam_ts[f])
    # IMU samples containing the edges of every consecutive pair
    ti = sc["imu_t_ns"] * 1e-9
    imu = []
    for f in range(F - 1):
        t0, t1 = ts_imu[f], ts_imu[f + 1]
        if ti.size == 0 or ti[0] > t0 or ti[-1] < t1:
            continue
        k0 = int(np.searchsorted(ti, t0, side="right")) - 1       # last sample with t <= t0
        k1 = int(np.searchsorted(ti, t1, side="right"))            # first sample with t > t1
        k1 = min(k1, ti.size - 1)
        imu.append((f, ti[k0:k1 + 1].copy(), sc["imu_gyr"][k0:k1 + 1].copy(), sc["imu_acc"][k0:k1 + 1].copy()))
    # global positions
    t_gp = sc["gp_t_ns"] * 1e-9
    t_gp_imu_init = sc["t_offset_cam_imu"] - sc["t_s_gp"]
    f0 = np.searchsorted(ts_imu, t_gp + t_gp_imu_init, side="right") - 1     # first state with ts > t, minus one
    ok = (f0 >= 0) & (f0 < F - 1)
    return dict(cam_model=sc["cam_model"], intr=sc["intr"], dist=sc["dist"], q_bc=sc["q_bc"], p_bc=sc["p_bc"], t_offset_cam_imu=sc["t_offset_cam_imu"],
                width=sc["width"], height=sc["height"], poses=poses, cam_ts=cam_ts, gt_poses=np.concatenate([sc["gt_frame_q"], sc["gt_frame_p"]], axis=1),
                landmarks=sc["landmarks"], scale=sc["scale"], p_bp=sc["p_bp"], t_offset_gs_imu=t_gp_imu_init, g_W=sc["g"],
                state_ids=np.arange(F, dtype=np.int32), obs_begin=ob, uv=sc["uv"], vel=vel, lm_index=sc["lm_index"], sigma_reproj_px=sc["sigma_reproj_px"],
                imu=imu, gyro_noise_density=sc["sigma_omega_c"], acc_noise_density=sc["sigma_acc_c"], gyro_random_walk=gyro_random_walk, acc_random_walk=acc_random_walk,
                gp_t=t_gp[ok], gp_pos=np.ascontiguousarray(sc["gp_pos"][ok]), gp_frame0=f0[ok].astype(np.int32), gp_std=sc["gp_std"],
                flags=(sc["flags"][0], sc["flags"][1], sc["flags"][3], sc["flags"][4]))
