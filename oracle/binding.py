"""ORACLE — TEST INFRASTRUCTURE ONLY.  ctypes binding of oracle/liboracle.so.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  The method names mirror BSplineSO3R3 (/root/reference include/spline/bspline_so3r3.h)
so that parity tests read like calls into the reference.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
i64p = np.ctypeslib.ndpointer(dtype=np.int64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, f) for f in ("api.cpp", "lm.h", "sparse_lm.h", "dt_problem.h", "problem.h", "functors.h", "ref_math.h", "jet.h", "select.h")]
    if force or not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        so = os.path.join(_HERE, "liboracle.so")
        if not os.path.exists(so):
            build()
        L = C.CDLL(so)
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_int64, C.c_int64]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_time_evaluate.restype = C.c_double
        _LIB = L
    return _LIB


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class OracleSpline:
    """CPU restatement of BSplineSO3R3<N> + ceres::Solve (see oracle/problem.h, oracle/lm.h)."""

    def __init__(self, order, dt_ns, start_t_ns=0):
        self.L = lib()
        self.N = order
        self.h = C.c_void_p(self.L.orc_create(order, int(dt_ns), int(start_t_ns)))
        if not self.h:
            raise ValueError("unsupported spline order")

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    def setKnots(self, quat_xyzw, pos):
        q, p = _d(quat_xyzw), _d(pos)
        self.L.orc_set_knots(self.h, len(q), q.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p))

    def numKnots(self):
        return self.L.orc_num_knots(self.h)

    def getKnots(self):
        n = self.numKnots()
        q, p = np.zeros((n, 4)), np.zeros((n, 3))
        self.L.orc_get_knots(self.h, q.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p))
        return q, p

    def initializeBiases(self, dt_s_bias, acc0, gyr0):
        a, g = _d(acc0), _d(gyr0)
        return self.L.orc_initialize_biases(self.h, C.c_double(dt_s_bias), a.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p))

    def numBiasKnots(self):
        return self.L.orc_num_bias_knots(self.h)

    def setBiases(self, acc, gyr):
        a, g = _d(acc), _d(gyr)
        self.L.orc_set_biases(self.h, a.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p))

    def getBiases(self):
        n = self.numBiasKnots()
        a, g = np.zeros((n, 3)), np.zeros((n, 3))
        self.L.orc_get_biases(self.h, a.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p))
        return a, g

    def set3DPoints(self, xyz):
        x = _d(xyz)
        self._nlm = len(x)
        self.L.orc_set_landmarks(self.h, len(x), x.ctypes.data_as(C.c_void_p))

    def get3DPoints(self):
        x = np.zeros((self._nlm, 3))
        self.L.orc_get_landmarks(self.h, x.ctypes.data_as(C.c_void_p))
        return x

    def setCamera(self, model, intr, dist, q_bc, p_bc, t_offset_cam_imu):
        a = [_d(v) for v in (intr, dist, q_bc, p_bc)]
        self.L.orc_set_camera(self.h, int(model), *[v.ctypes.data_as(C.c_void_p) for v in a], C.c_double(t_offset_cam_imu))

    def setGlobals(self, scale, p_bp, g, t_s_gp_s):
        a, b = _d(p_bp), _d(g)
        self.L.orc_set_globals(self.h, C.c_double(scale), a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), C.c_double(t_s_gp_s))

    def setCorrections(self, corr10):
        c = _d(corr10)
        assert c.size == 10
        self.L.orc_set_corrections(self.h, c.ctypes.data_as(C.c_void_p))

    def getGlobals(self):
        o = np.zeros(19)
        self.L.orc_get_globals(self.h, o.ctypes.data_as(C.c_void_p))
        return o

    def setFlags(self, optimize_scale, optimize_pbp, optimize_gravity, optimize_camimu_extrinsics, optimize_t_offset_s_gs):
        self.L.orc_set_flags(self.h, int(optimize_scale), int(optimize_pbp), int(optimize_gravity), int(optimize_camimu_extrinsics), int(optimize_t_offset_s_gs))

    def addFrames(self, t_ns, obs_begin, uv, lm_index, sigma_px):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        ob = np.ascontiguousarray(obs_begin, dtype=np.int32)
        u = _d(uv)
        li = np.ascontiguousarray(lm_index, dtype=np.int32)
        return self.L.orc_add_frames(self.h, len(t), t.ctypes.data_as(C.c_void_p), ob.ctypes.data_as(C.c_void_p), u.ctypes.data_as(C.c_void_p), li.ctypes.data_as(C.c_void_p), C.c_double(sigma_px))

    def addImu(self, t_ns, acc, gyr, acc_noise_density, gyro_noise_density):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        a, g = _d(acc), _d(gyr)
        return self.L.orc_add_imu(self.h, len(t), t.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), C.c_double(acc_noise_density), C.c_double(gyro_noise_density))

    def addGp(self, t_ns, pos, noise):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        p = _d(pos)
        return self.L.orc_add_gp(self.h, len(t), t.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p), C.c_double(noise))

    def addImuBiasCost(self, freq_hz, dt_s, acc_rw, gyr_rw):
        return self.L.orc_add_bias_cost(self.h, C.c_double(freq_hz), C.c_double(dt_s), C.c_double(acc_rw), C.c_double(gyr_rw))

    def addGravityNormError(self):
        self.L.orc_add_gravity_norm(self.h)

    def relinearize(self):
        self.L.orc_relinearize(self.h)

    def relinearizeMask(self, mask):
        """bit 0: relinearizeSplineImuTimeOffset only, bit 1: relinearizeSplineGlobSensTimeOffset only"""
        self.L.orc_relinearize_mask(self.h, int(mask))

    def counts(self):
        c = np.zeros(8, dtype=np.int64)
        self.L.orc_counts(self.h, c.ctypes.data_as(C.c_void_p))
        return dict(zip(["frames", "obs", "accel", "gyro", "gp", "bias", "gravity", "cols"], [int(v) for v in c]))

    def jac_width(self, kind):
        N = self.N
        return [6 * N + 10, 6 * N + 16, 3 * N + 13, 6 * N + 5, 24, 3, 3 * N, 6 * N + 11][kind]   # 7: world alignment, all blocks of the functor (knots first)

    def evaluate(self, kind, jac=True):
        c = self.counts()
        nres = [2 * c["obs"], 3 * c["accel"], 3 * c["gyro"], 3 * c["gp"], 3 * c["bias"], c["gravity"], 3 * self.numPriors(), 3 * self.numAlignmentMeasurements()][kind]
        r = np.zeros(nres)
        J = np.zeros((nres, self.jac_width(kind))) if jac else None
        aux = np.zeros(1)
        self.L.orc_evaluate(self.h, kind, r.ctypes.data_as(C.c_void_p), J.ctypes.data_as(C.c_void_p) if jac else None, aux.ctypes.data_as(C.c_void_p))
        return r, J, float(aux[0])

    def cost(self):
        return self.costTerms()[:7]

    def costTerms(self):
        o = np.zeros(8)
        self.L.orc_cost(self.h, o.ctypes.data_as(C.c_void_p))
        return o

    def addPositionMeasurements(self, t_ns, pos, noise):
        t = np.ascontiguousarray(t_ns, dtype=np.int64); m = np.ascontiguousarray(pos, dtype=np.float64)
        return self.L.orc_add_priors(self.h, 0, len(t), t.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p), C.c_double(noise))

    def addOrientationMeasurements(self, t_ns, quat_xyzw, noise):
        t = np.ascontiguousarray(t_ns, dtype=np.int64); m = np.ascontiguousarray(quat_xyzw, dtype=np.float64)
        return self.L.orc_add_priors(self.h, 1, len(t), t.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p), C.c_double(noise))

    def numPriors(self):
        return self.L.orc_num_priors(self.h)

    # ---- alignment of a frozen spline to the world frame (align_spline_to_global_frame.cpp) ----
    def setWorldColmapTransformation(self, T_wg, corr=None):
        T = _d(T_wg); c = _d(corr) if corr is not None else None
        self.L.orc_set_alignment(self.h, T.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p) if c is not None else None)

    def addWorldColmapAlignmentMeasurements(self, t_ns, pos, noise):
        t = np.ascontiguousarray(t_ns, dtype=np.int64); m = _d(pos)
        return self.L.orc_add_align(self.h, len(t), t.ctypes.data_as(C.c_void_p), m.ctypes.data_as(C.c_void_p), C.c_double(noise))

    def numAlignmentMeasurements(self):
        return self.L.orc_num_align(self.h)

    def getWorldColmapTransformation(self):
        T = np.zeros(7); c = np.zeros(7)
        self.L.orc_get_alignment(self.h, T.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p))
        return T, c

    def optimize(self, max_num_iterations, num_threads=1):
        s = np.zeros(6)
        cap = max_num_iterations + 2
        log = np.zeros((cap, 7))
        self.L.orc_solve(self.h, int(max_num_iterations), int(num_threads), s.ctypes.data_as(C.c_void_p), log.ctypes.data_as(C.c_void_p), cap)
        n = int(s[5])
        return dict(termination_type=int(s[0]), initial_cost=s[1], final_cost=s[2], num_successful_steps=int(s[3]),
                    num_unsuccessful_steps=int(s[4]), iterations=log[:min(n, cap)].copy())

    def optimizeSparse(self, max_num_iterations, num_threads=1, no_termination=False):
        """The same LM as optimize() with the sparse exact solve of sparse_lm.h (landmark Schur complement + bordered band Cholesky on
        host threads): whole BASELINE configs fit.  Adds the phase timings of the run."""
        s = np.zeros(6)
        cap = max_num_iterations + 2
        log = np.zeros((cap, 7))
        tm = np.zeros(12)
        self.L.orc_solve_sparse(self.h, int(max_num_iterations), int(num_threads), int(bool(no_termination)), s.ctypes.data_as(C.c_void_p),
                                log.ctypes.data_as(C.c_void_p), cap, tm.ctypes.data_as(C.c_void_p))
        n = int(s[5])
        keys = ["eval_jac_s", "eval_res_s", "build_s", "factor_s", "total_s", "jac_evals", "res_evals", "solves", "n_band", "half_bandwidth", "n_globals", "n_landmarks"]
        return dict(termination_type=int(s[0]), initial_cost=s[1], final_cost=s[2], num_successful_steps=int(s[3]),
                    num_unsuccessful_steps=int(s[4]), iterations=log[:min(n, cap)].copy(), timings=dict(zip(keys, [float(v) for v in tm])))

    def computeValidReprojections(self, t_ns, pt_begin, uv, lm_index, width, height, cell_size_px):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        pb = np.ascontiguousarray(pt_begin, dtype=np.int32)
        u = _d(uv)
        li = np.ascontiguousarray(lm_index, dtype=np.int32)
        valid = np.zeros(max(li.size, 1), dtype=np.uint8)
        nv = np.zeros(max(t.size, 1), dtype=np.int32)
        lib().orc_compute_valid_reprojections(self.h, int(t.size), t.ctypes.data_as(C.c_void_p), pb.ctypes.data_as(C.c_void_p),
                                              u.ctypes.data_as(C.c_void_p), li.ctypes.data_as(C.c_void_p), int(width), int(height),
                                              int(cell_size_px), valid.ctypes.data_as(C.c_void_p), nv.ctypes.data_as(C.c_void_p))
        return valid[:li.size].astype(bool), nv[:t.size]

    def timeEvaluate(self, num_threads, reps=1, frame_limit=0):
        obs = C.c_int64(0)
        t = self.L.orc_time_evaluate(self.h, int(num_threads), int(reps), int(frame_limit), C.byref(obs))
        return float(t), int(obs.value)

    def sampleAt(self, what, t_ns):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        out = np.zeros((t.size, 7 if what == 0 else 3))
        rc = self.L.orc_sample_at(self.h, int(what), C.c_int64(t.size), t.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p))
        if rc:
            raise ValueError("sample time outside the spline")
        return out

    def sample(self, s, u):
        o = np.zeros(19)
        self.L.orc_sample(self.h, C.c_int64(int(s)), C.c_double(u), o.ctypes.data_as(C.c_void_p))
        return o


class OracleDiscrete:
    """CPU restatement of DiscreteTimeEstimator + its three cost terms (oracle/dt_problem.h).  Method names follow
    include/discrete_time_estimator/estimator.h; poses are Sophus SE3 blocks [qx qy qz qw tx ty tz]."""

    def __init__(self):
        self.L = lib()
        self.L.orc_dt_create.restype = C.c_void_p
        self.h = C.c_void_p(self.L.orc_dt_create())
        self.F = 0
        self._nlm = 0

    def __del__(self):
        try:
            self.L.orc_dt_destroy(self.h)
        except Exception:
            pass

    def setCamera(self, model, intr, dist, q_bc, p_bc, t_offset_cam_imu, width, height):
        a = [_d(v) for v in (intr, dist, q_bc, p_bc)]
        self.L.orc_dt_set_camera(self.h, int(model), *[v.ctypes.data_as(C.c_void_p) for v in a], C.c_double(t_offset_cam_imu), C.c_double(width), C.c_double(height))

    def setImu(self, gyro_noise_density, acc_noise_density, gyro_random_walk, acc_random_walk, g_W, g_max=7.8, a_max=150.0):
        g = _d(g_W)
        self.L.orc_dt_set_imu_params(self.h, C.c_double(gyro_noise_density), C.c_double(acc_noise_density), C.c_double(gyro_random_walk), C.c_double(acc_random_walk),
                                     C.c_double(g_max), C.c_double(a_max), g.ctypes.data_as(C.c_void_p))

    def initializeStates(self, poses, cam_ts):
        P, t = _d(poses), _d(cam_ts)
        self.F = len(t)
        self.L.orc_dt_set_states(self.h, self.F, P.ctypes.data_as(C.c_void_p), t.ctypes.data_as(C.c_void_p))

    def initializeLinearVelocity(self):
        self.L.orc_dt_init_velocity(self.h)

    def initializeStatesKeep(self, poses, sb):
        """overwrite poses and speed-and-biases of the existing states (measurements and IMU caches stay)"""
        P, x = _d(poses), _d(sb)
        assert P.shape == (self.F, 7) and x.shape == (self.F, 9)
        self.L.orc_dt_overwrite_states(self.h, P.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p))

    def setPbp(self, p_bp):
        a = _d(p_bp)
        self.L.orc_dt_set_pbp(self.h, a.ctypes.data_as(C.c_void_p))

    def setSpeedAndBias(self, sb):
        x = _d(sb)
        assert x.shape == (self.F, 9)
        self.L.orc_dt_set_speed_bias(self.h, x.ctypes.data_as(C.c_void_p))

    def getStates(self):
        P, sb = np.zeros((self.F, 7)), np.zeros((self.F, 9))
        self.L.orc_dt_get_states(self.h, P.ctypes.data_as(C.c_void_p), sb.ctypes.data_as(C.c_void_p))
        return P, sb

    def set3DPoints(self, xyz):
        x = _d(xyz); self._nlm = len(x)
        self.L.orc_dt_set_landmarks(self.h, len(x), x.ctypes.data_as(C.c_void_p))

    def get3DPoints(self):
        x = np.zeros((self._nlm, 3))
        self.L.orc_dt_get_landmarks(self.h, x.ctypes.data_as(C.c_void_p))
        return x

    def setGlobals(self, scale, p_bp, t_offset_globsens_imu):
        a = _d(p_bp)
        self.L.orc_dt_set_globals(self.h, C.c_double(scale), a.ctypes.data_as(C.c_void_p), C.c_double(t_offset_globsens_imu))

    def setCorrections(self, corr9, t_gs_corr=0.0):
        c = _d(corr9); assert c.size == 9
        self.L.orc_dt_set_corrections(self.h, c.ctypes.data_as(C.c_void_p), C.c_double(t_gs_corr))

    def getGlobals(self):
        o = np.zeros(13)
        self.L.orc_dt_get_globals(self.h, o.ctypes.data_as(C.c_void_p))
        return o

    def setFlags(self, optimize_scale, optimize_pbp, optimize_camimu_extrinsics, optimize_t_offset_globsens_imu):
        self.L.orc_dt_set_flags(self.h, int(optimize_scale), int(optimize_pbp), int(optimize_camimu_extrinsics), int(optimize_t_offset_globsens_imu))

    def addFrames(self, state_ids, obs_begin, uv, vel, lm_index, sigma_px):
        sid = np.ascontiguousarray(state_ids, dtype=np.int32); ob = np.ascontiguousarray(obs_begin, dtype=np.int32)
        u, v = _d(uv), _d(vel); li = np.ascontiguousarray(lm_index, dtype=np.int32)
        return self.L.orc_dt_add_frames(self.h, len(sid), sid.ctypes.data_as(C.c_void_p), ob.ctypes.data_as(C.c_void_p), u.ctypes.data_as(C.c_void_p),
                                        v.ctypes.data_as(C.c_void_p), li.ctypes.data_as(C.c_void_p), C.c_double(sigma_px))

    def addImuMeasurements(self, frame0, t, gyr, acc):
        t, g, a = _d(t), _d(gyr), _d(acc)
        self.L.orc_dt_add_imu_factor(self.h, int(frame0), len(t), t.ctypes.data_as(C.c_void_p), g.ctypes.data_as(C.c_void_p), a.ctypes.data_as(C.c_void_p))

    def addGlobalSensMeasurements(self, t, pos, frame0, sigma):
        t, x = _d(t), _d(pos); f0 = np.ascontiguousarray(frame0, dtype=np.int32)
        return self.L.orc_dt_add_gp(self.h, len(t), t.ctypes.data_as(C.c_void_p), x.ctypes.data_as(C.c_void_p), f0.ctypes.data_as(C.c_void_p), C.c_double(sigma))

    def counts(self):
        c = np.zeros(5, dtype=np.int64)
        self.L.orc_dt_counts(self.h, c.ctypes.data_as(C.c_void_p))
        return dict(zip(["frames", "obs", "imu", "gp", "cols"], [int(v) for v in c]))

    def evaluate(self, kind, jac=True):
        c = self.counts()
        nres, w = [(2 * c["obs"], 17), (15 * c["imu"], 30), (3 * c["gp"], 17)][kind]
        r = np.zeros(nres); J = np.zeros((nres, w)) if jac else None
        self.L.orc_dt_evaluate(self.h, int(kind), r.ctypes.data_as(C.c_void_p), J.ctypes.data_as(C.c_void_p) if jac else None)
        return r, J

    def cost(self):
        o = np.zeros(4)
        self.L.orc_dt_cost(self.h, o.ctypes.data_as(C.c_void_p))
        return o

    def optimize(self, max_num_iterations, num_threads=1):
        s = np.zeros(6); cap = max_num_iterations + 2; log = np.zeros((cap, 7))
        self.L.orc_dt_solve(self.h, int(max_num_iterations), int(num_threads), s.ctypes.data_as(C.c_void_p), log.ctypes.data_as(C.c_void_p), cap)
        n = int(s[5])
        return dict(termination_type=int(s[0]), initial_cost=s[1], final_cost=s[2], num_successful_steps=int(s[3]), num_unsuccessful_steps=int(s[4]),
                    iterations=log[:min(n, cap)].copy())

    def computeValidReprojections(self, pt_begin, uv, lm_index, width, height, cell_size_px, cam_ts=None):
        """computeValidReprojections (+ computeFeatureVelocity when cam_ts is given: returns (valid, n_valid, velocity))"""
        pb = np.ascontiguousarray(pt_begin, dtype=np.int32); u = _d(uv); li = np.ascontiguousarray(lm_index, dtype=np.int32)
        F = pb.size - 1
        valid = np.zeros(max(li.size, 1), dtype=np.uint8); nv = np.zeros(max(F, 1), dtype=np.int32)
        ts = _d(cam_ts) if cam_ts is not None else np.zeros(max(F, 1))
        vel = np.zeros((max(li.size, 1), 2)) if cam_ts is not None else None
        self.L.orc_dt_compute_valid_reprojections(self.h, int(F), ts.ctypes.data_as(C.c_void_p), pb.ctypes.data_as(C.c_void_p), u.ctypes.data_as(C.c_void_p),
                                                  li.ctypes.data_as(C.c_void_p), int(width), int(height), int(cell_size_px), valid.ctypes.data_as(C.c_void_p),
                                                  nv.ctypes.data_as(C.c_void_p), vel.ctypes.data_as(C.c_void_p) if vel is not None else None)
        if vel is None:
            return valid[:li.size].astype(bool), nv[:F]
        return valid[:li.size].astype(bool), nv[:F], vel[:li.size]

    def optimizeSparse(self, max_num_iterations, num_threads=1, no_termination=False):
        """sparse LM (oracle/sparse_lm.h) on the discrete-time problem: what bench.py's CPU legs time at BASELINE sizes"""
        s = np.zeros(6); cap = max_num_iterations + 2; log = np.zeros((cap, 7)); tm = np.zeros(12)
        self.L.orc_dt_solve_sparse(self.h, int(max_num_iterations), int(num_threads), int(bool(no_termination)), s.ctypes.data_as(C.c_void_p), log.ctypes.data_as(C.c_void_p), cap,
                                   tm.ctypes.data_as(C.c_void_p))
        n = int(s[5])
        keys = ["eval_jac_s", "eval_res_s", "build_s", "factor_s", "total_s", "jac_evals", "res_evals", "solves", "n_band", "half_bandwidth", "n_globals", "n_landmarks"]
        return dict(termination_type=int(s[0]), initial_cost=s[1], final_cost=s[2], num_successful_steps=int(s[3]), num_unsuccessful_steps=int(s[4]),
                    iterations=log[:min(n, cap)].copy(), timings=dict(zip(keys, [float(v) for v in tm])))

    def imuMinimal(self, k):
        r = np.zeros(15); J = np.zeros((15, 30))
        self.L.orc_dt_imu_minimal(self.h, int(k), r.ctypes.data_as(C.c_void_p), J.ctypes.data_as(C.c_void_p))
        return r, J

    def imuRedoCount(self, k):
        return self.L.orc_dt_imu_redo_count(self.h, int(k))


def se3_plus(T, d):
    o = np.zeros(7); T = _d(T); d = _d(d)
    lib().orc_se3_plus(T.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p)); return o


def se3_plus_jacobian(T):
    J = np.zeros((7, 6)); T = _d(T)
    lib().orc_se3_plus_jacobian(T.ctypes.data_as(C.c_void_p), J.ctypes.data_as(C.c_void_p)); return J


# ---- primitives for known-answer tests ----
def so3_exp(w):
    q = np.zeros(4); w = _d(w)
    lib().orc_so3_exp(w.ctypes.data_as(C.c_void_p), q.ctypes.data_as(C.c_void_p)); return q


def so3_log(q):
    w = np.zeros(3); q = _d(q)
    lib().orc_so3_log(q.ctypes.data_as(C.c_void_p), w.ctypes.data_as(C.c_void_p)); return w


def so3_mul(a, b):
    c = np.zeros(4); a = _d(a); b = _d(b)
    lib().orc_so3_mul(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), c.ctypes.data_as(C.c_void_p)); return c


def so3_matrix(q):
    R = np.zeros((3, 3)); q = _d(q)
    lib().orc_so3_matrix(q.ctypes.data_as(C.c_void_p), R.ctypes.data_as(C.c_void_p)); return R


def so3_act(q, p):
    o = np.zeros(3); q = _d(q); p = _d(p)
    lib().orc_so3_act(q.ctypes.data_as(C.c_void_p), p.ctypes.data_as(C.c_void_p), o.ctypes.data_as(C.c_void_p)); return o


def so3_plus_jacobian(q):
    J = np.zeros((4, 3)); q = _d(q)
    lib().orc_so3_plus_jacobian(q.ctypes.data_as(C.c_void_p), J.ctypes.data_as(C.c_void_p)); return J


def spline_table(N, which):
    o = np.zeros((N, N))
    lib().orc_spline_table(int(N), int(which), o.ctypes.data_as(C.c_void_p)); return o
