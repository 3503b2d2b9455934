"""rpg_vision-based_slam_b200 — host-side mirror of the reference's BSplineSO3R3 interface over the
CUDA C-ABI library (include/ctba.h, csrc/libctba.so).

`BSplineSO3R3` keeps the method names and argument meaning of
/root/reference/include/spline/bspline_so3r3.h so that tests read like calls into the reference;
every method forwards to one C-ABI entry point.  There is NO CPU fallback: importing works without
a GPU (so the symbol table can be checked), but constructing a problem raises if the CUDA library
or a CUDA device is missing.

The directory name contains a hyphen, so import it with
    importlib.import_module("rpg_vision-based_slam_b200")
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CTBA_LIB_PATH") or os.path.join(_HERE, "csrc", "libctba.so")   # the override loads instrumented debug builds (profiles/tools)
_LIB = None

CTBA_REPROJ, CTBA_ACCEL, CTBA_GYRO, CTBA_GP, CTBA_BIAS, CTBA_GRAVITY, CTBA_PRIOR = range(7)
(CTBA_SAMPLE_POSE, CTBA_SAMPLE_ACCEL_BODY, CTBA_SAMPLE_ACCEL_WORLD, CTBA_SAMPLE_GYRO_BODY, CTBA_SAMPLE_GYRO_WORLD, CTBA_SAMPLE_ACCEL_BIAS,
 CTBA_SAMPLE_GYRO_BIAS) = range(7)

EXPORTS = [
    "ctba_create", "ctba_destroy", "ctba_last_error", "ctba_set_stream", "ctba_set_knots", "ctba_num_knots", "ctba_get_knots",
    "ctba_initialize_biases", "ctba_num_bias_knots", "ctba_set_biases", "ctba_get_biases", "ctba_set_landmarks", "ctba_get_landmarks",
    "ctba_set_camera", "ctba_set_globals", "ctba_set_corrections", "ctba_get_globals", "ctba_set_flags", "ctba_add_frames", "ctba_add_imu",
    "ctba_add_gp", "ctba_add_position_measurements", "ctba_add_orientation_measurements", "ctba_num_priors", "ctba_cost_terms", "ctba_add_bias_cost", "ctba_add_gravity_norm", "ctba_counts", "ctba_evaluate", "ctba_cost", "ctba_solve",
    "ctba_relinearize_time_offsets", "ctba_bench_evaluate", "ctba_set_option", "ctba_get_stat", "ctba_invalidate_device", "ctba_transfer_bytes",
    "ctba_bench_reproj", "ctba_compute_valid_reprojections", "ctba_evaluate_errors", "ctba_sample", "ctba_comm_unique_id", "ctba_comm_init_nccl", "ctba_comm_init_local", "ctba_shard_plan",
    "ctba_set_world_alignment", "ctba_get_world_alignment", "ctba_add_alignment_measurements", "ctba_num_alignment_measurements",
    "ctba_align_system_noalloc",   # allocation-free entry point (no reference counterpart; tests/test_zz_no_alloc.py binds it by hand)
]


EXPORTS_DT = [
    "ctba_dt_create", "ctba_dt_destroy", "ctba_dt_last_error", "ctba_dt_set_stream", "ctba_dt_invalidate_device", "ctba_dt_set_camera", "ctba_dt_set_imu", "ctba_dt_initialize_states",
    "ctba_dt_initialize_linear_velocity", "ctba_dt_set_speed_and_bias", "ctba_dt_set_states", "ctba_dt_get_states", "ctba_dt_set_landmarks",
    "ctba_dt_get_landmarks", "ctba_dt_set_globals", "ctba_dt_set_corrections", "ctba_dt_get_globals", "ctba_dt_set_flags", "ctba_dt_add_frames",
    "ctba_dt_add_imu", "ctba_dt_add_gp", "ctba_dt_compute_valid_reprojections", "ctba_dt_counts", "ctba_dt_evaluate", "ctba_dt_cost", "ctba_dt_evaluate_errors", "ctba_dt_imu_redo_count",
    "ctba_dt_solve", "ctba_dt_set_option", "ctba_dt_get_stat",
]


class CtbaError(RuntimeError):
    pass


class _Iteration(C.Structure):
    _fields_ = [("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double), ("step_norm", C.c_double),
                ("tr_ratio", C.c_double), ("tr_radius", C.c_double), ("successful", C.c_int)]


class _Summary(C.Structure):
    _fields_ = [("termination_type", C.c_int), ("initial_cost", C.c_double), ("final_cost", C.c_double),
                ("num_successful_steps", C.c_int), ("num_unsuccessful_steps", C.c_int), ("num_iterations", C.c_int),
                ("iterations", C.POINTER(_Iteration)), ("iterations_capacity", C.c_int),
                ("time_total_s", C.c_double), ("time_eval_s", C.c_double), ("time_build_s", C.c_double), ("time_solve_s", C.c_double),
                ("kernel_launches", C.c_int64)]


def lib():
    """Load csrc/libctba.so (built by __graft_entry__.build()).  Fails loudly when it is missing."""
    global _LIB
    if _LIB is None:
        if not os.path.exists(LIB_PATH):
            raise CtbaError(f"CUDA extension missing: {LIB_PATH} (run __graft_entry__.build()); there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.ctba_last_error.restype = C.c_char_p
        L.ctba_last_error.argtypes = [C.c_void_p]
        _LIB = L
    return _LIB


def _d(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class BSplineSO3R3:
    """GPU-resident continuous-time problem: state + measurements + LM solve (ctba.h)."""

    def __init__(self, order, dt_ns, start_t_ns=0, device=0):
        self.L = lib()
        self.N = int(order)
        self.h = C.c_void_p()
        rc = self.L.ctba_create(self.N, C.c_int64(int(dt_ns)), C.c_int64(int(start_t_ns)), int(device), C.byref(self.h))
        if rc != 0:
            self.h = None
            raise CtbaError(f"ctba_create failed with code {rc} (no CUDA device or bad arguments); there is no CPU fallback")
        self._nlm = 0
        self.dt_ns, self.start_t_ns, self.dt_ns_bias = int(dt_ns), int(start_t_ns), None

    def close(self):
        if getattr(self, "h", None):
            self.L.ctba_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc < 0:
            msg = self.L.ctba_last_error(self.h)
            raise CtbaError(f"ctba error {rc}: {msg.decode() if msg else ''}")
        return rc

    def setStream(self, cuda_stream_ptr):
        self._chk(self.L.ctba_set_stream(self.h, C.c_void_p(cuda_stream_ptr)))

    # ---- state ----
    def setKnots(self, quat_xyzw, pos):
        q, p = _d(quat_xyzw), _d(pos)
        assert q.shape[0] == p.shape[0]
        self._chk(self.L.ctba_set_knots(self.h, int(q.shape[0]), _p(q), _p(p)))

    def numKnots(self):
        return self.L.ctba_num_knots(self.h)

    def getKnots(self):
        n = self.numKnots()
        q, p = np.zeros((n, 4)), np.zeros((n, 3))
        self._chk(self.L.ctba_get_knots(self.h, _p(q), _p(p)))
        return q, p

    def initializeBiases(self, dt_s_bias, acc0, gyr0):
        a, g = _d(acc0), _d(gyr0)
        return self._chk(self.L.ctba_initialize_biases(self.h, C.c_double(dt_s_bias), _p(a), _p(g)))

    def numBiasKnots(self):
        return self.L.ctba_num_bias_knots(self.h)

    def setBiases(self, acc, gyr):
        a, g = _d(acc), _d(gyr)
        self._chk(self.L.ctba_set_biases(self.h, _p(a), _p(g)))

    def getBiases(self):
        n = self.numBiasKnots()
        a, g = np.zeros((n, 3)), np.zeros((n, 3))
        self._chk(self.L.ctba_get_biases(self.h, _p(a), _p(g)))
        return a, g

    def set3DPoints(self, xyz):
        x = _d(xyz)
        self._nlm = int(x.shape[0])
        self._chk(self.L.ctba_set_landmarks(self.h, self._nlm, _p(x)))

    def get3DPoints(self):
        x = np.zeros((self._nlm, 3))
        self._chk(self.L.ctba_get_landmarks(self.h, _p(x)))
        return x

    def setCamera(self, model, intr, dist, q_bc, p_bc, t_offset_cam_imu):
        a = [_d(v) for v in (intr, dist, q_bc, p_bc)]
        self._chk(self.L.ctba_set_camera(self.h, int(model), *[_p(v) for v in a], C.c_double(t_offset_cam_imu)))

    def setGlobals(self, scale, p_bp, g, t_s_gp_s):
        a, b = _d(p_bp), _d(g)
        self._chk(self.L.ctba_set_globals(self.h, C.c_double(scale), _p(a), _p(b), C.c_double(t_s_gp_s)))

    def setCorrections(self, corr10):
        c = _d(corr10)
        assert c.size == 10
        self._chk(self.L.ctba_set_corrections(self.h, _p(c)))

    def getGlobals(self):
        o = np.zeros(19)
        self._chk(self.L.ctba_get_globals(self.h, _p(o)))
        return o

    def setFlags(self, optimize_scale, optimize_pbp, optimize_gravity, optimize_camimu_extrinsics, optimize_t_offset_s_gs):
        self._chk(self.L.ctba_set_flags(self.h, int(optimize_scale), int(optimize_pbp), int(optimize_gravity),
                                        int(optimize_camimu_extrinsics), int(optimize_t_offset_s_gs)))

    # ---- measurements ----
    def addFrames(self, t_ns, obs_begin, uv, lm_index, sigma_px):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        ob = np.ascontiguousarray(obs_begin, dtype=np.int32)
        u = _d(uv)
        li = np.ascontiguousarray(lm_index, dtype=np.int32)
        assert ob.size == t.size + 1
        return self._chk(self.L.ctba_add_frames(self.h, int(t.size), _p(t), _p(ob), _p(u), _p(li), C.c_double(sigma_px)))

    def addImu(self, t_ns, acc, gyr, acc_noise_density, gyro_noise_density):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        a, g = _d(acc), _d(gyr)
        return self._chk(self.L.ctba_add_imu(self.h, int(t.size), _p(t), _p(a), _p(g), C.c_double(acc_noise_density), C.c_double(gyro_noise_density)))

    def addGp(self, t_ns, pos, noise):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        x = _d(pos)
        return self._chk(self.L.ctba_add_gp(self.h, int(t.size), _p(t), _p(x), C.c_double(noise)))

    def addPositionMeasurements(self, t_ns, pos, noise):
        """addPositionMeasurement per sample (bspline_so3r3.h:1048-1085)"""
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        x = _d(pos)
        return self._chk(self.L.ctba_add_position_measurements(self.h, int(t.size), _p(t), _p(x), C.c_double(noise)))

    def addOrientationMeasurements(self, t_ns, quat_xyzw, noise):
        """addOrientationMeasurement per sample (bspline_so3r3.h:1087-1123)"""
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        x = _d(quat_xyzw)
        return self._chk(self.L.ctba_add_orientation_measurements(self.h, int(t.size), _p(t), _p(x), C.c_double(noise)))

    def numPriors(self):
        return self.L.ctba_num_priors(self.h)

    # ---- alignment of a frozen spline to the world frame (src/align_spline_to_global_frame.cpp:153-226) ----
    def setWorldColmapTransformation(self, T_wg, corr=None):
        """setWorldColmapTransformation (bspline_so3r3.h:1825); T_wg = qx qy qz qw tx ty tz"""
        T = _d(T_wg); c = _d(corr) if corr is not None else None
        self._chk(self.L.ctba_set_world_alignment(self.h, _p(T), _p(c) if c is not None else None))

    def addWorldColmapAlignmentMeasurements(self, t_ns, pos, noise):
        """addWorldColmapAlignmentMeasurement per sample (bspline_so3r3.h:1121-1211)"""
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        x = _d(pos)
        return self._chk(self.L.ctba_add_alignment_measurements(self.h, int(t.size), _p(t), _p(x), C.c_double(noise)))

    def numAlignmentMeasurements(self):
        return self.L.ctba_num_alignment_measurements(self.h)

    def getWorldColmapTransformation(self):
        """(getWorldColmapTransformation, getWorldColmapTransformationCorr)  bspline_so3r3.h:1827-1840"""
        T = np.zeros(7); c = np.zeros(7)
        self._chk(self.L.ctba_get_world_alignment(self.h, _p(T), _p(c)))
        return T, c

    def evaluateWorldColmapAlignmentErrors(self):
        """bspline_so3r3.h:1682-1790: |residual| of every alignment measurement with unit weight"""
        return self.evaluateErrors(7)

    def addImuBiasCost(self, freq_hz, dt_s, acc_rw, gyr_rw):
        return self._chk(self.L.ctba_add_bias_cost(self.h, C.c_double(freq_hz), C.c_double(dt_s), C.c_double(acc_rw), C.c_double(gyr_rw)))

    def addGravityNormError(self):
        self._chk(self.L.ctba_add_gravity_norm(self.h))

    def counts(self):
        c = np.zeros(8, dtype=np.int64)
        self._chk(self.L.ctba_counts(self.h, _p(c)))
        return dict(zip(["frames", "obs", "accel", "gyro", "gp", "bias", "gravity", "cols"], [int(v) for v in c]))

    # ---- evaluation / solve ----
    def jac_width(self, kind):
        N = self.N
        return [6 * N + 10, 6 * N + 16, 3 * N + 13, 6 * N + 5, 24, 3, 3 * N, 11][kind]

    def evaluate(self, kind, jac=True):
        c = self.counts()
        nres = [2 * c["obs"], 3 * c["accel"], 3 * c["gyro"], 3 * c["gp"], 3 * c["bias"], c["gravity"], 3 * self.numPriors(),
                3 * self.numAlignmentMeasurements()][kind]
        r = np.zeros(nres)
        J = np.zeros((nres, self.jac_width(kind))) if jac else None
        n = self._chk(self.L.ctba_evaluate(self.h, int(kind), _p(r), _p(J) if jac else None))
        assert n == nres
        return r, J

    def cost(self):
        o = np.zeros(7)
        self._chk(self.L.ctba_cost(self.h, _p(o)))
        return o

    def costTerms(self):
        """cost() plus [7] = the pose priors (position / orientation measurements)"""
        o = np.zeros(8)
        self._chk(self.L.ctba_cost_terms(self.h, _p(o)))
        return o

    def optimize(self, max_num_iterations):
        cap = max_num_iterations + 2
        its = (_Iteration * cap)()
        s = _Summary()
        s.iterations = C.cast(its, C.POINTER(_Iteration))
        s.iterations_capacity = cap
        self._chk(self.L.ctba_solve(self.h, int(max_num_iterations), C.byref(s)))
        n = min(s.num_iterations, cap)
        log = np.array([[its[i].cost, its[i].cost_change, its[i].gradient_max_norm, its[i].step_norm, its[i].tr_ratio, its[i].tr_radius,
                         its[i].successful] for i in range(n)]).reshape(n, 7)
        return dict(termination_type=s.termination_type, initial_cost=s.initial_cost, final_cost=s.final_cost,
                    num_successful_steps=s.num_successful_steps, num_unsuccessful_steps=s.num_unsuccessful_steps, iterations=log,
                    time_total_s=s.time_total_s, time_eval_s=s.time_eval_s, time_build_s=s.time_build_s, time_solve_s=s.time_solve_s,
                    kernel_launches=int(s.kernel_launches))

    def relinearize(self):
        self._chk(self.L.ctba_relinearize_time_offsets(self.h))

    def setOption(self, name, value):
        self._chk(self.L.ctba_set_option(self.h, name.encode(), C.c_double(value)))

    def getStat(self, name):
        v = C.c_double(0.0)
        self._chk(self.L.ctba_get_stat(self.h, name.encode(), C.byref(v)))
        return float(v.value)

    def invalidateDevice(self):
        self._chk(self.L.ctba_invalidate_device(self.h))

    def transferBytes(self):
        b = np.zeros(2, dtype=np.int64)
        self._chk(self.L.ctba_transfer_bytes(self.h, _p(b)))
        return int(b[0]), int(b[1])

    def computeValidReprojections(self, t_ns, pt_begin, uv, lm_index, width, height, cell_size_px):
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        pb = np.ascontiguousarray(pt_begin, dtype=np.int32)
        u = _d(uv)
        li = np.ascontiguousarray(lm_index, dtype=np.int32)
        valid = np.zeros(max(li.size, 1), dtype=np.uint8)
        nv = np.zeros(max(t.size, 1), dtype=np.int32)
        self._chk(self.L.ctba_compute_valid_reprojections(self.h, int(t.size), _p(t), _p(pb), _p(u), _p(li), int(width), int(height),
                                                          int(cell_size_px), _p(valid), _p(nv)))
        return valid[:li.size].astype(bool), nv[:t.size]

    def evaluateErrors(self, kind):
        n = self._chk(self.L.ctba_evaluate_errors(self.h, int(kind), None))
        e = np.zeros(max(n, 1))
        self._chk(self.L.ctba_evaluate_errors(self.h, int(kind), _p(e)))
        return e[:n]

    def sampleAt(self, what, t_ns):
        """getPose / getAccel / getGyro / get*Bias at many timestamps (bspline_so3r3.h:122-327; traceSpline, save.h:99-153)"""
        t = np.ascontiguousarray(t_ns, dtype=np.int64)
        out = np.zeros((t.size, 7 if what == CTBA_SAMPLE_POSE else 3))
        self._chk(self.L.ctba_sample(self.h, int(what), C.c_int64(t.size), _p(t), _p(out)))
        return out

    def getEndTns(self):
        """bspline_so3r3.h:1862-1867 (1 ms before the end of the last window)"""
        return int(float((self.numKnots() - self.N + 1) * self.dt_ns + self.start_t_ns) - 1e6)   # `t_end_ns - 1e6` goes through double

    def traceSpline(self, dt_ns, t_end_ns=None):
        """traceSpline (save.h:99-121): rows 'timestamp_s tx ty tz qx qy qz qw' from getStartTns() to getEndTns() (exclusive)"""
        t = np.arange(self.start_t_ns, int(self.getEndTns() if t_end_ns is None else t_end_ns), int(dt_ns), dtype=np.int64)
        return np.concatenate([(t * 1e-9)[:, None], self.sampleAt(CTBA_SAMPLE_POSE, t)], axis=1)

    def benchReproj(self, reps, flush_l2=True):
        ms = C.c_double(0.0)
        self._chk(self.L.ctba_bench_reproj(self.h, int(reps), int(bool(flush_l2)), C.byref(ms)))
        return float(ms.value)

    # ---- multi-GPU: one problem sharded by time segment (ctba.h, "multi-GPU") ----
    @staticmethod
    def commUniqueId():
        buf = C.create_string_buffer(128)
        rc = lib().ctba_comm_unique_id(buf)
        if rc != 0:
            raise CtbaError(f"ctba_comm_unique_id failed with code {rc} (libnccl.so.2 missing?)")
        return bytes(buf.raw)

    def initCommNccl(self, world, rank, unique_id):
        assert len(unique_id) == 128
        self._chk(self.L.ctba_comm_init_nccl(self.h, int(world), int(rank), C.c_char_p(unique_id)))

    @staticmethod
    def initCommLocal(problems):
        arr = (C.c_void_p * len(problems))(*[p.h for p in problems])
        problems[0]._chk(problems[0].L.ctba_comm_init_local(arr, len(problems)))

    def shardPlan(self, world, rank):
        c = self.counts()
        plan = np.zeros(80, dtype=np.int64)
        lm = np.zeros(max(self._nlm, 1), dtype=np.int32)
        imu = np.zeros(max(c["accel"], 1), dtype=np.int32)
        gp = np.zeros(max(c["gp"], 1), dtype=np.int32)
        bias = np.zeros(max(c["bias"], 1), dtype=np.int32)
        self._chk(self.L.ctba_shard_plan(self.h, int(world), int(rank), _p(plan), _p(lm), _p(imu), _p(gp), _p(bias)))
        keys = ["P", "nbc", "nbw", "q_lo", "q_hi", "bc_lo", "bc_hi", "own_lo", "own_hi", "F", "f_first", "f_cost_begin", "nodes", "hbn", "T", "KD"]
        d = {k: int(v) for k, v in zip(keys, plan[:16])}
        d["c0"] = [int(plan[16 + 2 * q]) for q in range(d["P"])]
        d["nI"] = [int(plan[17 + 2 * q]) for q in range(d["P"])]
        d.update(lm_owner=lm[:self._nlm], imu_owner=imu[:c["accel"]], gp_owner=gp[:c["gp"]], bias_owner=bias[:c["bias"]])
        return d

    def benchEvaluate(self, reps):
        n = C.c_int64(0)
        self._chk(self.L.ctba_bench_evaluate(self.h, int(reps), C.byref(n)))
        return int(n.value)


def load_scene(prob, sc, with_imu=True, with_gp=True, with_frames=True, with_bias_cost=True):
    """Feed a synthetic scene (scene.make_scene) into any object with the BSplineSO3R3 method set
    (the CUDA problem or the CPU oracle), in the reference driver's order
    (src/optimize_continuous_time.cpp:478-601)."""
    prob.setCamera(sc["cam_model"], sc["intr"], sc["dist"], sc["q_bc"], sc["p_bc"], sc["t_offset_cam_imu"])
    prob.setKnots(sc["knot_q"], sc["knot_p"])
    prob.setGlobals(sc["scale"], sc["p_bp"], sc["g"], sc["t_s_gp"])
    prob.setFlags(*sc["flags"])
    prob.set3DPoints(sc["landmarks"])
    if sc.get("gravity_norm", False):
        prob.addGravityNormError()
    n = {}
    if with_imu and sc["imu_t_ns"].size:
        prob.initializeBiases(sc["dt_s_bias"], sc["init_acc_bias"], sc["init_gyro_bias"])
    if with_gp and sc["gp_t_ns"].size:
        n["gp"] = prob.addGp(sc["gp_t_ns"], sc["gp_pos"], sc["gp_std"])
    if with_imu and sc["imu_t_ns"].size:
        n["imu"] = prob.addImu(sc["imu_t_ns"], sc["imu_acc"], sc["imu_gyr"], sc["sigma_acc_c"], sc["sigma_omega_c"])
    if with_frames and sc["frame_t_ns"].size:
        n["frames"] = prob.addFrames(sc["frame_t_ns"], sc["obs_begin"], sc["uv"], sc["lm_index"], sc["sigma_reproj_px"])
    if with_imu and with_bias_cost and sc["imu_t_ns"].size:
        n["bias"] = prob.addImuBiasCost(sc["bias_err_freq_hz"], sc["bias_err_dt_s"], sc["sigma_acc_bias_c"], sc["sigma_omega_bias_c"])
    return n


class DiscreteTimeEstimator:
    """GPU-resident discrete-time problem (include/ctba_dt.h): per-frame SE3 + [v, b_g, b_a] states, reprojection / pre-integrated IMU /
    interpolated global-position terms, LM solve.  Method names follow include/discrete_time_estimator/estimator.h; poses are Sophus SE3
    blocks [qx qy qz qw tx ty tz]."""

    def __init__(self, device=0):
        self.L = lib()
        self.L.ctba_dt_last_error.restype = C.c_char_p
        self.L.ctba_dt_last_error.argtypes = [C.c_void_p]
        self.h = C.c_void_p()
        rc = self.L.ctba_dt_create(int(device), C.byref(self.h))
        if rc != 0:
            self.h = None
            raise CtbaError(f"ctba_dt_create failed with code {rc} (no CUDA device); there is no CPU fallback")
        self.F = 0
        self._nlm = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.ctba_dt_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc):
        if rc < 0:
            raise CtbaError(f"ctba_dt error {rc}: {self.L.ctba_dt_last_error(self.h).decode()}")
        return rc

    def setCamera(self, model, intr, dist, q_bc, p_bc, t_offset_cam_imu, width, height):
        a = [_d(v) for v in (intr, dist, q_bc, p_bc)]
        self._chk(self.L.ctba_dt_set_camera(self.h, int(model), *[_p(v) for v in a], C.c_double(t_offset_cam_imu), C.c_double(width), C.c_double(height)))

    def setImu(self, gyro_noise_density, acc_noise_density, gyro_random_walk, acc_random_walk, g_W, g_max=7.8, a_max=150.0):
        g = _d(g_W)
        self._chk(self.L.ctba_dt_set_imu(self.h, C.c_double(gyro_noise_density), C.c_double(acc_noise_density), C.c_double(gyro_random_walk), C.c_double(acc_random_walk),
                                         C.c_double(g_max), C.c_double(a_max), _p(g)))

    def initializeStates(self, poses, cam_ts):
        P, t = _d(poses), _d(cam_ts)
        assert P.shape == (len(t), 7)
        self.F = len(t)
        self._chk(self.L.ctba_dt_initialize_states(self.h, self.F, _p(P), _p(t)))

    def initializeLinearVelocity(self):
        self._chk(self.L.ctba_dt_initialize_linear_velocity(self.h))

    def initializeStatesKeep(self, poses, sb):
        P, x = _d(poses), _d(sb)
        assert P.shape == (self.F, 7) and x.shape == (self.F, 9)
        self._chk(self.L.ctba_dt_set_states(self.h, _p(P), _p(x)))

    def setSpeedAndBias(self, sb):
        x = _d(sb)
        assert x.shape == (self.F, 9)
        self._chk(self.L.ctba_dt_set_speed_and_bias(self.h, _p(x)))

    def getStates(self):
        P, sb = np.zeros((self.F, 7)), np.zeros((self.F, 9))
        self._chk(self.L.ctba_dt_get_states(self.h, _p(P), _p(sb)))
        return P, sb

    def set3DPoints(self, xyz):
        x = _d(xyz)
        self._nlm = len(x)
        self._chk(self.L.ctba_dt_set_landmarks(self.h, len(x), _p(x)))

    def get3DPoints(self):
        x = np.zeros((self._nlm, 3))
        self._chk(self.L.ctba_dt_get_landmarks(self.h, _p(x)))
        return x

    def setGlobals(self, scale, p_bp, t_offset_globsens_imu):
        self._scale, self._t_gs = float(scale), float(t_offset_globsens_imu)
        self._chk(self.L.ctba_dt_set_globals(self.h, C.c_double(scale), _p(_d(p_bp)), C.c_double(t_offset_globsens_imu)))

    def setPbp(self, p_bp):
        """setPbp (estimator.h:731): scale and the time offset keep the values of the last setGlobals"""
        self._chk(self.L.ctba_dt_set_globals(self.h, C.c_double(self._scale), _p(_d(p_bp)), C.c_double(self._t_gs)))

    def setCorrections(self, corr9, t_gs_corr=0.0):
        c = _d(corr9)
        assert c.size == 9
        self._chk(self.L.ctba_dt_set_corrections(self.h, _p(c), C.c_double(t_gs_corr)))

    def getGlobals(self):
        o = np.zeros(13)
        self._chk(self.L.ctba_dt_get_globals(self.h, _p(o)))
        return o

    def setFlags(self, optimize_scale, optimize_pbp, optimize_camimu_extrinsics, optimize_t_offset_globsens_imu):
        self._chk(self.L.ctba_dt_set_flags(self.h, int(optimize_scale), int(optimize_pbp), int(optimize_camimu_extrinsics), int(optimize_t_offset_globsens_imu)))

    def addFrames(self, state_ids, obs_begin, uv, vel, lm_index, sigma_px):
        sid = np.ascontiguousarray(state_ids, dtype=np.int32); ob = np.ascontiguousarray(obs_begin, dtype=np.int32)
        u, v = _d(uv), _d(vel); li = np.ascontiguousarray(lm_index, dtype=np.int32)
        return self._chk(self.L.ctba_dt_add_frames(self.h, len(sid), _p(sid), _p(ob), _p(u), _p(v), _p(li), C.c_double(sigma_px)))

    def addImuMeasurements(self, frame0, t, gyr, acc):
        t, g, a = _d(t), _d(gyr), _d(acc)
        self._chk(self.L.ctba_dt_add_imu(self.h, int(frame0), len(t), _p(t), _p(g), _p(a)))

    def addGlobalSensMeasurements(self, t, pos, frame0, sigma):
        t, x = _d(t), _d(pos); f0 = np.ascontiguousarray(frame0, dtype=np.int32)
        return self._chk(self.L.ctba_dt_add_gp(self.h, len(t), _p(t), _p(x), _p(f0), C.c_double(sigma)))

    def computeValidReprojections(self, pt_begin, uv, lm_index, width, height, cell_size_px):
        pb = np.ascontiguousarray(pt_begin, dtype=np.int32); u = _d(uv); li = np.ascontiguousarray(lm_index, dtype=np.int32)
        F = pb.size - 1
        valid = np.zeros(max(li.size, 1), dtype=np.uint8); nv = np.zeros(max(F, 1), dtype=np.int32)
        self._chk(self.L.ctba_dt_compute_valid_reprojections(self.h, int(F), _p(pb), _p(u), _p(li), int(width), int(height), int(cell_size_px), _p(valid), _p(nv)))
        return valid[:li.size].astype(bool), nv[:F]

    def counts(self):
        c = np.zeros(5, dtype=np.int64)
        self._chk(self.L.ctba_dt_counts(self.h, _p(c)))
        return dict(zip(["frames", "obs", "imu", "gp", "cols"], [int(v) for v in c]))

    def evaluate(self, kind, jac=True):
        c = self.counts()
        nres, w = [(2 * c["obs"], 17), (15 * c["imu"], 30), (3 * c["gp"], 17)][kind]
        r = np.zeros(nres); J = np.zeros((nres, w)) if jac else None
        self._chk(self.L.ctba_dt_evaluate(self.h, int(kind), _p(r), _p(J) if jac else None))
        return r, J

    def cost(self):
        o = np.zeros(4)
        self._chk(self.L.ctba_dt_cost(self.h, _p(o)))
        return o

    def evaluateReprojectionErrors(self):
        o = np.zeros(self.counts()["obs"]); self._chk(self.L.ctba_dt_evaluate_errors(self.h, 0, _p(o))); return o

    def evaluateImuErrors(self):
        o = np.zeros((self.counts()["imu"], 15)); self._chk(self.L.ctba_dt_evaluate_errors(self.h, 1, _p(o))); return o

    def evaluateGlobalPosErrors(self):
        o = np.zeros(self.counts()["gp"]); self._chk(self.L.ctba_dt_evaluate_errors(self.h, 2, _p(o))); return o

    def setStream(self, cuda_stream):
        self._chk(self.L.ctba_dt_set_stream(self.h, C.c_void_p(int(cuda_stream))))

    def invalidateDevice(self):
        self._chk(self.L.ctba_dt_invalidate_device(self.h))

    def imuRedoCount(self, k):
        v = C.c_int(0)
        self._chk(self.L.ctba_dt_imu_redo_count(self.h, int(k), C.byref(v)))
        return v.value

    def optimize(self, max_num_iterations):
        cap = max_num_iterations + 2
        its = (_Iteration * cap)()
        s = _Summary()
        s.iterations = C.cast(its, C.POINTER(_Iteration))
        s.iterations_capacity = cap
        self._chk(self.L.ctba_dt_solve(self.h, int(max_num_iterations), C.byref(s)))
        n = min(s.num_iterations, cap)
        log = np.array([[its[i].cost, its[i].cost_change, its[i].gradient_max_norm, its[i].step_norm, its[i].tr_ratio, its[i].tr_radius,
                         its[i].successful] for i in range(n)]).reshape(n, 7)
        return dict(termination_type=s.termination_type, initial_cost=s.initial_cost, final_cost=s.final_cost,
                    num_successful_steps=s.num_successful_steps, num_unsuccessful_steps=s.num_unsuccessful_steps, iterations=log,
                    time_total_s=s.time_total_s, time_eval_s=s.time_eval_s, time_build_s=s.time_build_s, time_solve_s=s.time_solve_s,
                    kernel_launches=int(s.kernel_launches))

    def setOption(self, name, value):
        self._chk(self.L.ctba_dt_set_option(self.h, name.encode(), C.c_double(value)))

    def getStat(self, name):
        v = C.c_double(0.0)
        self._chk(self.L.ctba_dt_get_stat(self.h, name.encode(), C.byref(v)))
        return v.value


def load_dt_scene(prob, d, with_imu=True, with_gp=True, with_frames=True):
    """Feed a discrete-time scene (scene.dt_scene_from) into any object with the DiscreteTimeEstimator method set (the CUDA problem or the
    CPU oracle), in the reference driver's order (src/optimize_discrete_time.cpp:436-570)."""
    prob.setCamera(d["cam_model"], d["intr"], d["dist"], d["q_bc"], d["p_bc"], d["t_offset_cam_imu"], d["width"], d["height"])
    prob.setImu(d["gyro_noise_density"], d["acc_noise_density"], d["gyro_random_walk"], d["acc_random_walk"], d["g_W"])
    prob.initializeStates(d["poses"], d["cam_ts"])
    prob.initializeLinearVelocity()
    prob.setGlobals(d["scale"], d["p_bp"], d["t_offset_gs_imu"])
    prob.setFlags(*d["flags"])
    prob.set3DPoints(d["landmarks"])
    if with_imu:
        for (f0, t, gyr, acc) in d["imu"]:
            prob.addImuMeasurements(f0, t, gyr, acc)
    if with_frames and d["uv"].shape[0]:
        prob.addFrames(d["state_ids"], d["obs_begin"], d["uv"], d["vel"], d["lm_index"], d["sigma_reproj_px"])
    if with_gp and d["gp_t"].size:
        prob.addGlobalSensMeasurements(d["gp_t"], d["gp_pos"], d["gp_frame0"], d["gp_std"])
