"""align_spline_to_global_frame: the frozen-spline Sim(3) alignment problem.
CPU part: the shared host/device pieces of csrc/ctba_align.h (per-sample normal-equation accumulation, the <= 11-unknown LM step and the
trust-region loop) compiled for the host by the tests-only harness, against the oracle's Jet-autodiff problem + LM restatement:
per-iteration cost 1e-8 relative, final cost 1e-6, recovered alignment close to the truth.  The GPU part runs the same problem through
the C ABI (ctba_solve on a handle that holds alignment measurements)."""
import ctypes as C

import numpy as np
import pytest
from scipy.spatial.transform import Rotation as Rot

from align_case import load_alignment, make_alignment_case

P = lambda a: a.ctypes.data_as(C.c_void_p)


class _It(C.Structure):
    _fields_ = [("cost", C.c_double), ("cost_change", C.c_double), ("gradient_max_norm", C.c_double), ("step_norm", C.c_double),
                ("tr_ratio", C.c_double), ("tr_radius", C.c_double), ("successful", C.c_int)]


class _Sum(C.Structure):
    _fields_ = [("termination_type", C.c_int), ("initial_cost", C.c_double), ("final_cost", C.c_double), ("num_successful_steps", C.c_int),
                ("num_unsuccessful_steps", C.c_int), ("num_iterations", C.c_int), ("iterations", C.POINTER(_It)), ("iterations_capacity", C.c_int),
                ("time_total_s", C.c_double), ("time_eval_s", C.c_double), ("time_build_s", C.c_double), ("time_solve_s", C.c_double),
                ("kernel_launches", C.c_int64)]


def _free_mask(optimize_pbp, optimize_t):
    m = 0b00011111110                     # scale, R_corr (3), p_corr (3)
    if optimize_t: m |= 1
    if optimize_pbp: m |= 0b111 << 8
    return m


def _host_lm(hostdbg, case, optimize_pbp, optimize_t, max_iters):
    N = case["N"]; dt_ns = case["dt_ns"]
    cfg = np.zeros(19); cfg[0] = 4; cfg[1:5] = [1, 1, 0, 0]; cfg[12] = 1.0; cfg[16] = dt_ns * 1e-9; cfg[17] = 1.0; cfg[18] = case["scale_init"]
    t_init_ns = int(case["t_off_init"] * 1e9)
    st_ns = case["t_ns"] - case["start_t_ns"] - t_init_ns
    keep = (st_ns >= 0) & (st_ns // dt_ns + N <= len(case["q"]))
    st_ns = st_ns[keep]
    s = (st_ns // dt_ns).astype(np.int32); st_s = st_ns * 1e-9; k = (st_ns // dt_ns).astype(np.float64)
    meas = np.ascontiguousarray(case["z"][keep]); inv_std = np.full(len(s), 1.0 / case["noise_arg"])
    x12 = np.zeros(12); x12[5] = 1.0
    its = (_It * (max_iters + 2))()
    sm = _Sum(); sm.iterations = its; sm.iterations_capacity = max_iters + 2
    sys_out = np.zeros(78)
    q = np.ascontiguousarray(case["q"]); p = np.ascontiguousarray(case["p"]); T = np.ascontiguousarray(case["T_wg_init"])
    rc = hostdbg.dbg_align_lm(N, P(cfg), P(T), P(x12), C.c_uint(_free_mask(optimize_pbp, optimize_t)), P(q), P(p), len(s), P(s), P(st_s), P(k),
                              P(meas), P(inv_std), C.c_double(0.10), max_iters, C.byref(sm), P(sys_out))
    assert rc == 0
    log = np.array([[i.cost, i.cost_change, i.gradient_max_norm, i.step_norm, i.tr_ratio, i.tr_radius, i.successful] for i in its[:sm.num_iterations]])
    return sm, log, x12


@pytest.mark.parametrize("N", [4, 6])
@pytest.mark.parametrize("optimize_pbp,optimize_t", [(True, True), (False, False), (True, False)])
def test_alignment_lm_host_pieces_vs_oracle(oracle, scene, hostdbg, N, optimize_pbp, optimize_t):
    case = make_alignment_case(scene, oracle, N, seed=40 + N)
    o = oracle.OracleSpline(N, case["dt_ns"], case["start_t_ns"])
    n_added = load_alignment(o, case, optimize_pbp, optimize_t)
    assert n_added == len(case["t_ns"])
    max_iters = 30
    so = o.optimize(max_iters)
    sm, log, x12 = _host_lm(hostdbg, case, optimize_pbp, optimize_t, max_iters)
    lo = so["iterations"]
    assert sm.termination_type == so["termination_type"] == 0
    assert len(log) == len(lo) and sm.num_successful_steps == so["num_successful_steps"]
    np.testing.assert_allclose(log[:, 0], lo[:, 0], rtol=1e-8)                       # cost of every iteration
    np.testing.assert_array_equal(log[:, 6], lo[:, 6])                                # same accept / reject decisions
    np.testing.assert_allclose(log[:, 5], lo[:, 5], rtol=1e-6)                        # trust-region radius
    assert abs(sm.final_cost - so["final_cost"]) <= 1e-6 * so["final_cost"]
    # the optimised blocks: oracle state vs the packed state of the shared LM
    T_o, corr_o = o.getWorldColmapTransformation()
    g = o.getGlobals()
    np.testing.assert_allclose(x12[2:6], corr_o[:4], atol=1e-9)
    np.testing.assert_allclose(x12[6:9], corr_o[4:], atol=1e-8)
    np.testing.assert_allclose(x12[1], g[0], atol=1e-9)                               # scale_corr
    np.testing.assert_allclose(x12[9:12], g[10:13], atol=1e-8)                        # p_bp
    np.testing.assert_allclose(x12[0], g[9], atol=1e-9)                               # t_s_gp_corr
    if not optimize_pbp: assert np.all(x12[9:12] == 0.0)
    if not optimize_t: assert x12[0] == 0.0
    # and the alignment is actually recovered (5 % outliers of ~3 m are rejected by the loss)
    tr = case["truth"]
    ang = (Rot.from_quat(T_o[:4]).inv() * Rot.from_quat(tr["R_wg"])).magnitude()
    assert ang < (2e-3 if optimize_pbp and optimize_t else 2e-2)      # a frozen p_bp = 0 / time offset = 0 biases the fit
    assert abs(case["scale_init"] + g[0] - tr["scale"]) < (5e-3 if optimize_pbp and optimize_t else 3e-2) * tr["scale"]
    if optimize_pbp and optimize_t:
        assert abs(g[9] - tr["t_off"]) < 8e-3


def test_alignment_host_logic_on_planning_handle(pkg, oracle, scene):
    """Host side of the C ABI without a GPU (device -1): skip rules of addWorldColmapAlignmentMeasurement, get/setWorldColmapTransformation
    composition (bspline_so3r3.h:1827-1840) and the loud failure of the compute entry points."""
    N = 4
    case = make_alignment_case(scene, oracle, N, seed=3, duration_s=5.0)
    g = pkg.BSplineSO3R3(N, case["dt_ns"], case["start_t_ns"], device=-1)
    o = oracle.OracleSpline(N, case["dt_ns"], case["start_t_ns"])
    # measurements before the spline start and beyond the last window are skipped by both
    t_ns = np.r_[case["start_t_ns"] - 5, case["t_ns"], case["start_t_ns"] + int(1e12)].astype(np.int64)
    z = np.vstack([np.zeros((1, 3)), case["z"], np.zeros((1, 3))])
    case2 = dict(case, t_ns=t_ns, z=z)
    n_g = load_alignment(g, case2, True, True)
    n_o = load_alignment(o, case2, True, True)
    assert n_g == n_o == len(case["t_ns"]) and g.numAlignmentMeasurements() == n_o
    corr = np.r_[Rot.from_rotvec([0.1, 0.2, -0.3]).as_quat(), [0.5, -0.25, 1.0]]
    g.setWorldColmapTransformation(case["T_wg_init"], corr); o.setWorldColmapTransformation(case["T_wg_init"], corr)
    Tg, cg = g.getWorldColmapTransformation(); To, co = o.getWorldColmapTransformation()
    np.testing.assert_allclose(Tg, To, rtol=0, atol=1e-15)
    np.testing.assert_array_equal(cg, co)
    with pytest.raises(pkg.CtbaError):
        g.optimize(3)
    with pytest.raises(pkg.CtbaError):
        g.evaluate(7)


@pytest.mark.gpu
@pytest.mark.parametrize("N", [4, 6])
@pytest.mark.parametrize("optimize_pbp,optimize_t", [(True, True), (False, False)])
def test_alignment_gpu_vs_oracle(pkg, oracle, scene, N, optimize_pbp, optimize_t):
    """The device path through the C ABI: residuals / Jacobians 1e-10, unit-weight errors, LM trace 1e-8, final cost 1e-6, same state."""
    case = make_alignment_case(scene, oracle, N, seed=60 + N, duration_s=40.0, gp_hz=20.0)
    g = pkg.BSplineSO3R3(N, case["dt_ns"], case["start_t_ns"], device=0)
    o = oracle.OracleSpline(N, case["dt_ns"], case["start_t_ns"])
    assert load_alignment(g, case, optimize_pbp, optimize_t) == load_alignment(o, case, optimize_pbp, optimize_t) == len(case["t_ns"])
    r_g, J_g = g.evaluate(7)
    r_o, J_o, _ = o.evaluate(7)
    np.testing.assert_allclose(r_g, r_o, rtol=1e-10, atol=1e-10 * np.abs(r_o).max())
    np.testing.assert_allclose(J_g, J_o[:, 6 * N:], rtol=1e-10, atol=1e-10 * np.abs(J_o).max())
    e_g = g.evaluateWorldColmapAlignmentErrors()
    np.testing.assert_allclose(e_g, np.linalg.norm(r_o.reshape(-1, 3), axis=1) * case["noise_arg"], rtol=1e-10)
    sg = g.optimize(30)
    so = o.optimize(30)
    lg, lo = sg["iterations"], so["iterations"]
    assert sg["termination_type"] == so["termination_type"] == 0 and len(lg) == len(lo)
    np.testing.assert_allclose(lg[:, 0], lo[:, 0], rtol=1e-8)
    np.testing.assert_array_equal(lg[:, 6], lo[:, 6])
    assert abs(sg["final_cost"] - so["final_cost"]) <= 1e-6 * so["final_cost"]
    Tg, cg = g.getWorldColmapTransformation(); To, co = o.getWorldColmapTransformation()
    np.testing.assert_allclose(Tg, To, atol=1e-8)
    gg, go = g.getGlobals(), o.getGlobals()
    np.testing.assert_allclose(gg[[0, 9, 10, 11, 12]], go[[0, 9, 10, 11, 12]], atol=1e-8)
    # a second solve from the optimum terminates at once with the same cost (state round trip host <-> device)
    s2 = g.optimize(5)
    assert abs(s2["initial_cost"] - sg["final_cost"]) <= 1e-9 * sg["final_cost"]
    # relinearizeSplineGlobSensTimeOffset after the solve (align_spline_to_global_frame.cpp:217)
    g.relinearize(); o.relinearize()
    np.testing.assert_allclose(g.getGlobals()[17], o.getGlobals()[17], atol=1e-9)
