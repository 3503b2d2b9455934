"""ctba_align_system_noalloc (include/ctba.h): the allocation-free entry point, scratch in a file-scope static __device__ array.
No counterpart in the reference.  Checked through the C ABI against the EXISTING entry
points on the same problem: ctba_evaluate(CTBA_ALIGN) gives the weighted residuals / Jacobians from which the normal equations follow,
ctba_solve reports the cost k_align_accumulate + k_align_reduce computed at the same state.  Two (and more) calls in a row share the
static scratch; the later ones run at a different state and with fewer CTAs, so anything a call left behind would show.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from align_case import load_alignment, make_alignment_case

CAUCHY_A = 0.10          # CauchyLoss(0.10), bspline_so3r3.h:1189
MAX_MEAS = 1 << 20       # CTBA_NOALLOC_MAX_MEASUREMENTS


def _bind(L):
    f = L.ctba_align_system_noalloc
    f.restype = C.c_int
    f.argtypes = [C.c_int, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_double, C.c_double, C.c_void_p, C.c_void_p, C.c_double, C.c_int,
                  C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_char_p, C.c_int]
    return f


def _system_from_rows(r, J, b):
    """Normal equations of the robustified problem from the unrobustified weighted rows, per 3-row block as Ceres' corrector does it
    for CauchyLoss (rho'' < 0: residual and Jacobian scaled by sqrt(rho')).  Returns H (11 x 11), g (11), cost."""
    r3 = r.reshape(-1, 3); J3 = J.reshape(-1, 3, J.shape[1])
    s = (r3 * r3).sum(axis=1)
    w = np.sqrt(1.0 / (1.0 + s / b))
    rc = r3 * w[:, None]; Jc = J3 * w[:, None, None]
    Jf = Jc.reshape(-1, J.shape[1])
    return Jf.T @ Jf, Jf.T @ rc.reshape(-1), float(0.5 * b * np.log1p(s / b).sum())


def _unpack(sys78):
    H = np.zeros((11, 11))
    iu = np.triu_indices(11)
    H[iu] = sys78[:66]
    H = H + np.triu(H, 1).T
    return H, sys78[66:77], float(sys78[77])


def _check(sys78, r, J, cost_existing=None):
    H, g, cost = _unpack(sys78)
    H0, g0, cost0 = _system_from_rows(r, J, CAUCHY_A ** 2)
    d = np.sqrt(np.diag(H0)); d[d == 0.0] = 1.0
    rn = np.sqrt(2.0 * cost0) if cost0 > 0 else 1.0
    # entries relative to the scale of their rows / columns (Cauchy-Schwarz bound of the entry): the test suite's 1e-10 gate
    np.testing.assert_allclose(H / np.outer(d, d), H0 / np.outer(d, d), rtol=1e-10, atol=1e-10)
    np.testing.assert_allclose(g / d, g0 / d, rtol=1e-10, atol=1e-10 * rn)
    assert abs(cost - cost0) <= 1e-10 * cost0
    if cost_existing is not None:   # the cost the existing kernels (k_align_accumulate + k_align_reduce) computed at the same state
        assert abs(cost - cost_existing) <= 1e-10 * cost_existing


@pytest.mark.gpu
@pytest.mark.parametrize("N", [6, 4])
def test_no_alloc_alignment_system_matches_existing_entry_points(pkg, oracle, scene, N):
    import torch
    case = make_alignment_case(scene, oracle, N, seed=70 + N, duration_s=60.0, gp_hz=20.0)
    g = pkg.BSplineSO3R3(N, case["dt_ns"], case["start_t_ns"], device=0)
    n = load_alignment(g, case, True, True)
    assert n == len(case["t_ns"])
    f = _bind(g.L)
    dev = torch.device("cuda:0")
    d_q = torch.from_numpy(np.ascontiguousarray(case["q"])).to(dev)
    d_p = torch.from_numpy(np.ascontiguousarray(case["p"])).to(dev)
    # two samples the window arithmetic must skip (before the spline start, beyond the last window) ride along; they contribute nothing
    t_all = np.r_[case["start_t_ns"] - 5, case["t_ns"], case["start_t_ns"] + int(1e12)].astype(np.int64)
    z_all = np.vstack([np.full((1, 3), 1e3), case["z"], np.full((1, 3), -1e3)])
    d_t = torch.from_numpy(t_all).to(dev)
    d_z = torch.from_numpy(np.ascontiguousarray(z_all)).to(dev)
    d_sys = torch.full((78,), float("nan"), dtype=torch.float64, device=dev)
    err = C.create_string_buffer(256)
    T0 = np.ascontiguousarray(case["T_wg_init"], dtype=np.float64)
    P = lambda a: a.ctypes.data_as(C.c_void_p)

    def call(x12, scale, t_gp, n_meas, t_ptr, z_ptr):
        d_sys.fill_(float("nan"))
        torch.cuda.synchronize()
        rc = f(N, case["dt_ns"], case["start_t_ns"], 0, None, scale, t_gp, P(T0), P(x12), CAUCHY_A, len(case["q"]), d_q.data_ptr(), d_p.data_ptr(),
               n_meas, t_ptr, z_ptr, case["noise_arg"], d_sys.data_ptr(), err, 256)
        assert rc == 0, err.value
        torch.cuda.synchronize()
        return d_sys.cpu().numpy().copy()

    # ---- call 1: the initial state, against ctba_evaluate and the initial cost ctba_solve reports
    r1, J1 = g.evaluate(7)
    x1 = np.zeros(12); x1[5] = 1.0
    gl = g.getGlobals()
    s1 = call(x1, gl[18], gl[17], len(t_all), d_t.data_ptr(), d_z.data_ptr())
    sg = g.optimize(30)
    _check(s1, r1, J1, sg["initial_cost"])
    # ---- call 2, right after: the optimised state (different values everywhere), same scratch
    r2, J2 = g.evaluate(7)
    _, corr = g.getWorldColmapTransformation()
    gl = g.getGlobals()
    x2 = np.r_[gl[9], gl[0], corr[:4], corr[4:], gl[10:13]]
    s2 = call(x2, gl[18], gl[17], len(t_all), d_t.data_ptr(), d_z.data_ptr())
    _check(s2, r2, J2, sg["final_cost"])
    assert abs(_unpack(s2)[2] - _unpack(s1)[2]) > 1e-3 * _unpack(s1)[2]          # the two states really differ
    # ---- call 3: a single CTA (the rows of the other CTAs keep call 2's partial sums and must not be read)
    m = 100
    s3 = call(x2, gl[18], gl[17], m, d_t.data_ptr() + 8, d_z.data_ptr() + 24)
    _check(s3, r2[:3 * m], J2[:3 * m])
    # ---- call 4 repeats call 1 bit for bit (fixed-order reductions; nothing of calls 2 and 3 leaks)
    s4 = call(x1, g.getGlobals()[18], g.getGlobals()[17], len(t_all), d_t.data_ptr(), d_z.data_ptr())
    np.testing.assert_array_equal(s4, s1)
    # ---- no measurements at all: an all-zero system, not stale scratch
    s5 = call(x1, gl[18], gl[17], 0, None, None)
    np.testing.assert_array_equal(s5, np.zeros(78))


@pytest.mark.gpu
def test_no_alloc_rejects_sizes_beyond_the_static_scratch(pkg):
    import torch
    f = _bind(pkg.lib())
    dev = torch.device("cuda:0")
    d = torch.zeros(128, dtype=torch.float64, device=dev)
    T = np.r_[0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0]; x = np.zeros(12); x[5] = 1.0
    err = C.create_string_buffer(256)
    P = lambda a: a.ctypes.data_as(C.c_void_p)
    rc = f(4, 100_000_000, 0, 0, None, 1.0, 0.0, P(T), P(x), CAUCHY_A, 8, d.data_ptr(), d.data_ptr(), MAX_MEAS + 1, d.data_ptr(), d.data_ptr(), 1.0,
           d.data_ptr(), err, 256)
    assert rc == -1 and b"exceed the fixed limit" in err.value, err.value
    rc = f(5, 100_000_000, 0, 0, None, 1.0, 0.0, P(T), P(x), CAUCHY_A, 8, d.data_ptr(), d.data_ptr(), 1, d.data_ptr(), d.data_ptr(), 1.0,
           d.data_ptr(), err, 256)
    assert rc == -1 and b"order" in err.value
