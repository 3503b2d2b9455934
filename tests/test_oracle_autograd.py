"""Independent checks of the ORACLE's Jet Jacobians: a torch-float64 restatement of the spline and of the four
residuals with `torch.autograd.functional.jacobian` over the right-perturbation tangent space — EVERY column block (rotation knots,
position knots, bias knots, gravity, IMU / GP time offsets, scale, camera-IMU extrinsics, p_bp, landmarks), not only the position-knot
columns that the central-difference test covers — and the Sophus sympy statement of SO3 exp / Dx_this_mul_exp_x_at_0
(/root/reference/thirdparty/Sophus/py/sophus/so3.py, restated with plain sympy so that the test also runs where /root/reference is absent).
Reference semantics: include/optimization/continuous_time/ceres_spline_helper.h:97-330, imu_cost_functor.h:48-77, 106-139,
global_position_cost_functor.h:45-73, reprojection_cost_functor.h:47-109, lie_local_parametrization.h:34-55.  CPU only."""
from math import comb, factorial

import numpy as np
import pytest
import torch

torch.set_default_dtype(torch.float64)


def _blend(N, cumulative):
    M = np.zeros((N, N))
    for s in range(N):
        for n in range(N):
            acc = 0.0
            for l in range(s, N):
                acc += (-1) ** (l - s) * comb(N, l - s) * (N - 1 - l) ** (N - 1 - n)
            M[s, n] = comb(N - 1, n) * acc / factorial(N - 1)
    if cumulative:
        M = np.cumsum(M[::-1], axis=0)[::-1].copy()
    return torch.tensor(M)


# ---- quaternions [x y z w] ----
def qmul(a, b):
    av, aw, bv, bw = a[:3], a[3], b[:3], b[3]
    return torch.cat([aw * bv + bw * av + torch.linalg.cross(av, bv), (aw * bw - av @ bv).reshape(1)])


def qconj(q):
    return torch.cat([-q[:3], q[3:]])


def qrot(q, v):
    u = 2.0 * torch.linalg.cross(q[:3], v)
    return v + q[3] * u + torch.linalg.cross(q[:3], u)


def qexp(w):
    th = torch.sqrt(w @ w)
    return torch.cat([torch.sin(th / 2) / th * w, torch.cos(th / 2).reshape(1)])


def qlog(q):
    n = torch.sqrt(q[:3] @ q[:3])
    return 2.0 * torch.atan(n / q[3]) / n * q[:3]


def plus(q, d):
    """q * exp(d) to first order in d (the derivative at d = 0 is that of the exponential)"""
    e = torch.cat([0.5 * d, torch.ones(1)])
    r = qmul(q, e)
    return r / torch.sqrt(r @ r)


def powers(u, N, deriv):
    out = []
    for j in range(N):
        c = 1.0
        for m in range(deriv):
            c *= (j - m)
        out.append(c * u ** (j - deriv) if j >= deriv else u * 0.0)
    return torch.stack(out)


def spline(qs, ps, u, inv_dt, N):
    """R (quat), body omega, position, world acceleration of the cumulative SO(3) x R^3 B-spline at u (u may carry a gradient)"""
    Mc, M = _blend(N, True), _blend(N, False)
    lam = Mc @ powers(u, N, 0)
    lam1 = inv_dt * (Mc @ powers(u, N, 1))
    R = qs[0]
    w = torch.zeros(3)
    for i in range(1, N):
        d = qlog(qmul(qconj(qs[i - 1]), qs[i]))
        A = qexp(lam[i] * d)
        R = qmul(R, A)
        w = qrot(qconj(A), w) + lam1[i] * d
    pos = (M @ powers(u, N, 0)) @ ps
    acc = inv_dt ** 2 * ((M @ powers(u, N, 2)) @ ps)
    return R, w, pos, acc


def bias_at(B4, ub):
    return (_blend(4, False) @ powers(ub, 4, 0)) @ B4


def project_radtan(intr, dist, pc):
    x, y = pc[0] / pc[2], pc[1] / pc[2]
    k1, k2, p1, p2 = dist
    r2 = x * x + y * y
    rad = k1 * r2 + k2 * r2 * r2
    xd = x + x * rad + 2 * p1 * x * y + p2 * (r2 + 2 * x * x)
    yd = y + y * rad + 2 * p2 * x * y + p1 * (r2 + 2 * y * y)
    return torch.stack([intr[0] * xd + intr[2], intr[1] * yd + intr[3]])


def _T(a):
    return torch.tensor(np.asarray(a, dtype=np.float64))


def _colscale_err(J, Jo):
    cs = np.maximum(np.abs(Jo).max(axis=0, keepdims=True), 1e-300)
    return float((np.abs(J - Jo) / (1e-9 * np.abs(Jo) + 1e-9 * cs + 1e-11 * np.abs(Jo).max())).max())


@pytest.mark.parametrize("N", [4, 6])
def test_jet_jacobians_equal_torch_autograd(oracle, pkg, scene, N):
    sc = scene.make_scene(order=N, duration_s=2.0, n_landmarks=40, track_len=6, imu_hz=50, gp_hz=10, dt_s_bias=0.1, seed=31 + N,
                          flags=(True, True, True, True, True), cam="euroc")
    o = oracle.OracleSpline(N, sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(o, sc, with_bias_cost=False)
    corr = np.array([0.002, 0.001, -0.002, 0.0015, 0.0, 0.002, -0.003, 0.001, 0.004, -0.003])   # scale, R_bc (xyzw), p_bc, t_cam_imu, t_gp (small enough that observations stay inside the 10 px gate)
    corr[4] = np.sqrt(1 - np.sum(corr[1:4] ** 2))
    o.setCorrections(corr)
    rng = np.random.default_rng(5)
    nb = o.numBiasKnots()
    ba, bg = rng.normal(size=(nb, 3)) * 0.05, rng.normal(size=(nb, 3)) * 0.01
    o.setBiases(ba, bg)
    q, p = _T(sc["knot_q"]), _T(sc["knot_p"])
    dt_ns, dt_s, inv_dt = sc["dt_ns"], sc["dt_ns"] * 1e-9, 1e9 / sc["dt_ns"]
    dt_ns_b = int(np.uint64(sc["dt_s_bias"] / 1e-9))
    K = q.shape[0]
    g_w, p_bp, scale = _T(sc["g"]), _T(sc["p_bp"]), sc["scale"]
    t_corr_imu, t_corr_gp, scale_corr = corr[8], corr[9], corr[0]

    def knots(s, d_rot, d_pos):
        return [plus(q[s + i], d_rot[3 * i:3 * i + 3]) for i in range(N)], p[s:s + N] + d_pos.reshape(N, 3)

    # ---------------- gyro + accel ----------------
    r_a, J_a, _ = o.evaluate(1); r_g, J_g, _ = o.evaluate(2)
    rows = 0
    checked = 0
    for t_ns, am, gm in zip(sc["imu_t_ns"], sc["imu_acc"], sc["imu_gyr"]):
        st = int(t_ns + int(-sc["t_offset_cam_imu"] * 1e9) - sc["start_t_ns"])
        if st < 0 or st // dt_ns + N > K:
            continue
        s, k, st_s = st // dt_ns, float(st // dt_ns), st * 1e-9
        stb = int(t_ns - sc["start_t_ns"]); sb, ub = stb // dt_ns_b, float(stb % dt_ns_b) / dt_ns_b
        i0 = rows; rows += 3
        if sb + N > nb or checked >= 6 or (i0 // 3) % 7:
            continue
        checked += 1

        def f_gyr(x):   # [3N rot | 12 bias | t]
            qs, _ = knots(s, x[:3 * N], torch.zeros(3 * N))
            B4 = _T(bg[sb:sb + 4]) + x[3 * N:3 * N + 12].reshape(4, 3)
            u = (st_s - (t_corr_imu + x[3 * N + 12]) - k * dt_s) / dt_s
            _, w, _, _ = spline(qs, p[s:s + N], u, inv_dt, N)
            return (w - _T(gm) + bias_at(B4, torch.tensor(ub))) / sc["sigma_omega_c"]

        def f_acc(x):   # [3N rot | 3N pos | 12 bias | 3 g | t]
            qs, ps = knots(s, x[:3 * N], x[3 * N:6 * N])
            B4 = _T(ba[sb:sb + 4]) + x[6 * N:6 * N + 12].reshape(4, 3)
            u = (st_s - (t_corr_imu + x[6 * N + 15]) - k * dt_s) / dt_s
            R, _, _, a_w = spline(qs, ps, u, inv_dt, N)
            return (qrot(qconj(R), a_w + g_w + x[6 * N + 12:6 * N + 15]) - _T(am) + bias_at(B4, torch.tensor(ub))) / sc["sigma_acc_c"]

        for f, nx, r_o, J_o in ((f_gyr, 3 * N + 13, r_g, J_g), (f_acc, 6 * N + 16, r_a, J_a)):
            x0 = torch.zeros(nx)
            np.testing.assert_allclose(f(x0).numpy(), r_o[i0:i0 + 3], rtol=1e-9, atol=1e-9)
            J = torch.autograd.functional.jacobian(f, x0).numpy()
            assert _colscale_err(J, J_o[i0:i0 + 3]) <= 1.0, _colscale_err(J, J_o[i0:i0 + 3])
            assert np.abs(J_o[i0:i0 + 3, -1]).max() > 0 and np.abs(J_o[i0:i0 + 3, :3 * N]).max() > 0    # time-offset and rotation columns are live
    assert checked >= 3

    # ---------------- global position ----------------
    r_p, J_p, _ = o.evaluate(3)
    rows = 0; checked = 0
    for t_ns, z in zip(sc["gp_t_ns"], sc["gp_pos"]):
        st = int(t_ns - sc["start_t_ns"] - int(sc["t_s_gp"] * 1e9))
        if st < 0 or st // dt_ns + N > K:
            continue
        s, k, st_s = st // dt_ns, float(st // dt_ns), st * 1e-9
        i0 = rows; rows += 3
        if checked >= 5 or (i0 // 3) % 3:
            continue
        checked += 1

        def f_gp(x):   # [3N rot | 3N pos | p_bp 3 | t | scale]
            qs, ps = knots(s, x[:3 * N], x[3 * N:6 * N])
            u = (st_s - (t_corr_gp + x[6 * N + 3]) - k * dt_s) / dt_s
            R, _, pos, _ = spline(qs, ps, u, inv_dt, N)
            return ((scale + scale_corr + x[6 * N + 4]) * qrot(R, p_bp + x[6 * N:6 * N + 3]) + pos - _T(z)) / sc["gp_std"]

        x0 = torch.zeros(6 * N + 5)
        np.testing.assert_allclose(f_gp(x0).numpy(), r_p[i0:i0 + 3], rtol=1e-9, atol=1e-9)
        J = torch.autograd.functional.jacobian(f_gp, x0).numpy()
        assert _colscale_err(J, J_p[i0:i0 + 3]) <= 1.0
    assert checked >= 3

    # ---------------- reprojection ----------------
    r_r, J_r, dropped = o.evaluate(0)
    assert dropped == 0.0
    q_bc0, p_bc0 = _T(sc["q_bc"]), _T(sc["p_bc"])
    R_corr0, p_corr0 = _T(corr[1:5]), _T(corr[5:8])
    intr, dist, lm = _T(sc["intr"]), _T(sc["dist"]), _T(sc["landmarks"])
    row = 0; checked = 0
    for fi, t_ns in enumerate(sc["frame_t_ns"]):
        st = int(t_ns - sc["start_t_ns"])
        b, e = int(sc["obs_begin"][fi]), int(sc["obs_begin"][fi + 1])
        if st < 0 or e == b or st // dt_ns + N > K:
            continue
        s, u = st // dt_ns, float(st % dt_ns) / dt_ns
        for ob in range(b, e):
            i0 = row; row += 2
            if checked >= 8 or (i0 // 2) % 29:
                continue
            if np.abs(r_r[i0:i0 + 2]).max() == 0.0:
                continue   # gated observation: zero rows on both sides by definition
            checked += 1
            uv, li = _T(sc["uv"][ob]), int(sc["lm_index"][ob])

            def f_rp(x):   # [3N rot | 3N pos | scale | R_bc 3 | p_bc 3 | landmark 3]
                qs, ps = knots(s, x[:3 * N], x[3 * N:6 * N])
                R_wb, _, p_wb, _ = spline(qs, ps, torch.tensor(u), inv_dt, N)
                sfac = scale + scale_corr + x[6 * N]
                R_bc = qmul(q_bc0, plus(R_corr0, x[6 * N + 1:6 * N + 4]))
                p_bc = qrot(q_bc0, p_corr0 + x[6 * N + 4:6 * N + 7]) + p_bc0
                R_wc = qmul(R_wb, R_bc)
                p_wc = sfac * qrot(R_wb, p_bc) + p_wb
                pc = qrot(qconj(R_wc), lm[li] + x[6 * N + 7:6 * N + 10] - p_wc)
                return (project_radtan(intr, dist, pc) - uv) / sc["sigma_reproj_px"]

            x0 = torch.zeros(6 * N + 10)
            np.testing.assert_allclose(f_rp(x0).numpy(), r_r[i0:i0 + 2], rtol=1e-9, atol=1e-9)
            J = torch.autograd.functional.jacobian(f_rp, x0).numpy()
            assert _colscale_err(J, J_r[i0:i0 + 2]) <= 1.0, _colscale_err(J, J_r[i0:i0 + 2])
    assert checked >= 4


def test_so3_against_sympy_statement(oracle):
    """Sophus' sympy statement (py/sophus/so3.py: exp :36-46, Dx_this_mul_exp_x_at_0 = d(q * exp(x))/dx at 0 :96-104, test_derivatives
    :205-235) restated with plain sympy: the oracle's exp and its 4 x 3 plus-Jacobian equal the symbolic expressions evaluated in
    50-digit arithmetic."""
    sympy = pytest.importorskip("sympy")
    x = sympy.symbols("x0:3", real=True)
    qs = sympy.symbols("qx qy qz qw", real=True)
    theta = sympy.sqrt(sum(v * v for v in x))
    ex = [sympy.sin(theta / 2) / theta * v for v in x] + [sympy.cos(theta / 2)]

    def mul(a, b):
        av, aw, bv, bw = sympy.Matrix(a[:3]), a[3], sympy.Matrix(b[:3]), b[3]
        v = aw * bv + bw * av + av.cross(bv)
        return [v[0], v[1], v[2], aw * bw - av.dot(bv)]

    prod = mul(qs, ex)
    rng = np.random.default_rng(0)
    for _ in range(4):
        w = rng.normal(size=3) * 0.7
        q = oracle.so3_exp(w)
        val = [sympy.N(e.subs(dict(zip(x, [sympy.Float(float(v), 50) for v in w]))), 30) for e in ex]
        np.testing.assert_allclose(q, np.array(val, dtype=np.float64), rtol=0, atol=1e-15)
        # Dx_this_mul_exp_x_at_0: the limit x -> 0 of d(q * exp(x))/dx, taken along a small symbolic x with the series of sin / cos
        J = oracle.so3_plus_jacobian(q)
        half = [v / 2 for v in x] + [1]                      # exp(x) = (x/2, 1) + O(|x|^2): the derivative at 0 is that of this polynomial
        lin = mul(qs, half)
        Js = sympy.Matrix(lin).jacobian(sympy.Matrix(x)).subs(dict(zip(qs, [sympy.Float(float(v), 50) for v in q])))
        np.testing.assert_allclose(J, np.array(Js.tolist(), dtype=np.float64), rtol=0, atol=1e-15)
        # and numerically: the full symbolic product differentiated at a tiny x agrees with it to O(|x|)
        tiny = dict(zip(x, [sympy.Float("1e-20"), sympy.Float("2e-20"), sympy.Float("-1e-20")]))
        Jf = sympy.Matrix(prod).jacobian(sympy.Matrix(x)).subs(dict(zip(qs, [sympy.Float(float(v), 50) for v in q]))).subs(tiny)
        np.testing.assert_allclose(np.array(sympy.N(Jf, 30).tolist(), dtype=np.float64), J, rtol=0, atol=1e-12)
