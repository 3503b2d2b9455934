"""Pins the oracle against the only numeric anchors the reference tree holds for this path
(SURVEY §8c): Appendix-E spline constants (closed form of spline_common.h:54-120), the Sophus SO3
fixture set (thirdparty/Sophus/test/core/test_so3.cpp:28-61) with Sophus' own test properties
(tests.hpp: expLogTest, expMapTest, groupActionTest, productTest, derivativeTest), and the generated
plus-Jacobian (thirdparty/Sophus/py/cpp_gencode/So3_Dx_this_mul_exp_x_at_0.cpp).  CPU only."""
import numpy as np
import pytest

from conftest import relerr

PI = np.pi


def test_spline_constants_appendix_e(oracle):
    M4 = np.array([[1, -3, 3, -1], [4, 0, -6, 3], [1, 3, 3, -3], [0, 0, 0, 1]]) / 6.0
    C4 = np.array([[6, 0, 0, 0], [5, 3, -3, 1], [1, 3, 3, -2], [0, 0, 0, 1]]) / 6.0
    B4 = np.array([[1, 1, 1, 1], [0, 1, 2, 3], [0, 0, 2, 6], [0, 0, 0, 6]], dtype=float)
    M6 = np.array([[1, -5, 10, -10, 5, -1], [26, -50, 20, 20, -20, 5], [66, 0, -60, 0, 30, -10], [26, 50, 20, -20, -20, 10],
                   [1, 5, 10, 10, 5, -5], [0, 0, 0, 0, 0, 1]]) / 120.0
    C6 = np.array([[120, 0, 0, 0, 0, 0], [119, 5, -10, 10, -5, 1], [93, 55, -30, -10, 15, -4], [27, 55, 30, -10, -15, 6],
                   [1, 5, 10, 10, 5, -4], [0, 0, 0, 0, 0, 1]]) / 120.0
    B6 = np.array([[1, 1, 1, 1, 1, 1], [0, 1, 2, 3, 4, 5], [0, 0, 2, 6, 12, 20], [0, 0, 0, 6, 24, 60], [0, 0, 0, 0, 24, 120], [0, 0, 0, 0, 0, 120]], dtype=float)
    for N, M, Cm, B in ((4, M4, C4, B4), (6, M6, C6, B6)):
        np.testing.assert_allclose(oracle.spline_table(N, 0), M, rtol=0, atol=1e-15)
        np.testing.assert_allclose(oracle.spline_table(N, 1), Cm, rtol=0, atol=1e-15)
        np.testing.assert_allclose(oracle.spline_table(N, 2), B, rtol=0, atol=0)
        assert abs(oracle.spline_table(N, 0)[:, 0].sum() - 1.0) < 1e-15  # partition of unity


def _sophus_fixture(oracle):
    def q_wxyz(w, x, y, z):
        q = np.array([x, y, z, w], dtype=float)
        return q / np.linalg.norm(q)
    e = oracle.so3_exp
    m = oracle.so3_mul
    so3 = [q_wxyz(0.1e-11, 0.0, 1.0, 0.0), q_wxyz(-1, 0.00001, 0.0, 0.0), e([0.2, 0.5, 0.0]), e([0.2, 0.5, -1.0]), e([0, 0, 0]), e([0, 0, 0.00001]),
           e([PI, 0, 0]), m(m(e([0.2, 0.5, 0.0]), e([PI, 0, 0])), e([-0.2, -0.5, -0.0])), m(m(e([0.3, 0.5, 0.1]), e([PI, 0, 0])), e([-0.3, -0.5, -0.1]))]
    tangent = [np.zeros(3), np.array([1.0, 0, 0]), np.array([0, 1.0, 0]), np.array([PI / 2, PI / 2, 0]), np.array([-1.0, 1, 0]),
               np.array([20.0, -1, 0]), np.array([30.0, 5, -1])]
    points = [np.array([1.0, 2, 4]), np.array([1.0, -3, 0.5])]
    return so3, tangent, points


def test_sophus_exp_log(oracle):
    so3, tangent, _ = _sophus_fixture(oracle)
    for q in so3:  # expLogTest: exp(log(g)) == g as rotation matrices
        R1 = oracle.so3_matrix(q)
        R2 = oracle.so3_matrix(oracle.so3_exp(oracle.so3_log(q)))
        np.testing.assert_allclose(R1, R2, atol=1e-9)
    for w in tangent:  # expMapTest: log(exp(w)) gives the same rotation
        q = oracle.so3_exp(w)
        assert abs(np.linalg.norm(q) - 1.0) < 1e-10   # SOPHUS_ENSURE in expAndTheta
        np.testing.assert_allclose(oracle.so3_matrix(oracle.so3_exp(oracle.so3_log(q))), oracle.so3_matrix(q), atol=1e-9)
    # exp against scipy for the non-degenerate tangents
    from scipy.spatial.transform import Rotation as Rot
    for w in tangent[1:]:
        np.testing.assert_allclose(oracle.so3_matrix(oracle.so3_exp(w)), Rot.from_rotvec(w).as_matrix(), atol=1e-12)


def test_sophus_group_action_and_product(oracle):
    so3, _, points = _sophus_fixture(oracle)
    for q in so3:
        R = oracle.so3_matrix(q)
        for p in points:  # groupActionTest
            np.testing.assert_allclose(oracle.so3_act(q, p), R @ p, atol=1e-10)
    for a in so3:
        for b in so3:  # productTest: matrix(a*b) == matrix(a) matrix(b)
            np.testing.assert_allclose(oracle.so3_matrix(oracle.so3_mul(a, b)), oracle.so3_matrix(a) @ oracle.so3_matrix(b), atol=1e-9)


def test_plus_jacobian_known_answer_and_numeric(oracle):
    so3, _, _ = _sophus_fixture(oracle)
    for q in so3:
        x, y, z, w = q
        ref = 0.5 * np.array([[w, -z, y], [z, w, -x], [-y, x, w], [-x, -y, -z]])
        J = oracle.so3_plus_jacobian(q)
        np.testing.assert_allclose(J, ref, atol=1e-15)
        # derivativeTest: numeric derivative of q*exp(x) at 0
        h = 1e-6
        Jn = np.zeros((4, 3))
        for k in range(3):
            d = np.zeros(3); d[k] = h
            Jn[:, k] = (oracle.so3_mul(q, oracle.so3_exp(d)) - oracle.so3_mul(q, oracle.so3_exp(-d))) / (2 * h)
        np.testing.assert_allclose(J, Jn, atol=1e-8)


def test_log_small_angle_and_negative_w_branches(oracle):
    # n^2 < eps^2 -> series branch (so3.hpp:262-270)
    q = np.array([1e-12, 0, 0, 1.0]); q /= np.linalg.norm(q)
    np.testing.assert_allclose(oracle.so3_log(q), [2e-12, 0, 0], rtol=1e-12)
    # w < 0: atan(n/w) gives the negative-angle branch (kept)
    q = np.array([np.sin(0.1), 0, 0, -np.cos(0.1)])
    np.testing.assert_allclose(oracle.so3_log(q), [-0.2, 0, 0], rtol=1e-12)


@pytest.mark.parametrize("N", [4, 6])
def test_oracle_jacobians_vs_central_differences(oracle, pkg, scene, N):
    """The Jet-autodiff oracle against central differences of its own residuals (independent of the Jet code path)."""
    sc = scene.make_scene(order=N, duration_s=1.5, n_landmarks=40, track_len=6, imu_hz=50, gp_hz=10, dt_s_bias=0.1, seed=7,
                          flags=(True, True, True, True, True))
    o = oracle.OracleSpline(N, sc["dt_ns"], sc["start_t_ns"])
    pkg.load_scene(o, sc)
    q0, p0 = o.getKnots()
    for kind in (0, 1, 2, 3):
        r0, J, dropped = o.evaluate(kind)
        assert dropped == 0.0
        # perturb position knot 3 (Euclidean) and check the matching column of the first block that sees it
        h = 1e-6
        cols_checked = 0
        for axis in range(3):
            pp = p0.copy(); pp[3, axis] += h; o.setKnots(q0, pp); rp, _, _ = o.evaluate(kind, jac=False)
            pm = p0.copy(); pm[3, axis] -= h; o.setKnots(q0, pm); rm, _, _ = o.evaluate(kind, jac=False)
            o.setKnots(q0, p0)
            num = (rp - rm) / (2 * h)
            if kind == 2:
                assert np.abs(num).max() == 0.0  # gyro does not see positions
                continue
            # blocks whose window starts at s see knot 3 at local index 3-s; compare on blocks with s == 0..3
            nz = np.nonzero(np.abs(num) > 0)[0]
            assert nz.size > 0
            for ridx in nz[:6]:
                row = J[ridx]
                posblock = row[3 * N:6 * N].reshape(N, 3)[:, axis]
                assert np.min(np.abs(posblock - num[ridx])) < 1e-6 * max(1.0, abs(num[ridx]))
                cols_checked += 1
        if kind != 2:
            assert cols_checked > 0
