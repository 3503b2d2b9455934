// ctba_dt_math.h — device math of the discrete-time estimator (BASELINE configs[3]).
//
// What each function replaces in the reference:
//   dt_se3_plus / dt_se3_plus_jacobian   LieLocalParameterization<Sophus::SE3d>: include/optimization/lie_local_parametrization.h:34-55
//                                        -> T * exp(delta) (se3.hpp:761-782) and Dx_this_mul_exp_x_at_0 (se3.hpp:135-204)
//   dt_reproj_frame / dt_reproj_point    ReprojectionCostFunctor (discrete): include/optimization/discrete_time/reprojection_cost_functor.h:36-116
//   dt_gp_eval                           GlobalPositionSensorCostFunctor (discrete): include/optimization/discrete_time/global_position_cost_functor.h:24-51
//   dt_imu_redo / dt_imu_eval            ImuCostFunctor: include/optimization/discrete_time/imu_cost_functor.h:377-594 (redoPreintegration),
//                                        :667-856 (EvaluateWithMinimalJacobians), liftJacobian pose_local_parameterization.h:20-31,
//                                        helpers util/matrix_operations.h:17-80, sinc sensor/imu.h:24-40
// Unknowns per pose are the SE3 tangent [upsilon, omega] of the right perturbation T * exp(delta); the reprojection and global-position
// Jacobians are written analytically in that tangent (what Ceres obtains as autodiff Jacobian x plus-Jacobian); the IMU factor hands
// Ceres its own 15 x 7 blocks built with liftJacobian on the OKVIS layout, which Ceres then multiplies by the Sophus plus-Jacobian: that
// product is reproduced as it is (DESIGN.md "discrete-time estimator").
#pragma once
#include "ctba_math.h"

namespace ctba {

// ---- SE3 on the Sophus layout [qx qy qz qw tx ty tz] ----
CTBA_HD void dt_se3_plus(const double* T, const double* d, double* out) {
  const v3 ups{d[0], d[1], d[2]}, om{d[3], d[4], d[5]};
  const double th2 = dot(om, om), theta = sqrt(th2);
  const quat e = so3_exp(om);
  v3 te;
  if (theta < kSophusEps) {
    te = mul(quat_matrix(e), ups);
  } else {
    const double a = (1.0 - cos(theta)) / th2, b = (theta - sin(theta)) / (th2 * theta);
    const v3 c1 = cross(om, ups), c2 = cross(om, c1);
    te = ups + a * c1 + b * c2;
  }
  const quat q{T[0], T[1], T[2], T[3]};
  const quat r = quat_mul(q, e);
  const v3 tr = quat_act(q, te);
  out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
  out[4] = T[4] + tr.x; out[5] = T[5] + tr.y; out[6] = T[6] + tr.z;
}
// rows = parameters [qx qy qz qw tx ty tz], columns = tangent [upsilon, omega]; J is 7 x 6 row-major
CTBA_HD void dt_se3_plus_jacobian(const double* T, double* J) {
  const double qx = T[0], qy = T[1], qz = T[2], qw = T[3];
  for (int i = 0; i < 42; ++i) J[i] = 0.0;
  const double hw = 0.5 * qw, hx = 0.5 * qx, hy = 0.5 * qy, hz = 0.5 * qz;
  J[0 * 6 + 3] = hw;  J[0 * 6 + 4] = -hz; J[0 * 6 + 5] = hy;
  J[1 * 6 + 3] = hz;  J[1 * 6 + 4] = hw;  J[1 * 6 + 5] = -hx;
  J[2 * 6 + 3] = -hy; J[2 * 6 + 4] = hx;  J[2 * 6 + 5] = hw;
  J[3 * 6 + 3] = -hx; J[3 * 6 + 4] = -hy; J[3 * 6 + 5] = -hz;
  const double ww = qw * qw, xx = qx * qx, yy = qy * qy, zz = qz * qz;
  const double wz = 2.0 * qw * qz, xy = 2.0 * qx * qy, wy = 2.0 * qw * qy, xz = 2.0 * qx * qz, wx = 2.0 * qw * qx, yz = 2.0 * qy * qz;
  J[4 * 6 + 0] = -yy - zz + ww + xx; J[4 * 6 + 1] = -wz + xy;           J[4 * 6 + 2] = wy + xz;
  J[5 * 6 + 0] = wz + xy;            J[5 * 6 + 1] = -zz + (ww - xx) + yy; J[5 * 6 + 2] = -wx + yz;
  J[6 * 6 + 0] = -wy + xz;           J[6 * 6 + 1] = wx + yz;            J[6 * 6 + 2] = -yy + zz + (ww - xx);
}

// ---- reprojection (discrete time) ----
// scalar problem data of the discrete-time estimator, passed to the kernels by value
struct DtConst {
  Camera cam;
  double width, height, t_offset_cam_imu;
  double scale;
  double q_bc0[4], p_bc0[3];
  double t_offset_gs_imu;
  double gyro_noise_density, acc_noise_density, gyro_random_walk, acc_random_walk, g_max, a_max, g_W[3];
  double cauchy_a;
};
// device layout of the global unknowns (13 doubles), the order of DiscreteTimeEstimator's members (estimator.h:767-790)
enum { DG_SCALE = 0, DG_RBC = 1, DG_PBC = 5, DG_TCAM = 8, DG_TGS = 9, DG_PBP = 10, DG_SIZE = 13 };
// tangent index of each global inside the border (<= GT_SIZE = 15 rows)
enum { DGT_SCALE = 0, DGT_RBC = 1, DGT_PBC = 4, DGT_TCAM = 7, DGT_TGS = 8, DGT_PBP = 9, DGT_COUNT = 12 };

// frame-level part of the reprojection: everything that does not depend on the landmark
struct DtFrameCam {
  m3 R_cb;      // R_BC^T
  m3 R_cw;      // R_WC^T
  m3 R_wb_t;    // R_WB^T
  m3 R_corr_t;  // R_BC_corr^T
  v3 t_wb, sp_bc /* s * p_BC */, p_bc;
  double s, t_off;
};
CTBA_HD void dt_reproj_frame(const DtConst& K, const double* pose, const double* glb, DtFrameCam* fc) {
  const quat q_wb{pose[0], pose[1], pose[2], pose[3]};
  const m3 R_wb = quat_matrix(q_wb);
  const m3 R_bc0 = quat_matrix(quat{K.q_bc0[0], K.q_bc0[1], K.q_bc0[2], K.q_bc0[3]});
  const m3 R_corr = quat_matrix(quat{glb[DG_RBC], glb[DG_RBC + 1], glb[DG_RBC + 2], glb[DG_RBC + 3]});
  const m3 R_bc = mul(R_bc0, R_corr);
  fc->p_bc = mul(R_bc0, v3{glb[DG_PBC], glb[DG_PBC + 1], glb[DG_PBC + 2]}) + v3{K.p_bc0[0], K.p_bc0[1], K.p_bc0[2]};
  fc->s = K.scale + glb[DG_SCALE];
  fc->sp_bc = fc->s * fc->p_bc;
  fc->R_cb = transpose(R_bc);
  fc->R_wb_t = transpose(R_wb);
  fc->R_cw = mul(fc->R_cb, fc->R_wb_t);
  fc->R_corr_t = transpose(R_corr);
  fc->t_wb = v3{pose[4], pose[5], pose[6]};
  fc->t_off = K.t_offset_cam_imu + glb[DG_TCAM];
}
// one valid reprojection: r (2) and, if JAC, J (2 x 17 row-major: pose 6 | scale 1 | R_bc 3 | p_bc 3 | t_cam_imu 1 | landmark 3).
// Gates as the reference: shifted feature outside the image or reprojection error above 10 px -> zero residual, zero Jacobian.
template <bool JAC>
CTBA_HD void dt_reproj_point(const DtConst& K, const DtFrameCam& fc, v3 p_w, const double* uv, const double* vel, double inv_std, double* r, double* J) {
  const v3 q_b = mul(fc.R_wb_t, p_w - fc.t_wb);
  const v3 p_c = mul(fc.R_cb, q_b - fc.sp_bc);
  double proj[2], Jp[6];
  camera_project<JAC>(K.cam, p_c, proj, Jp);
  const double u = uv[0] + vel[0] * fc.t_off, v = uv[1] + vel[1] * fc.t_off;
  const double eu = proj[0] - u, ev = proj[1] - v;
  const bool gated = (u < 0.0) || (u > K.width) || (v < 0.0) || (v > K.height) || (eu * eu + ev * ev > 100.0);
  if (gated) {
    r[0] = r[1] = 0.0;
    if (JAC) for (int i = 0; i < 34; ++i) J[i] = 0.0;
    return;
  }
  r[0] = inv_std * eu; r[1] = inv_std * ev;
  if (!JAC) return;
  // d p_c / d(upsilon) = -R_cb, d p_c / d(omega) = R_cb [q_b]x, d p_c / d s = -R_cb p_bc, d p_c / d theta_corr = [p_c]x,
  // d p_c / d p_corr = -s R_corr^T, d p_c / d p_w = R_cw ; d e / d t_corr = -vel
  const m3 Qx = hat(q_b), Px = hat(p_c);
  const m3 dW = mul(fc.R_cb, Qx);
  const v3 ds = -1.0 * mul(fc.R_cb, fc.p_bc);
#pragma unroll
  for (int row = 0; row < 2; ++row) {
    const double a = inv_std * Jp[3 * row], b = inv_std * Jp[3 * row + 1], c = inv_std * Jp[3 * row + 2];
    double* o = J + 17 * row;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      o[k] = -(a * fc.R_cb.a[k] + b * fc.R_cb.a[3 + k] + c * fc.R_cb.a[6 + k]);
      o[3 + k] = a * dW.a[k] + b * dW.a[3 + k] + c * dW.a[6 + k];
      o[7 + k] = a * Px.a[k] + b * Px.a[3 + k] + c * Px.a[6 + k];
      o[10 + k] = -fc.s * (a * fc.R_corr_t.a[k] + b * fc.R_corr_t.a[3 + k] + c * fc.R_corr_t.a[6 + k]);
      o[14 + k] = a * fc.R_cw.a[k] + b * fc.R_cw.a[3 + k] + c * fc.R_cw.a[6 + k];
    }
    o[6] = a * ds.x + b * ds.y + c * ds.z;
    o[13] = -inv_std * vel[row];
  }
}

// ---- global position (discrete time): linear interpolation between the two states around the measurement ----
// r (3), J (3 x 17 row-major: pose0 6 | pose1 6 | scale 1 | t_gs 1 | p_bp 3)
template <bool JAC>
CTBA_HD void dt_gp_eval(const DtConst& K, const double* pose0, const double* pose1, const double* glb, double t_meas, const double* z, double t_k0, double t_k1,
                        double inv_std, double* r, double* J) {
  const m3 R0 = quat_matrix(quat{pose0[0], pose0[1], pose0[2], pose0[3]}), R1 = quat_matrix(quat{pose1[0], pose1[1], pose1[2], pose1[3]});
  const double s = K.scale + glb[DG_SCALE];
  const v3 p_bp{glb[DG_PBP], glb[DG_PBP + 1], glb[DG_PBP + 2]};
  const double lambda = ((t_meas + (K.t_offset_gs_imu + glb[DG_TGS])) - t_k0) / (t_k1 - t_k0), one_m = 1.0 - lambda;
  const v3 Rp0 = mul(R0, p_bp), Rp1 = mul(R1, p_bp);
  const v3 p0 = s * Rp0 + v3{pose0[4], pose0[5], pose0[6]}, p1 = s * Rp1 + v3{pose1[4], pose1[5], pose1[6]};
  const v3 pi = one_m * p0 + lambda * p1;
  r[0] = inv_std * (pi.x - z[0]); r[1] = inv_std * (pi.y - z[1]); r[2] = inv_std * (pi.z - z[2]);
  if (!JAC) return;
  const m3 Px = hat(p_bp);
  const m3 A0 = mul(R0, Px), A1 = mul(R1, Px);
  const double idt = 1.0 / (t_k1 - t_k0);
  const v3 dsc = one_m * Rp0 + lambda * Rp1, dt = idt * (p1 - p0);
  const double dscv[3] = {dsc.x, dsc.y, dsc.z}, dtv[3] = {dt.x, dt.y, dt.z};
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    double* o = J + 17 * i;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      o[k] = inv_std * one_m * R0.a[3 * i + k];
      o[3 + k] = -inv_std * one_m * s * A0.a[3 * i + k];
      o[6 + k] = inv_std * lambda * R1.a[3 * i + k];
      o[9 + k] = -inv_std * lambda * s * A1.a[3 * i + k];
      o[14 + k] = inv_std * s * (one_m * R0.a[3 * i + k] + lambda * R1.a[3 * i + k]);
    }
    o[12] = inv_std * dscv[i];
    o[13] = inv_std * dtv[i];
  }
}

// ---- IMU pre-integration factor ----
// Per factor cache (the reference's mutable members, imu_cost_functor.h:893-935): written by the redo, read by every evaluation.
struct DtImuCache {
  double Delta_q[4];
  double C_integral[9], C_doubleintegral[9], acc_integral[3], acc_doubleintegral[3];
  double dalpha_db_g[9], dv_db_g[9], dp_db_g[9];
  double sqrt_info[225];
  double sb_ref[9];
  int redo, redo_counter;
};

namespace dtm {
// Eigen::Quaterniond algebra (no normalisation), coefficient order x y z w
CTBA_HD quat qmul(quat a, quat b) {
  return {a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x,
          a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}
CTBA_HD quat qinv(quat q) { const double n2 = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w; return {-q.x / n2, -q.y / n2, -q.z / n2, q.w / n2}; }
CTBA_HD double sinc(double x) {
  if (fabs(x) > 1e-6) return sin(x) / x;
  const double x2 = x * x, x4 = x2 * x2, x6 = x2 * x2 * x2;
  return 1.0 - (1.0 / 6.0) * x2 + (1.0 / 120.0) * x4 - (1.0 / 5040.0) * x6;
}
CTBA_HD quat deltaQ(v3 a) {
  const double halfnorm = 0.5 * sqrt(dot(a, a)), sc = sinc(halfnorm) * 0.5;
  return {sc * a.x, sc * a.y, sc * a.z, cos(halfnorm)};
}
CTBA_HD m3 expmap_derivative(v3 om) {   // util/matrix_operations.h:36-51
  const double theta2 = dot(om, om);
  m3 J = identity3();
  if (theta2 <= 2.220446049250313e-16) return J;
  const double theta = sqrt(theta2);
  m3 Y = hat(om);
  for (int i = 0; i < 9; ++i) Y.a[i] /= theta;
  const m3 Y2 = mul(Y, Y);
  const double a = (1.0 - cos(theta)) / theta, b = 1.0 - sin(theta) / theta;
  for (int i = 0; i < 9; ++i) J.a[i] = J.a[i] - a * Y.a[i] + b * Y2.a[i];
  return J;
}
// top-left 3 x 3 of products of the 4 x 4 quaternion matrices: Q+(q) (left product matrix) and Q(+)(q) (right product matrix)
CTBA_HD void qplus(quat q, double* Q) {   // 4 x 4 row-major, matrix_operations.h:58-68
  const double c[4] = {q.x, q.y, q.z, q.w};
  Q[0] = c[3];  Q[1] = -c[2]; Q[2] = c[1];  Q[3] = c[0];
  Q[4] = c[2];  Q[5] = c[3];  Q[6] = -c[0]; Q[7] = c[1];
  Q[8] = -c[1]; Q[9] = c[0];  Q[10] = c[3]; Q[11] = c[2];
  Q[12] = -c[0]; Q[13] = -c[1]; Q[14] = -c[2]; Q[15] = c[3];
}
CTBA_HD void qoplus(quat q, double* Q) {  // matrix_operations.h:71-80
  const double c[4] = {q.x, q.y, q.z, q.w};
  Q[0] = c[3];  Q[1] = c[2];  Q[2] = -c[1]; Q[3] = c[0];
  Q[4] = -c[2]; Q[5] = c[3];  Q[6] = c[0];  Q[7] = c[1];
  Q[8] = c[1];  Q[9] = -c[0]; Q[10] = c[3]; Q[11] = c[2];
  Q[12] = -c[0]; Q[13] = -c[1]; Q[14] = -c[2]; Q[15] = c[3];
}
CTBA_HD void m4mul(const double* A, const double* B, double* C) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += A[4 * i + k] * B[4 * k + j]; C[4 * i + j] = s; }
}
}  // namespace dtm

// redoPreintegration: integrates the samples [t0, t1] at the speed-and-bias sb (v, b_g, b_a) into the cache; scratch = 3 x 225 doubles.
// Returns the number of integration steps, -1 if the samples do not reach t1, -2 if the covariance cannot be inverted.
CTBA_HD int dt_imu_redo(const DtConst& K, DtImuCache* Cc, const double* ts, const double* gyr, const double* acc, int n, double t0, double t1, const double* sb,
                        double* scratch) {
  using namespace dtm;
  if (!(ts[n - 1] >= t1)) return -1;
  double* P = scratch;          // 15 x 15 covariance
  double* F = scratch + 225;    // 15 x 15 transition, later the inverse work area
  double* W2 = scratch + 450;   // 15 x 15 product
  quat Dq{0, 0, 0, 1};
  m3 C_int = zero3(), C_dint = zero3(), crossm = zero3(), dalpha = zero3(), dv = zero3(), dp = zero3();
  v3 a_int{0, 0, 0}, a_dint{0, 0, 0};
  for (int i = 0; i < 225; ++i) P[i] = 0.0;
  double time = t0;
  bool has_started = false, last_iteration = false;
  int n_integrated = 0;
  for (int i = 0; i + 1 < n; ++i) {
    v3 om0{gyr[3 * i], gyr[3 * i + 1], gyr[3 * i + 2]}, ac0{acc[3 * i], acc[3 * i + 1], acc[3 * i + 2]};
    v3 om1{gyr[3 * i + 3], gyr[3 * i + 4], gyr[3 * i + 5]}, ac1{acc[3 * i + 3], acc[3 * i + 4], acc[3 * i + 5]};
    double nexttime = ts[i + 1];
    double dt = nexttime - time;
    if (t1 < nexttime) {
      const double interval = nexttime - ts[i];
      nexttime = t1;
      last_iteration = true;
      dt = nexttime - time;
      const double r = dt / interval;
      om1 = (1.0 - r) * om0 + r * om1; ac1 = (1.0 - r) * ac0 + r * ac1;
    }
    if (dt <= 0.0) continue;
    if (!has_started) {
      has_started = true;
      const double r = dt / (nexttime - ts[i]);
      om0 = r * om0 + (1.0 - r) * om1; ac0 = r * ac0 + (1.0 - r) * ac1;
    }
    double sigma_g_c = K.gyro_noise_density, sigma_a_c = K.acc_noise_density;
    const bool gsat = fabs(om0.x) > K.g_max || fabs(om0.y) > K.g_max || fabs(om0.z) > K.g_max || fabs(om1.x) > K.g_max || fabs(om1.y) > K.g_max || fabs(om1.z) > K.g_max;
    const bool asat = fabs(ac0.x) > K.a_max || fabs(ac0.y) > K.a_max || fabs(ac0.z) > K.a_max || fabs(ac1.x) > K.a_max || fabs(ac1.y) > K.a_max || fabs(ac1.z) > K.a_max;
    if (gsat) sigma_g_c *= 100;
    if (asat) sigma_a_c *= 100;
    const v3 om_true = 0.5 * (om0 + om1) - v3{sb[3], sb[4], sb[5]}, ac_true = 0.5 * (ac0 + ac1) - v3{sb[6], sb[7], sb[8]};
    const double theta_half = sqrt(dot(om_true, om_true)) * 0.5 * dt;
    const double sinc_th = sinc(theta_half), cos_th = cos(theta_half);
    const quat dq{sinc_th * om_true.x * 0.5 * dt, sinc_th * om_true.y * 0.5 * dt, sinc_th * om_true.z * 0.5 * dt, cos_th};
    const quat Dq1 = qmul(Dq, dq);
    const m3 C = quat_matrix(Dq), C1 = quat_matrix(Dq1);
    m3 CC; for (int k = 0; k < 9; ++k) CC.a[k] = C.a[k] + C1.a[k];
    const v3 CCa = mul(CC, ac_true);
    m3 C_int1; for (int k = 0; k < 9; ++k) C_int1.a[k] = C_int.a[k] + 0.5 * CC.a[k] * dt;
    const v3 a_int1 = a_int + (0.5 * dt) * CCa;   // 0.5 * CCa * dt
    for (int k = 0; k < 9; ++k) C_dint.a[k] += C_int.a[k] * dt + 0.25 * CC.a[k] * dt * dt;
    const v3 adi_step = dt * a_int + (0.25 * dt * dt) * CCa;
    a_dint = a_dint + adi_step;
    const m3 Jr = expmap_derivative(dt * om_true);
    m3 T1 = mul(C1, Jr);
    for (int k = 0; k < 9; ++k) dalpha.a[k] += T1.a[k] * dt;
    const m3 Rdqi = quat_matrix(qinv(dq));
    T1 = mul(Rdqi, crossm);
    m3 cross1; for (int k = 0; k < 9; ++k) cross1.a[k] = T1.a[k] + Jr.a[k] * dt;
    const m3 acx = hat(ac_true);
    const m3 A0 = mul(mul(C, acx), crossm), A1 = mul(mul(C1, acx), cross1);
    m3 Wm; for (int k = 0; k < 9; ++k) Wm.a[k] = A0.a[k] + A1.a[k];
    m3 dv1; for (int k = 0; k < 9; ++k) dv1.a[k] = dv.a[k] + 0.5 * dt * Wm.a[k];
    // covariance propagation (dv, C_int, a_int BEFORE their shift)
    for (int k = 0; k < 225; ++k) F[k] = 0.0;
    for (int k = 0; k < 15; ++k) F[16 * k] = 1.0;
    const m3 S1 = hat(adi_step), S2 = hat((0.5 * dt) * CCa);
    for (int a = 0; a < 3; ++a)
      for (int b = 0; b < 3; ++b) {
        const int ab = 3 * a + b;
        F[15 * a + 3 + b] = -S1.a[ab];
        F[15 * a + 6 + b] = (a == b ? 1.0 : 0.0) * dt;
        F[15 * a + 9 + b] = dt * dv.a[ab] + 0.25 * dt * dt * Wm.a[ab];
        F[15 * a + 12 + b] = -C_int.a[ab] * dt + 0.25 * CC.a[ab] * dt * dt;
        F[15 * (3 + a) + 9 + b] = -dt * C1.a[ab];
        F[15 * (6 + a) + 3 + b] = -S2.a[ab];
        F[15 * (6 + a) + 9 + b] = 0.5 * dt * Wm.a[ab];
        F[15 * (6 + a) + 12 + b] = -0.5 * CC.a[ab] * dt;
      }
    for (int k = 0; k < 9; ++k) dp.a[k] += dt * dv.a[k] + 0.25 * dt * dt * Wm.a[k];
#pragma unroll 1
    for (int a = 0; a < 15; ++a)
#pragma unroll 1
      for (int b = 0; b < 15; ++b) { double s = 0; for (int k = 0; k < 15; ++k) s += F[15 * a + k] * P[15 * k + b]; W2[15 * a + b] = s; }
#pragma unroll 1
    for (int a = 0; a < 15; ++a)
#pragma unroll 1
      for (int b = 0; b < 15; ++b) { double s = 0; for (int k = 0; k < 15; ++k) s += W2[15 * a + k] * F[15 * b + k]; P[15 * a + b] = s; }
    const double sigma2_dalpha = dt * sigma_g_c * sigma_g_c, sigma2_v = dt * sigma_a_c * sigma_a_c, sigma2_p = 0.5 * dt * dt * sigma2_v;
    const double sigma2_b_g = dt * K.gyro_random_walk * K.gyro_random_walk, sigma2_b_a = dt * K.acc_random_walk * K.acc_random_walk;
    for (int a = 0; a < 3; ++a) {
      P[16 * (3 + a)] += sigma2_dalpha; P[16 * (6 + a)] += sigma2_v; P[16 * a] += sigma2_p; P[16 * (9 + a)] += sigma2_b_g; P[16 * (12 + a)] += sigma2_b_a;
    }
    Dq = Dq1; C_int = C_int1; a_int = a_int1; crossm = cross1; dv = dv1;
    time = nexttime;
    ++n_integrated;
    if (last_iteration) break;
  }
  Cc->Delta_q[0] = Dq.x; Cc->Delta_q[1] = Dq.y; Cc->Delta_q[2] = Dq.z; Cc->Delta_q[3] = Dq.w;
  for (int k = 0; k < 9; ++k) {
    Cc->C_integral[k] = C_int.a[k]; Cc->C_doubleintegral[k] = C_dint.a[k]; Cc->dalpha_db_g[k] = dalpha.a[k]; Cc->dv_db_g[k] = dv.a[k]; Cc->dp_db_g[k] = dp.a[k];
    Cc->sb_ref[k] = sb[k];
  }
  Cc->acc_integral[0] = a_int.x; Cc->acc_integral[1] = a_int.y; Cc->acc_integral[2] = a_int.z;
  Cc->acc_doubleintegral[0] = a_dint.x; Cc->acc_doubleintegral[1] = a_dint.y; Cc->acc_doubleintegral[2] = a_dint.z;
  // information = inverse of the symmetrised covariance (LU with partial pivoting, as Eigen's fixed-size inverse), symmetrised again;
  // square-root information = L^T of its Cholesky factor
  for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) F[15 * a + b] = 0.5 * P[15 * a + b] + 0.5 * P[15 * b + a];
  double* X = W2;
  for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) X[15 * a + b] = a == b ? 1.0 : 0.0;
#pragma unroll 1
  for (int c = 0; c < 15; ++c) {
    int piv = c; double best = fabs(F[16 * c]);
    for (int r = c + 1; r < 15; ++r) if (fabs(F[15 * r + c]) > best) { best = fabs(F[15 * r + c]); piv = r; }
    if (best == 0.0) return -2;
    if (piv != c)
      for (int j = 0; j < 15; ++j) {
        double t = F[15 * c + j]; F[15 * c + j] = F[15 * piv + j]; F[15 * piv + j] = t;
        t = X[15 * c + j]; X[15 * c + j] = X[15 * piv + j]; X[15 * piv + j] = t;
      }
    for (int r = c + 1; r < 15; ++r) {
      const double f = F[15 * r + c] / F[16 * c];
      if (f != 0.0) {
        for (int j = c; j < 15; ++j) F[15 * r + j] -= f * F[15 * c + j];
        for (int j = 0; j < 15; ++j) X[15 * r + j] -= f * X[15 * c + j];
      }
    }
  }
#pragma unroll 1
  for (int c = 14; c >= 0; --c)
    for (int j = 0; j < 15; ++j) { double s = X[15 * c + j]; for (int k = c + 1; k < 15; ++k) s -= F[15 * c + k] * X[15 * k + j]; X[15 * c + j] = s / F[16 * c]; }
  for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) P[15 * a + b] = 0.5 * X[15 * a + b] + 0.5 * X[15 * b + a];
  double* L = F;
  for (int k = 0; k < 225; ++k) L[k] = 0.0;
#pragma unroll 1
  for (int j = 0; j < 15; ++j) {
    double d = P[16 * j];
    for (int k = 0; k < j; ++k) d -= L[15 * j + k] * L[15 * j + k];
    d = sqrt(d);
    L[16 * j] = d;
    for (int i = j + 1; i < 15; ++i) { double s = P[15 * i + j]; for (int k = 0; k < j; ++k) s -= L[15 * i + k] * L[15 * j + k]; L[15 * i + j] = s / d; }
  }
  for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) Cc->sqrt_info[15 * a + b] = L[15 * b + a];
  return n_integrated;
}

// One evaluation of the factor: weighted residual (15) and, if J != nullptr, the LOCAL Jacobian 15 x 30 row-major
// [pose k 6 | speed-and-bias k 9 | pose k+1 6 | speed-and-bias k+1 9] = (reference's 15 x 7 block) x (SE3 plus-Jacobian).  The cache is
// redone first when the gyro bias moved away from its linearisation point (imu_cost_functor.h:693-704).  scratch = 3 x 225 doubles.
CTBA_HD int dt_imu_eval(const DtConst& K, DtImuCache* Cc, const double* ts, const double* gyr, const double* acc, int n, double t0, double t1, const double* T0,
                        const double* sb0, const double* T1p, const double* sb1, double* r, double* J, double* scratch) {
  using namespace dtm;
  const double delta_t = t1 - t0;
  double Delta_b[6];
  for (int c = 0; c < 6; ++c) Delta_b[c] = sb0[3 + c] - Cc->sb_ref[3 + c];
  const bool redo = Cc->redo || (sqrt(Delta_b[0] * Delta_b[0] + Delta_b[1] * Delta_b[1] + Delta_b[2] * Delta_b[2]) * delta_t > 0.0001);
  int rc = 0;
  if (redo) {
    rc = dt_imu_redo(K, Cc, ts, gyr, acc, n, t0, t1, sb0, scratch);
    Cc->redo_counter++;
    for (int c = 0; c < 6; ++c) Delta_b[c] = 0.0;
    Cc->redo = 0;
  }
  const quat q0{T0[0], T0[1], T0[2], T0[3]}, q1{T1p[0], T1p[1], T1p[2], T1p[3]};
  const m3 C_B0_W = transpose(quat_matrix(q0));
  double* F0 = scratch;        // 15 x 15
  double* F1 = scratch + 225;  // 15 x 15
  for (int k = 0; k < 225; ++k) { F0[k] = 0.0; F1[k] = 0.0; }
  for (int k = 0; k < 15; ++k) { F0[16 * k] = 1.0; F1[16 * k] = -1.0; }
  const v3 gW{K.g_W[0], K.g_W[1], K.g_W[2]};
  const v3 v0{sb0[0], sb0[1], sb0[2]}, v1{sb1[0], sb1[1], sb1[2]};
  const v3 dp_est = v3{T0[4] - T1p[4], T0[5] - T1p[5], T0[6] - T1p[6]} + delta_t * v0 - (0.5 * delta_t * delta_t) * gW;
  const v3 dv_est = v0 - v1 - delta_t * gW;
  m3 dalpha, dv, dp, C_int, C_dint;
  for (int k = 0; k < 9; ++k) { dalpha.a[k] = Cc->dalpha_db_g[k]; dv.a[k] = Cc->dv_db_g[k]; dp.a[k] = Cc->dp_db_g[k]; C_int.a[k] = Cc->C_integral[k]; C_dint.a[k] = Cc->C_doubleintegral[k]; }
  const v3 dbg{Delta_b[0], Delta_b[1], Delta_b[2]};
  const quat Dq = qmul(deltaQ(-1.0 * mul(dalpha, dbg)), quat{Cc->Delta_q[0], Cc->Delta_q[1], Cc->Delta_q[2], Cc->Delta_q[3]});
  const m3 TSp = mul(C_B0_W, hat(dp_est)), TSv = mul(C_B0_W, hat(dv_est));
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) {
      const int ab = 3 * a + b;
      F0[15 * a + b] = C_B0_W.a[ab]; F0[15 * a + 3 + b] = TSp.a[ab]; F0[15 * a + 6 + b] = C_B0_W.a[ab] * delta_t; F0[15 * a + 9 + b] = dp.a[ab]; F0[15 * a + 12 + b] = -C_dint.a[ab];
      F0[15 * (6 + a) + 3 + b] = TSv.a[ab]; F0[15 * (6 + a) + 6 + b] = C_B0_W.a[ab]; F0[15 * (6 + a) + 9 + b] = dv.a[ab]; F0[15 * (6 + a) + 12 + b] = -C_int.a[ab];
      F1[15 * a + b] = -C_B0_W.a[ab]; F1[15 * (6 + a) + 6 + b] = -C_B0_W.a[ab];
    }
  double QA[16], QB[16], QC[16], QD[16];
  const quat q1i = qinv(q1);
  qplus(qmul(Dq, q1i), QA); qoplus(q0, QB); m4mul(QA, QB, QC);
  for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F0[15 * (3 + a) + 3 + b] = QC[4 * a + b];
  qoplus(qmul(q1i, q0), QA); qoplus(Dq, QB); m4mul(QA, QB, QC);
  for (int a = 0; a < 3; ++a)
    for (int b = 0; b < 3; ++b) { double s = 0; for (int k = 0; k < 3; ++k) s += QC[4 * a + k] * (-dalpha.a[3 * k + b]); F0[15 * (3 + a) + 9 + b] = s; }
  qplus(Dq, QA); qoplus(q0, QB); m4mul(QA, QB, QC); qplus(q1i, QD); m4mul(QC, QD, QA);
  for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F1[15 * (3 + a) + 3 + b] = -QA[4 * a + b];
  double error[15];
  const v3 e_p = mul(C_B0_W, dp_est), e_v = mul(C_B0_W, dv_est);
  const double epv[3] = {e_p.x, e_p.y, e_p.z}, evv[3] = {e_v.x, e_v.y, e_v.z};
  for (int a = 0; a < 3; ++a) {
    double s = 0, s2 = 0;
    for (int k = 0; k < 6; ++k) { s += F0[15 * a + 9 + k] * Delta_b[k]; s2 += F0[15 * (6 + a) + 9 + k] * Delta_b[k]; }
    error[a] = epv[a] + Cc->acc_doubleintegral[a] + s;
    error[6 + a] = evv[a] + Cc->acc_integral[a] + s2;
  }
  const quat qe = qmul(Dq, qmul(q1i, q0));
  error[3] = 2.0 * qe.x; error[4] = 2.0 * qe.y; error[5] = 2.0 * qe.z;
  for (int c = 0; c < 6; ++c) error[9 + c] = sb0[3 + c] - sb1[3 + c];
  const double* S = Cc->sqrt_info;
  for (int a = 0; a < 15; ++a) { double s = 0; for (int k = 0; k < 15; ++k) s += S[15 * a + k] * error[k]; r[a] = s; }
  if (!J) return rc;
  // pose blocks: (sqrt_info F)[:, :6] * liftJacobian(x) (6 x 7), then * plusJacobian(x) (7 x 6)
  double* SF = scratch + 450;   // 15 x 15
  for (int side = 0; side < 2; ++side) {
    const double* Fm = side ? F1 : F0;
    const double* x = side ? T1p : T0;
#pragma unroll 1
    for (int a = 0; a < 15; ++a)
#pragma unroll 1
      for (int b = 0; b < 15; ++b) { double s = 0; for (int k = 0; k < 15; ++k) s += S[15 * a + k] * Fm[15 * k + b]; SF[15 * a + b] = s; }
    double Lm[42], PJ[42], Qp[16];
    for (int k = 0; k < 42; ++k) Lm[k] = 0.0;
    for (int a = 0; a < 3; ++a) Lm[7 * a + a] = 1.0;
    qoplus(quat{-x[3], -x[4], -x[5], x[6]}, Qp);   // reads x[3..6] as a quaternion: the OKVIS layout
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 4; ++b) Lm[7 * (3 + a) + 3 + b] = 2.0 * Qp[4 * a + b];
    dt_se3_plus_jacobian(x, PJ);
    for (int a = 0; a < 15; ++a) {
      double* row = J + 30 * a + 15 * side;
      // (SF[:, :6] * Lm) first, then * PJ, like Ceres: J_global (15 x 7) * plus-Jacobian
      double g7[7];
      for (int b = 0; b < 7; ++b) { double s = 0; for (int k = 0; k < 6; ++k) s += SF[15 * a + k] * Lm[7 * k + b]; g7[b] = s; }
      for (int b = 0; b < 6; ++b) { double s = 0; for (int k = 0; k < 7; ++k) s += g7[k] * PJ[6 * k + b]; row[b] = s; }
      for (int b = 0; b < 9; ++b) row[6 + b] = SF[15 * a + 6 + b];
    }
  }
  return rc;
}

}  // namespace ctba
