// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// Discrete-time estimator of the reference (BASELINE configs[3]), restated:
//   DiscreteTimeEstimator                     include/discrete_time_estimator/estimator.h:47-824 (states :88-103, 105-125; problem wiring :149-165,
//                                             373-519; optimize :699-713)
//   ReprojectionCostFunctor (discrete)        include/optimization/discrete_time/reprojection_cost_functor.h:36-116
//   GlobalPositionSensorCostFunctor (discr.)  include/optimization/discrete_time/global_position_cost_functor.h:24-51
//   ImuCostFunctor                            include/optimization/discrete_time/imu_cost_functor.h:36-942 (redoPreintegration :377-594,
//                                             EvaluateWithMinimalJacobians :667-856), liftJacobian pose_local_parameterization.h:20-31,
//                                             helpers util/matrix_operations.h:17-80, sinc sensor/imu.h:24-40
//   LieLocalParameterization<SE3d>            include/optimization/lie_local_parametrization.h:34-55 -> se3.hpp:135-204 (Dx_this_mul_exp_x_at_0),
//                                             se3.hpp:761-782 (exp)
// Parity is on what the reference COMPUTES: the IMU factor hands Ceres 15 x 7 Jacobians built with liftJacobian, which assumes the OKVIS
// parameter layout [t, q] while the Sophus SE3 block is [q, t]; Ceres multiplies them by the SE3 plus-Jacobian.  Both are restated as they are.
#pragma once
#include <cmath>
#include <cstring>
#include <memory>
#include <vector>
#include "lm.h"

namespace oracle {

// ---- SE3 on the Sophus layout [qx qy qz qw tx ty tz] ----
// se3.hpp:761-782 exp, SE3 product: (so3 * so3 (normalised), t + so3 * t_other)
inline void se3_plus(const double* T, const double* d, double* out) {
  const V3<double> ups{d[0], d[1], d[2]}, om{d[3], d[4], d[5]};
  const double theta = std::sqrt(om.x * om.x + om.y * om.y + om.z * om.z);
  SO3<double> e = SO3<double>::exp(om);
  M3<double> Om{}, Om2{}, V{};
  Om.m[0][0] = 0; Om.m[0][1] = -om.z; Om.m[0][2] = om.y; Om.m[1][0] = om.z; Om.m[1][1] = 0; Om.m[1][2] = -om.x; Om.m[2][0] = -om.y; Om.m[2][1] = om.x; Om.m[2][2] = 0;
  Om2 = mul(Om, Om);
  if (theta < kEps) {
    V = e.matrix();
  } else {
    const double th2 = theta * theta, a = (1.0 - std::cos(theta)) / th2, b = (theta - std::sin(theta)) / (th2 * theta);
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) V.m[i][j] = (i == j ? 1.0 : 0.0) + a * Om.m[i][j] + b * Om2.m[i][j];
  }
  const V3<double> te = mul(V, ups);
  SO3<double> q{T[0], T[1], T[2], T[3]};
  SO3<double> r = q * e;
  const V3<double> tr = q.act(te);
  out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
  out[4] = T[4] + tr.x; out[5] = T[5] + tr.y; out[6] = T[6] + tr.z;
}
// se3.hpp:135-204, rows = parameters [qx qy qz qw tx ty tz], columns = tangent [upsilon, omega]
inline void se3_plus_jacobian(const double* T, double J[7][6]) {
  const double qx = T[0], qy = T[1], qz = T[2], qw = T[3];
  const double c0 = 0.5 * qw, c1 = 0.5 * qz, c2 = -c1, c3 = 0.5 * qy, c4 = 0.5 * qx, c5 = -c4, c6 = -c3;
  const double c7 = qw * qw, c8 = qx * qx, c9 = qy * qy, c10 = -c9, c11 = qz * qz, c12 = -c11, c13 = 2.0 * qw, c14 = c13 * qz, c15 = 2.0 * qx, c16 = c15 * qy;
  const double c17 = c13 * qy, c18 = c15 * qz, c19 = c7 - c8, c20 = c13 * qx, c21 = 2.0 * qy * qz;
  for (int i = 0; i < 7; ++i) for (int j = 0; j < 6; ++j) J[i][j] = 0.0;
  J[0][3] = c0; J[0][4] = c2; J[0][5] = c3;
  J[1][3] = c1; J[1][4] = c0; J[1][5] = c5;
  J[2][3] = c6; J[2][4] = c4; J[2][5] = c0;
  J[3][3] = c5; J[3][4] = c6; J[3][5] = c2;
  J[4][0] = c10 + c12 + c7 + c8; J[4][1] = -c14 + c16; J[4][2] = c17 + c18;
  J[5][0] = c14 + c16; J[5][1] = c12 + c19 + c9; J[5][2] = -c20 + c21;
  J[6][0] = -c17 + c18; J[6][1] = c20 + c21; J[6][2] = c10 + c11 + c19;
}

inline void se3_plus_fwd(const double* T, const double* d, double* out) { se3_plus(T, d, out); }
inline void se3_plus_jacobian_fwd(const double* T, double J[7][6]) { se3_plus_jacobian(T, J); }

struct DtCamera { CameraModel cam; double w = 0, h = 0, t_offset_cam_imu = 0; };

// reprojection_cost_functor.h:36-116 (discrete time).  Parameters: T_WB(7), scale_corr(1), R_BC_corr(4), p_BC_corr(3), t_offset_cam_imu_corr(1),
// then one landmark per valid reprojection.
struct DtReprojectionCostFunctor {
  const DtCamera* camera;
  double scale, inv_std;
  int n_reproj;
  const double* uv;    // 2 n
  const double* vel;   // 2 n (Point2D::velocity)
  template <class T>
  bool operator()(T const* const* parameters, T* sResiduals) const {
    SO3<T> q_wb = SO3<T>::fromRaw(parameters[0]);
    V3<T> t_wb{parameters[0][4], parameters[0][5], parameters[0][6]};
    T scale_corr = parameters[1][0];
    T s = scale + scale_corr;
    SO3<T> R_BC_corr = SO3<T>::fromRaw(parameters[2]);
    V3<T> p_BC_corr{parameters[3][0], parameters[3][1], parameters[3][2]};
    T t_offset_cam_imu_corr = parameters[4][0];
    SO3<double> q_bc0{camera->cam.q_bc[0], camera->cam.q_bc[1], camera->cam.q_bc[2], camera->cam.q_bc[3]};
    M3<double> Rbc0d = q_bc0.matrix();
    M3<T> Rbc0;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Rbc0.m[i][j] = T(Rbc0d.m[i][j]);
    V3<T> pbc0{T(camera->cam.p_bc[0]), T(camera->cam.p_bc[1]), T(camera->cam.p_bc[2])};
    M3<T> R_BC = mul(Rbc0, R_BC_corr.matrix());
    V3<T> p_BC = mul(Rbc0, p_BC_corr) + pbc0;
    M3<T> Rwb = q_wb.matrix();
    M3<T> R_WC = mul(Rwb, R_BC);
    M3<T> sR;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) sR.m[i][j] = s * Rwb.m[i][j];
    V3<T> p_WC = mul(sR, p_BC) + t_wb;
    M3<T> R_CW = transpose(R_WC);
    M3<T> mR;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) mR.m[i][j] = -1.0 * R_CW.m[i][j];
    V3<T> p_CW = mul(mR, p_WC);
    for (int n = 0; n < n_reproj; ++n) {
      V3<T> p_WLi{parameters[5 + n][0], parameters[5 + n][1], parameters[5 + n][2]};
      V3<T> p_CLi = mul(R_CW, p_WLi) + p_CW;
      T reproj_uv[2];
      project(camera->cam, p_CLi, reproj_uv);
      T u = uv[2 * n] + vel[2 * n] * (camera->t_offset_cam_imu + t_offset_cam_imu_corr);
      T v = uv[2 * n + 1] + vel[2 * n + 1] * (camera->t_offset_cam_imu + t_offset_cam_imu_corr);
      if ((u < T(0.0)) || (u > T(camera->w)) || (v < T(0.0)) || (v > T(camera->h))) {
        sResiduals[2 * n + 0] = T(0); sResiduals[2 * n + 1] = T(0);
        continue;
      }
      T thr = T(100.0);
      T sq = (reproj_uv[0] - u) * (reproj_uv[0] - u) + (reproj_uv[1] - v) * (reproj_uv[1] - v);
      if (sq > thr) {
        sResiduals[2 * n + 0] = T(0); sResiduals[2 * n + 1] = T(0);
        continue;
      }
      sResiduals[2 * n + 0] = inv_std * (reproj_uv[0] - u);
      sResiduals[2 * n + 1] = inv_std * (reproj_uv[1] - v);
    }
    return true;
  }
};

// global_position_cost_functor.h:24-51 (discrete time).  Parameters: T_WB0(7), T_WB1(7), scale_corr, t_offset_corr, p_BP(3)
struct DtGlobalPositionCostFunctor {
  double meas_t, meas_pos[3], scale, t_k0, t_k1, t_offset_gp_imu_init, inv_std;
  template <class T>
  bool operator()(T const* const* parameters, T* r) const {
    SO3<T> q0 = SO3<T>::fromRaw(parameters[0]), q1 = SO3<T>::fromRaw(parameters[1]);
    V3<T> t0{parameters[0][4], parameters[0][5], parameters[0][6]}, t1{parameters[1][4], parameters[1][5], parameters[1][6]};
    T scale_corr = parameters[2][0];
    T s = scale + scale_corr;
    T t_offset_corr = parameters[3][0];
    T t_offset = t_offset_gp_imu_init + t_offset_corr;
    V3<T> p_BP{parameters[4][0], parameters[4][1], parameters[4][2]};
    T t_interp = meas_t + t_offset;
    T lambda = (t_interp - t_k0) / (t_k1 - t_k0);
    M3<T> R0 = q0.matrix(), R1 = q1.matrix(), sR0, sR1;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { sR0.m[i][j] = s * R0.m[i][j]; sR1.m[i][j] = s * R1.m[i][j]; }
    V3<T> p0 = mul(sR0, p_BP) + t0, p1 = mul(sR1, p_BP) + t1;
    T one_m = 1.0 - lambda;
    V3<T> pi{one_m * p0.x + lambda * p1.x, one_m * p0.y + lambda * p1.y, one_m * p0.z + lambda * p1.z};
    r[0] = inv_std * (pi.x - meas_pos[0]); r[1] = inv_std * (pi.y - meas_pos[1]); r[2] = inv_std * (pi.z - meas_pos[2]);
    return true;
  }
};

// ---- small dense helpers for the IMU factor (Eigen fixed-size algebra restated) ----
namespace dt_detail {
struct Q4 { double x, y, z, w; };   // Eigen::Quaterniond, coeffs() = [x y z w]
inline Q4 qmul(const Q4& a, const Q4& b) {   // Eigen quaternion product (no normalisation)
  return {a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y, a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z, a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x,
          a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z};
}
inline Q4 qinv(const Q4& q) { const double n2 = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w; return {-q.x / n2, -q.y / n2, -q.z / n2, q.w / n2}; }   // Eigen inverse(): conjugate / squaredNorm
inline void qrot(const Q4& q, double R[3][3]) { SO3<double> s{q.x, q.y, q.z, q.w}; M3<double> m = s.matrix(); for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[i][j] = m.m[i][j]; }
inline double sinc(double x) {   // sensor/imu.h:24-40
  if (std::fabs(x) > 1e-6) return std::sin(x) / x;
  const double c_2 = 1.0 / 6.0, c_4 = 1.0 / 120.0, c_6 = 1.0 / 5040.0;
  const double x_2 = x * x, x_4 = x_2 * x_2, x_6 = x_2 * x_2 * x_2;
  return 1.0 - c_2 * x_2 + c_4 * x_4 - c_6 * x_6;
}
inline Q4 deltaQ(const double a[3]) {   // imu_cost_functor.h:25-33
  const double halfnorm = 0.5 * std::sqrt(a[0] * a[0] + a[1] * a[1] + a[2] * a[2]), sc = sinc(halfnorm) * 0.5;
  return {sc * a[0], sc * a[1], sc * a[2], std::cos(halfnorm)};
}
inline void skew(const double w[3], double S[3][3]) { S[0][0] = 0; S[0][1] = -w[2]; S[0][2] = w[1]; S[1][0] = w[2]; S[1][1] = 0; S[1][2] = -w[0]; S[2][0] = -w[1]; S[2][1] = w[0]; S[2][2] = 0; }
inline void expmapDerivativeSO3(const double om[3], double J[3][3]) {   // matrix_operations.h:36-51
  const double theta2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) J[i][j] = i == j ? 1.0 : 0.0;
  if (theta2 <= std::numeric_limits<double>::epsilon()) return;
  const double theta = std::sqrt(theta2);
  double Y[3][3], Y2[3][3];
  skew(om, Y);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Y[i][j] /= theta;
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double s = 0; for (int k = 0; k < 3; ++k) s += Y[i][k] * Y[k][j]; Y2[i][j] = s; }
  const double a = (1.0 - std::cos(theta)) / theta, b = 1.0 - std::sin(theta) / theta;
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) J[i][j] = J[i][j] - a * Y[i][j] + b * Y2[i][j];
}
inline void qplus(const Q4& q, double Q[4][4]) {   // matrix_operations.h:58-68
  const double c[4] = {q.x, q.y, q.z, q.w};
  Q[0][0] = c[3]; Q[0][1] = -c[2]; Q[0][2] = c[1]; Q[0][3] = c[0];
  Q[1][0] = c[2]; Q[1][1] = c[3]; Q[1][2] = -c[0]; Q[1][3] = c[1];
  Q[2][0] = -c[1]; Q[2][1] = c[0]; Q[2][2] = c[3]; Q[2][3] = c[2];
  Q[3][0] = -c[0]; Q[3][1] = -c[1]; Q[3][2] = -c[2]; Q[3][3] = c[3];
}
inline void qoplus(const Q4& q, double Q[4][4]) {   // matrix_operations.h:71-80
  const double c[4] = {q.x, q.y, q.z, q.w};
  Q[0][0] = c[3]; Q[0][1] = c[2]; Q[0][2] = -c[1]; Q[0][3] = c[0];
  Q[1][0] = -c[2]; Q[1][1] = c[3]; Q[1][2] = c[0]; Q[1][3] = c[1];
  Q[2][0] = c[1]; Q[2][1] = -c[0]; Q[2][2] = c[3]; Q[2][3] = c[2];
  Q[3][0] = -c[0]; Q[3][1] = -c[1]; Q[3][2] = -c[2]; Q[3][3] = c[3];
}
inline void m3mul(const double A[3][3], const double B[3][3], double C[3][3]) { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { double s = 0; for (int k = 0; k < 3; ++k) s += A[i][k] * B[k][j]; C[i][j] = s; } }
inline void m3vec(const double A[3][3], const double v[3], double o[3]) { for (int i = 0; i < 3; ++i) o[i] = A[i][0] * v[0] + A[i][1] * v[1] + A[i][2] * v[2]; }
inline void m4mul(const double A[4][4], const double B[4][4], double C[4][4]) { for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) { double s = 0; for (int k = 0; k < 4; ++k) s += A[i][k] * B[k][j]; C[i][j] = s; } }
// inverse of a 15 x 15 by LU with partial pivoting (Eigen's inverse() of a fixed 15 x 15 goes through PartialPivLU)
inline bool inv15(const double A[15][15], double X[15][15]) {
  double M[15][30];
  for (int i = 0; i < 15; ++i) for (int j = 0; j < 15; ++j) { M[i][j] = A[i][j]; M[i][15 + j] = i == j ? 1.0 : 0.0; }
  for (int c = 0; c < 15; ++c) {
    int piv = c; double best = std::fabs(M[c][c]);
    for (int r = c + 1; r < 15; ++r) if (std::fabs(M[r][c]) > best) { best = std::fabs(M[r][c]); piv = r; }
    if (best == 0.0) return false;
    if (piv != c) for (int j = 0; j < 30; ++j) std::swap(M[c][j], M[piv][j]);
    for (int r = c + 1; r < 15; ++r) { const double f = M[r][c] / M[c][c]; if (f != 0.0) for (int j = c; j < 30; ++j) M[r][j] -= f * M[c][j]; }
  }
  for (int c = 14; c >= 0; --c) {
    for (int j = 15; j < 30; ++j) { double s = M[c][j]; for (int k = c + 1; k < 15; ++k) s -= M[c][k] * M[k][j]; M[c][j] = s / M[c][c]; }
  }
  for (int i = 0; i < 15; ++i) for (int j = 0; j < 15; ++j) X[i][j] = M[i][15 + j];
  return true;
}
}  // namespace dt_detail

struct DtImuParams { double gyro_noise_density = 0, acc_noise_density = 0, gyro_random_walk = 0, acc_random_walk = 0, g_max = 7.8, a_max = 150.0; double g_W[3] = {0, 0, 9.80665}; };
struct DtImuSample { double t; double gyr[3], acc[3]; };

// imu_cost_functor.h:36-942.  Parameters: pose k (7), speed-and-biases k (9: v, b_g, b_a), pose k+1 (7), speed-and-biases k+1 (9).
// The pre-integration is cached in mutable members and redone when the gyro bias moved (:693-704): the residual depends on the order of the
// Evaluate calls, which is part of what the reference computes.
struct DtImuFactor {
  std::vector<DtImuSample> meas;
  DtImuParams par;
  double t0 = 0, t1 = 0;
  mutable dt_detail::Q4 Delta_q{0, 0, 0, 1};
  mutable double C_integral[3][3] = {}, C_doubleintegral[3][3] = {}, acc_integral[3] = {}, acc_doubleintegral[3] = {}, cross[3][3] = {};
  mutable double dalpha_db_g[3][3] = {}, dv_db_g[3][3] = {}, dp_db_g[3][3] = {};
  mutable double P_delta[15][15] = {}, sqrt_info[15][15] = {};
  mutable double sb_ref[9] = {};
  mutable bool redo = true;
  mutable int redo_counter = 0;

  // :377-594
  int redoPreintegration(const double* sb) const {
    using namespace dt_detail;
    double time = t0;
    if (!(meas.back().t >= t1)) return -1;
    Delta_q = Q4{0, 0, 0, 1};
    std::memset(C_integral, 0, sizeof(C_integral)); std::memset(C_doubleintegral, 0, sizeof(C_doubleintegral));
    std::memset(acc_integral, 0, sizeof(acc_integral)); std::memset(acc_doubleintegral, 0, sizeof(acc_doubleintegral));
    std::memset(cross, 0, sizeof(cross)); std::memset(dalpha_db_g, 0, sizeof(dalpha_db_g)); std::memset(dv_db_g, 0, sizeof(dv_db_g)); std::memset(dp_db_g, 0, sizeof(dp_db_g));
    std::memset(P_delta, 0, sizeof(P_delta));
    bool has_started = false, last_iteration = false;
    int n_integrated = 0;
    for (size_t i = 0; i + 1 < meas.size(); ++i) {
      double omega_S_0[3], acc_S_0[3], omega_S_1[3], acc_S_1[3];
      for (int c = 0; c < 3; ++c) { omega_S_0[c] = meas[i].gyr[c]; acc_S_0[c] = meas[i].acc[c]; omega_S_1[c] = meas[i + 1].gyr[c]; acc_S_1[c] = meas[i + 1].acc[c]; }
      double nexttime = meas[i + 1].t;
      double dt = nexttime - time;
      if (t1 < nexttime) {
        const double interval = nexttime - meas[i].t;
        nexttime = t1;
        last_iteration = true;
        dt = nexttime - time;
        const double r = dt / interval;
        for (int c = 0; c < 3; ++c) { omega_S_1[c] = (1.0 - r) * omega_S_0[c] + r * omega_S_1[c]; acc_S_1[c] = (1.0 - r) * acc_S_0[c] + r * acc_S_1[c]; }
      }
      if (dt <= 0.0) continue;
      if (!has_started) {
        has_started = true;
        const double r = dt / (nexttime - meas[i].t);
        for (int c = 0; c < 3; ++c) { omega_S_0[c] = r * omega_S_0[c] + (1.0 - r) * omega_S_1[c]; acc_S_0[c] = r * acc_S_0[c] + (1.0 - r) * acc_S_1[c]; }
      }
      double sigma_g_c = par.gyro_noise_density, sigma_a_c = par.acc_noise_density;
      bool gsat = false, asat = false;
      for (int c = 0; c < 3; ++c) { gsat |= std::fabs(omega_S_0[c]) > par.g_max || std::fabs(omega_S_1[c]) > par.g_max; asat |= std::fabs(acc_S_0[c]) > par.a_max || std::fabs(acc_S_1[c]) > par.a_max; }
      if (gsat) sigma_g_c *= 100;
      if (asat) sigma_a_c *= 100;
      // orientation
      double omega_S_true[3], acc_S_true[3];
      for (int c = 0; c < 3; ++c) { omega_S_true[c] = 0.5 * (omega_S_0[c] + omega_S_1[c]) - sb[3 + c]; acc_S_true[c] = 0.5 * (acc_S_0[c] + acc_S_1[c]) - sb[6 + c]; }
      const double theta_half = std::sqrt(omega_S_true[0] * omega_S_true[0] + omega_S_true[1] * omega_S_true[1] + omega_S_true[2] * omega_S_true[2]) * 0.5 * dt;
      const double sinc_theta_half = sinc(theta_half), cos_theta_half = std::cos(theta_half);
      Q4 dq{sinc_theta_half * omega_S_true[0] * 0.5 * dt, sinc_theta_half * omega_S_true[1] * 0.5 * dt, sinc_theta_half * omega_S_true[2] * 0.5 * dt, cos_theta_half};
      Q4 Delta_q_1 = qmul(Delta_q, dq);
      double C[3][3], C_1[3][3], CC[3][3];
      qrot(Delta_q, C); qrot(Delta_q_1, C_1);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) CC[a][b] = C[a][b] + C_1[a][b];
      double C_integral_1[3][3], acc_integral_1[3], CCa[3];
      m3vec(CC, acc_S_true, CCa);
      for (int a = 0; a < 3; ++a) { for (int b = 0; b < 3; ++b) C_integral_1[a][b] = C_integral[a][b] + 0.5 * CC[a][b] * dt; acc_integral_1[a] = acc_integral[a] + 0.5 * CCa[a] * dt; }
      for (int a = 0; a < 3; ++a) { for (int b = 0; b < 3; ++b) C_doubleintegral[a][b] += C_integral[a][b] * dt + 0.25 * CC[a][b] * dt * dt; }
      double adi_step[3];
      for (int a = 0; a < 3; ++a) { adi_step[a] = acc_integral[a] * dt + 0.25 * CCa[a] * dt * dt; acc_doubleintegral[a] += adi_step[a]; }
      // Jacobian parts
      double om_dt[3] = {omega_S_true[0] * dt, omega_S_true[1] * dt, omega_S_true[2] * dt}, Jr[3][3], T1[3][3];
      expmapDerivativeSO3(om_dt, Jr);
      m3mul(C_1, Jr, T1);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) dalpha_db_g[a][b] += T1[a][b] * dt;
      double Rdqi[3][3], cross_1[3][3];
      qrot(qinv(dq), Rdqi);
      m3mul(Rdqi, cross, T1);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) cross_1[a][b] = T1[a][b] + Jr[a][b] * dt;
      double acc_S_x[3][3], A0[3][3], A1[3][3], T2[3][3], W[3][3];
      skew(acc_S_true, acc_S_x);
      m3mul(C, acc_S_x, T1); m3mul(T1, cross, A0);
      m3mul(C_1, acc_S_x, T2); m3mul(T2, cross_1, A1);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) W[a][b] = A0[a][b] + A1[a][b];
      double dv_db_g_1[3][3];
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) { dv_db_g_1[a][b] = dv_db_g[a][b] + 0.5 * dt * W[a][b]; }
      // covariance propagation (uses dv_db_g, C_integral, acc_integral BEFORE their shift)
      double F[15][15];
      for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) F[a][b] = a == b ? 1.0 : 0.0;
      double S1[3][3], S2[3][3], h[3];
      skew(adi_step, S1);
      for (int a = 0; a < 3; ++a) h[a] = 0.5 * CCa[a] * dt;
      skew(h, S2);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) {
        F[a][3 + b] = -S1[a][b];
        F[a][6 + b] = (a == b ? 1.0 : 0.0) * dt;
        F[a][9 + b] = dt * dv_db_g[a][b] + 0.25 * dt * dt * W[a][b];
        F[a][12 + b] = -C_integral[a][b] * dt + 0.25 * CC[a][b] * dt * dt;
        F[3 + a][9 + b] = -dt * C_1[a][b];
        F[6 + a][3 + b] = -S2[a][b];
        F[6 + a][9 + b] = 0.5 * dt * W[a][b];
        F[6 + a][12 + b] = -0.5 * CC[a][b] * dt;
      }
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) dp_db_g[a][b] += dt * dv_db_g[a][b] + 0.25 * dt * dt * W[a][b];
      double FP[15][15];
      for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) { double s = 0; for (int k = 0; k < 15; ++k) s += F[a][k] * P_delta[k][b]; FP[a][b] = s; }
      for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) { double s = 0; for (int k = 0; k < 15; ++k) s += FP[a][k] * F[b][k]; P_delta[a][b] = s; }
      const double sigma2_dalpha = dt * sigma_g_c * sigma_g_c;
      const double sigma2_v = dt * sigma_a_c * sigma_a_c;
      const double sigma2_p = 0.5 * dt * dt * sigma2_v;
      const double sigma2_b_g = dt * par.gyro_random_walk * par.gyro_random_walk;
      const double sigma2_b_a = dt * par.acc_random_walk * par.acc_random_walk;
      for (int a = 0; a < 3; ++a) { P_delta[3 + a][3 + a] += sigma2_dalpha; P_delta[6 + a][6 + a] += sigma2_v; P_delta[a][a] += sigma2_p; P_delta[9 + a][9 + a] += sigma2_b_g; P_delta[12 + a][12 + a] += sigma2_b_a; }
      // memory shift
      Delta_q = Delta_q_1;
      std::memcpy(C_integral, C_integral_1, sizeof(C_integral)); std::memcpy(acc_integral, acc_integral_1, sizeof(acc_integral));
      std::memcpy(cross, cross_1, sizeof(cross)); std::memcpy(dv_db_g, dv_db_g_1, sizeof(dv_db_g));
      time = nexttime;
      ++n_integrated;
      if (last_iteration) break;
    }
    std::memcpy(sb_ref, sb, sizeof(sb_ref));
    double Ps[15][15], info[15][15];
    for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) Ps[a][b] = 0.5 * P_delta[a][b] + 0.5 * P_delta[b][a];
    std::memcpy(P_delta, Ps, sizeof(Ps));
    if (!inv15(P_delta, info)) return -2;
    double infs[15][15];
    for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) infs[a][b] = 0.5 * info[a][b] + 0.5 * info[b][a];
    // LLT: information = L L^T; square_root_information = L^T
    double L[15][15] = {};
    for (int j = 0; j < 15; ++j) {
      double d = infs[j][j];
      for (int k = 0; k < j; ++k) d -= L[j][k] * L[j][k];
      d = std::sqrt(d);
      L[j][j] = d;
      for (int i = j + 1; i < 15; ++i) { double s = infs[i][j]; for (int k = 0; k < j; ++k) s -= L[i][k] * L[j][k]; L[i][j] = s / d; }
    }
    for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) sqrt_info[a][b] = L[b][a];
    return n_integrated;
  }

  // :667-856.  jac (may be null): 15 x 32 row-major = [15x7 | 15x9 | 15x7 | 15x9], the blocks the reference hands to Ceres
  void evaluate(const std::vector<const double*>& p, double* residuals, double* jac, double* jac_minimal = nullptr) const {
    using namespace dt_detail;
    const double* T0 = p[0]; const double* sb0 = p[1]; const double* T1p = p[2]; const double* sb1 = p[3];
    const Q4 q0{T0[0], T0[1], T0[2], T0[3]}, q1{T1p[0], T1p[1], T1p[2], T1p[3]};
    double C_WB_0[3][3], C_B0_W[3][3];
    qrot(q0, C_WB_0);
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) C_B0_W[a][b] = C_WB_0[b][a];
    const double delta_t = t1 - t0;
    double Delta_b[6];
    for (int c = 0; c < 6; ++c) Delta_b[c] = sb0[3 + c] - sb_ref[3 + c];
    redo = redo || (std::sqrt(Delta_b[0] * Delta_b[0] + Delta_b[1] * Delta_b[1] + Delta_b[2] * Delta_b[2]) * delta_t > 0.0001);
    if (redo) {
      redoPreintegration(sb0);
      redo_counter++;
      for (int c = 0; c < 6; ++c) Delta_b[c] = 0.0;
      redo = false;
    }
    const double* g_W = par.g_W;
    double F0[15][15], F1[15][15];
    for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) { F0[a][b] = a == b ? 1.0 : 0.0; F1[a][b] = a == b ? -1.0 : 0.0; }
    double delta_p_est_W[3], delta_v_est_W[3];
    for (int c = 0; c < 3; ++c) {
      delta_p_est_W[c] = T0[4 + c] - T1p[4 + c] + sb0[c] * delta_t - 0.5 * g_W[c] * delta_t * delta_t;
      delta_v_est_W[c] = sb0[c] - sb1[c] - g_W[c] * delta_t;
    }
    double mda[3];
    for (int a = 0; a < 3; ++a) mda[a] = -(dalpha_db_g[a][0] * Delta_b[0] + dalpha_db_g[a][1] * Delta_b[1] + dalpha_db_g[a][2] * Delta_b[2]);
    const Q4 Dq = qmul(deltaQ(mda), Delta_q);
    double Sp[3][3], Sv[3][3], T[3][3];
    skew(delta_p_est_W, Sp); skew(delta_v_est_W, Sv);
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) { F0[a][b] = C_B0_W[a][b]; F0[a][6 + b] = C_B0_W[a][b] * delta_t; F0[a][9 + b] = dp_db_g[a][b]; F0[a][12 + b] = -C_doubleintegral[a][b];
      F0[6 + a][6 + b] = C_B0_W[a][b]; F0[6 + a][9 + b] = dv_db_g[a][b]; F0[6 + a][12 + b] = -C_integral[a][b]; F1[a][b] = -C_B0_W[a][b]; F1[6 + a][6 + b] = -C_B0_W[a][b]; }
    m3mul(C_B0_W, Sp, T); for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F0[a][3 + b] = T[a][b];
    m3mul(C_B0_W, Sv, T); for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F0[6 + a][3 + b] = T[a][b];
    double QA[4][4], QB[4][4], QC[4][4], QD[4][4];
    const Q4 q1i = qinv(q1);
    qplus(qmul(Dq, q1i), QA); qoplus(q0, QB); m4mul(QA, QB, QC);
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F0[3 + a][3 + b] = QC[a][b];
    qoplus(qmul(q1i, q0), QA); qoplus(Dq, QB); m4mul(QA, QB, QC);
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) { double s = 0; for (int k = 0; k < 3; ++k) s += QC[a][k] * (-dalpha_db_g[k][b]); F0[3 + a][9 + b] = s; }
    qplus(Dq, QA); qoplus(q0, QB); m4mul(QA, QB, QC); qplus(q1i, QD); m4mul(QC, QD, QA);
    for (int a = 0; a < 3; ++a) for (int b = 0; b < 3; ++b) F1[3 + a][3 + b] = -QA[a][b];
    double error[15];
    double t3[3];
    m3vec(C_B0_W, delta_p_est_W, t3);
    for (int a = 0; a < 3; ++a) { double s = 0; for (int k = 0; k < 6; ++k) s += F0[a][9 + k] * Delta_b[k]; error[a] = t3[a] + acc_doubleintegral[a] + s; }
    const Q4 qe = qmul(Dq, qmul(q1i, q0));
    error[3] = 2.0 * qe.x; error[4] = 2.0 * qe.y; error[5] = 2.0 * qe.z;
    m3vec(C_B0_W, delta_v_est_W, t3);
    for (int a = 0; a < 3; ++a) { double s = 0; for (int k = 0; k < 6; ++k) s += F0[6 + a][9 + k] * Delta_b[k]; error[6 + a] = t3[a] + acc_integral[a] + s; }
    for (int c = 0; c < 6; ++c) error[9 + c] = sb0[3 + c] - sb1[3 + c];
    for (int a = 0; a < 15; ++a) { double s = 0; for (int k = 0; k < 15; ++k) s += sqrt_info[a][k] * error[k]; residuals[a] = s; }
    if (!jac && !jac_minimal) return;
    double SF0[15][15], SF1[15][15];
    for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) { double s0 = 0, s1 = 0; for (int k = 0; k < 15; ++k) { s0 += sqrt_info[a][k] * F0[k][b]; s1 += sqrt_info[a][k] * F1[k][b]; } SF0[a][b] = s0; SF1[a][b] = s1; }
    auto lift = [&](const double* x, double Jl[6][7]) {   // pose_local_parameterization.h:20-31 (reads x[3..6] as a quaternion: the OKVIS layout)
      for (int a = 0; a < 6; ++a) for (int b = 0; b < 7; ++b) Jl[a][b] = 0.0;
      for (int a = 0; a < 3; ++a) Jl[a][a] = 1.0;
      const Q4 q_inv{-x[3], -x[4], -x[5], x[6]};     // Eigen::Quaterniond(w = x[6], -x[3], -x[4], -x[5])
      double Qp[4][4];
      qoplus(q_inv, Qp);
      for (int a = 0; a < 3; ++a) for (int b = 0; b < 4; ++b) Jl[3 + a][3 + b] = 2.0 * Qp[a][b];
    };
    if (jac_minimal)   // jacobians_minimal of the reference (:799-806, 829-836): 15 x 30 = [15x6 | 15x9 | 15x6 | 15x9]
      for (int a = 0; a < 15; ++a) for (int b = 0; b < 15; ++b) { jac_minimal[(size_t)a * 30 + b] = SF0[a][b]; jac_minimal[(size_t)a * 30 + 15 + b] = SF1[a][b]; }
    if (!jac) return;
    double L0[6][7], L1[6][7];
    lift(T0, L0); lift(T1p, L1);
    for (int a = 0; a < 15; ++a) {
      double* row = jac + (size_t)a * 32;
      for (int b = 0; b < 7; ++b) { double s0 = 0, s1 = 0; for (int k = 0; k < 6; ++k) { s0 += SF0[a][k] * L0[k][b]; s1 += SF1[a][k] * L1[k][b]; } row[b] = s0; row[16 + b] = s1; }
      for (int b = 0; b < 9; ++b) { row[7 + b] = SF0[a][6 + b]; row[23 + b] = SF1[a][6 + b]; }
    }
  }
};

struct DtFrame { int state_id; double inv_std; std::vector<double> uv, vel; std::vector<int> lm; };
struct DtGp { double t, pos[3]; int frame0; double inv_std; };

struct DtProblem {
  // states (estimator.h:88-125, state.h)
  std::vector<double> pose;     // 7 F  [qx qy qz qw tx ty tz]
  std::vector<double> sb;       // 9 F  [v, b_g, b_a]
  std::vector<double> ts_cam, ts_imu;
  std::vector<double> lm;
  DtCamera camera;
  DtImuParams imu;
  double scale = 0, scale_corr = 0, R_bc_corr[4] = {0, 0, 0, 1}, p_bc_corr[3] = {0, 0, 0}, t_offset_cam_imu_corr = 0;
  double t_offset_gs_imu = 0, t_offset_gs_imu_corr = 0, p_bp[3] = {0, 0, 0};
  bool optimize_p_bp = true, optimize_scale = false, optimize_camimu_extrinsics = true, optimize_t_offset_gs_imu = true;   // ctor :59-66
  double cauchy_a = 3.0;
  std::vector<DtFrame> frames;
  std::vector<std::shared_ptr<DtImuFactor>> imu_factors;
  std::vector<int> imu_frame0;
  std::vector<DtGp> gps;
  MiniProblem mp;
  bool built = false;
  size_t numStates() const { return ts_cam.size(); }

  void initializeStates(int F, const double* poses, const double* cam_ts) {   // :88-103
    pose.assign(poses, poses + 7 * (size_t)F); ts_cam.assign(cam_ts, cam_ts + F); ts_imu.resize(F); sb.assign(9 * (size_t)F, 0.0);
    for (int i = 0; i < F; ++i) ts_imu[i] = cam_ts[i] + camera.t_offset_cam_imu;
    built = false;
  }
  void initializeLinearVelocity() {   // :105-125
    const size_t n = numStates();
    for (size_t i = 0; i + 1 < n; ++i) {
      const double dt = ts_imu[i + 1] - ts_imu[i];
      for (int c = 0; c < 3; ++c) sb[9 * i + c] = (pose[7 * (i + 1) + 4 + c] - pose[7 * i + 4 + c]) / dt;
    }
    for (int c = 0; c < 3; ++c) sb[9 * (n - 1) + c] = sb[9 * (n - 2) + c];
  }
  void build() {
    if (built) return;
    mp = MiniProblem();
    const size_t F = numStates();
    for (size_t i = 0; i < F; ++i) mp.AddParameterBlockSE3(&pose[7 * i]);    // addLocalParametrization :149-165
    mp.AddParameterBlock(R_bc_corr, 4, true);
    for (size_t k = 0; k < imu_factors.size(); ++k) {   // addImuMeasurements :443-459 (driver: before the frames)
      const int f0 = imu_frame0[k];
      ResidualBlock b; b.kind = 1; b.index = (int)k; b.nres = 15; b.cauchy_a = 0;
      b.pb = {mp.AddParameterBlockSE3(&pose[7 * f0]), mp.AddParameterBlock(&sb[9 * f0], 9), mp.AddParameterBlockSE3(&pose[7 * (f0 + 1)]), mp.AddParameterBlock(&sb[9 * (f0 + 1)], 9)};
      std::shared_ptr<DtImuFactor> fac = imu_factors[k];
      b.eval = [fac](const std::vector<const double*>& p, const std::vector<int>&, double* r, double* j) { fac->evaluate(p, r, j); };
      mp.blocks.push_back(std::move(b));
    }
    for (size_t i = 0; i < frames.size(); ++i) {        // addImagePointsMeasurement :373-441
      const DtFrame& m = frames[i];
      const int n = (int)m.lm.size();
      if (n == 0) continue;
      ResidualBlock b; b.kind = 0; b.index = (int)i; b.nres = 2 * n; b.cauchy_a = cauchy_a;
      b.pb = {mp.AddParameterBlockSE3(&pose[7 * m.state_id]), mp.AddParameterBlock(&scale_corr, 1), mp.AddParameterBlock(R_bc_corr, 4, true), mp.AddParameterBlock(p_bc_corr, 3),
              mp.AddParameterBlock(&t_offset_cam_imu_corr, 1)};
      for (int j = 0; j < n; ++j) b.pb.push_back(mp.AddParameterBlock(&lm[3 * (size_t)m.lm[j]], 3));
      DtReprojectionCostFunctor f{&camera, scale, m.inv_std, n, m.uv.data(), m.vel.data()};
      b.eval = [f, n](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 2 * n, r, j); };
      mp.blocks.push_back(std::move(b));
      if (!optimize_scale) mp.SetParameterBlockConstant(&scale_corr);
      if (!optimize_camimu_extrinsics) { mp.SetParameterBlockConstant(R_bc_corr); mp.SetParameterBlockConstant(p_bc_corr); mp.SetParameterBlockConstant(&t_offset_cam_imu_corr); }
    }
    for (size_t i = 0; i < gps.size(); ++i) {           // addGlobalSensMeasurement :461-519
      const DtGp& m = gps[i];
      ResidualBlock b; b.kind = 2; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
      b.pb = {mp.AddParameterBlockSE3(&pose[7 * m.frame0]), mp.AddParameterBlockSE3(&pose[7 * (m.frame0 + 1)]), mp.AddParameterBlock(&scale_corr, 1),
              mp.AddParameterBlock(&t_offset_gs_imu_corr, 1), mp.AddParameterBlock(p_bp, 3)};
      DtGlobalPositionCostFunctor f{m.t, {m.pos[0], m.pos[1], m.pos[2]}, scale, ts_imu[m.frame0], ts_imu[m.frame0 + 1], t_offset_gs_imu, m.inv_std};
      b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      mp.blocks.push_back(std::move(b));
      if (!optimize_scale) mp.SetParameterBlockConstant(&scale_corr);
      if (!optimize_p_bp) mp.SetParameterBlockConstant(p_bp);
      if (!optimize_t_offset_gs_imu) mp.SetParameterBlockConstant(&t_offset_gs_imu_corr);
    }
    mp.finalize();
    built = true;
  }
};

}  // namespace oracle
