// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// CPU restatement, templated on the scalar type (double or oracle::Jet<K>), of the
// reference arithmetic that the hot path executes inside ceres::Solve():
//   * Sophus SO3 subset        thirdparty/Sophus/sophus/so3.hpp
//   * uniform B-spline tables  include/spline/spline_common.h:37-120
//   * CeresSplineHelper<N>     include/optimization/continuous_time/ceres_spline_helper.h:47-330
//   * Camera::project + distortion models 4/5/7
//                              include/sensor/camera.h:79-90, include/sensor/camera_calib.h:52-246
// Eigen/Sophus are not available in this image so the same formulas are
// written on plain arrays.  Each function cites the lines it follows.
#pragma once
#include <cmath>
#include <cstdint>
#include "jet.h"

namespace oracle {

constexpr double kEps = 1e-10;  // Sophus Constants<double>::epsilon(), common.hpp:94
constexpr double kPi = 3.141592653589793238462643383279502884;

template <class T> struct V3 { T x, y, z; };
template <class T> struct M3 { T m[3][3]; };

template <class T> inline V3<T> operator+(const V3<T>& a, const V3<T>& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
template <class T> inline V3<T> operator-(const V3<T>& a, const V3<T>& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
template <class T, class S> inline V3<T> scale(const V3<T>& a, const S& s) { return {a.x * s, a.y * s, a.z * s}; }
template <class T> inline V3<T> cross(const V3<T>& a, const V3<T>& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <class T> inline V3<T> mul(const M3<T>& A, const V3<T>& v) {
  return {A.m[0][0] * v.x + A.m[0][1] * v.y + A.m[0][2] * v.z,
          A.m[1][0] * v.x + A.m[1][1] * v.y + A.m[1][2] * v.z,
          A.m[2][0] * v.x + A.m[2][1] * v.y + A.m[2][2] * v.z};
}
template <class T> inline M3<T> mul(const M3<T>& A, const M3<T>& B) {
  M3<T> C;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[i][0] * B.m[0][j] + A.m[i][1] * B.m[1][j] + A.m[i][2] * B.m[2][j];
  return C;
}
template <class T> inline M3<T> transpose(const M3<T>& A) {
  M3<T> C;
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) C.m[i][j] = A.m[j][i];
  return C;
}

// ---------------------------------------------------------------------------------
// SO3 as a unit quaternion stored [x y z w] (so3.hpp:185-187).
template <class T>
struct SO3 {
  T x, y, z, w;

  static SO3 fromRaw(const T* p) { return {p[0], p[1], p[2], p[3]}; }  // Eigen::Map, no normalisation

  // so3.hpp:297-303
  void normalize() {
    T len = sqrt(x * x + y * y + z * z + w * w);
    x = x / len; y = y / len; z = z / len; w = w / len;
  }
  // so3.hpp:229-231 : SO3(conjugate) -> the quaternion ctor normalises (:481-487)
  SO3 inverse() const { SO3 r{-x, -y, -z, w}; r.normalize(); return r; }

  // so3.hpp:325-339 : Hamilton product, then normalising ctor
  SO3 operator*(const SO3& b) const {
    const SO3& a = *this;
    SO3 r;
    r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
    r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
    r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
    r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
    r.normalize();
    return r;
  }

  // so3.hpp:310-312 -> Eigen::QuaternionBase::toRotationMatrix
  M3<T> matrix() const {
    M3<T> R;
    T tx = 2.0 * x, ty = 2.0 * y, tz = 2.0 * z;
    T twx = tx * w, twy = ty * w, twz = tz * w;
    T txx = tx * x, txy = ty * x, txz = tz * x;
    T tyy = ty * y, tyz = tz * y, tzz = tz * z;
    R.m[0][0] = 1.0 - (tyy + tzz); R.m[0][1] = txy - twz; R.m[0][2] = txz + twy;
    R.m[1][0] = txy + twz; R.m[1][1] = 1.0 - (txx + tzz); R.m[1][2] = tyz - twx;
    R.m[2][0] = txz - twy; R.m[2][1] = tyz + twx; R.m[2][2] = 1.0 - (txx + tyy);
    return R;
  }

  // so3.hpp:358-367 : p + w*(2 v x p) + v x (2 v x p)
  V3<T> act(const V3<T>& p) const {
    V3<T> v{x, y, z};
    V3<T> uv = cross(v, p);
    uv = uv + uv;
    return p + scale(uv, w) + cross(v, uv);
  }

  // so3.hpp:583-619 (expAndTheta), no renormalisation
  static SO3 exp(const V3<T>& om) {
    T theta_sq = om.x * om.x + om.y * om.y + om.z * om.z;
    T imag, real;
    if (theta_sq < kEps * kEps) {
      T theta_po4 = theta_sq * theta_sq;
      imag = T(0.5) - T(1.0 / 48.0) * theta_sq + T(1.0 / 3840.0) * theta_po4;
      real = T(1.0) - T(1.0 / 8.0) * theta_sq + T(1.0 / 384.0) * theta_po4;
    } else {
      T theta = sqrt(theta_sq);
      T half = T(0.5) * theta;
      imag = sin(half) / theta;
      real = cos(half);
    }
    return {imag * om.x, imag * om.y, imag * om.z, real};
  }

  // so3.hpp:247-290 (logAndTheta)
  V3<T> log() const {
    T squared_n = x * x + y * y + z * z;
    T f;
    if (squared_n < kEps * kEps) {
      T squared_w = w * w;
      f = T(2.0) / w - T(2.0 / 3.0) * (squared_n) / (w * squared_w);
    } else {
      T n = sqrt(squared_n);
      if (abs(w) < kEps) {
        if (w > 0.0) f = T(kPi) / n; else f = T(-kPi) / n;
      } else {
        f = T(2.0) * atan(n / w) / n;
      }
    }
    return {f * x, f * y, f * z};
  }

  // so3.hpp:191-216 : d(this*exp(x))/dx at 0, 4x3 (rows x,y,z,w), double only
  void Dx_this_mul_exp_x_at_0(double J[4][3]) const {
    double c0 = 0.5 * value_of(w), c1 = 0.5 * value_of(z), c2 = -c1, c3 = 0.5 * value_of(y),
           c4 = 0.5 * value_of(x), c5 = -c4, c6 = -c3;
    J[0][0] = c0; J[0][1] = c2; J[0][2] = c3;
    J[1][0] = c1; J[1][1] = c0; J[1][2] = c5;
    J[2][0] = c6; J[2][1] = c4; J[2][2] = c0;
    J[3][0] = c5; J[3][1] = c6; J[3][2] = c2;
  }
};

// ---------------------------------------------------------------------------------
// spline_common.h:37-47
constexpr inline uint64_t C_n_k(uint64_t n, uint64_t k) {
  if (k > n) return 0;
  uint64_t r = 1;
  for (uint64_t d = 1; d <= k; ++d) { r *= n--; r /= d; }
  return r;
}

template <int N>
struct SplineTables {
  double blend[N][N];       // spline_common.h:54-85, _Cumulative=false
  double cum_blend[N][N];   // same, _Cumulative=true
  double base[N][N];        // spline_common.h:104-120
  SplineTables() {
    for (int cum = 0; cum < 2; ++cum) {
      double m[N][N];
      for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) m[i][j] = 0.0;
      for (int i = 0; i < N; ++i) {
        for (int j = 0; j < N; ++j) {
          double sum = 0;
          for (int s = j; s < N; ++s)
            sum += std::pow(-1.0, s - j) * double(C_n_k(N, s - j)) * std::pow(N - s - 1.0, N - 1.0 - i);
          m[j][i] = double(C_n_k(N - 1, N - 1 - i)) * sum;
        }
      }
      if (cum) {
        for (int i = 0; i < N; i++) for (int j = i + 1; j < N; j++) for (int c = 0; c < N; ++c) m[i][c] += m[j][c];
      }
      uint64_t factorial = 1;
      for (int i = 2; i < N; ++i) factorial *= i;
      for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) (cum ? cum_blend : blend)[i][j] = m[i][j] / double(factorial);
    }
    for (int i = 0; i < N; ++i) for (int j = 0; j < N; ++j) base[i][j] = 0.0;
    for (int j = 0; j < N; ++j) base[0][j] = 1.0;
    const int DEG = N - 1;
    int order = DEG;
    for (int n = 1; n < N; n++) {
      for (int i = DEG - order; i < N; i++) base[n][i] = (order - DEG + i) * base[n - 1][i];
      order--;
    }
  }
  static const SplineTables& get() { static const SplineTables t; return t; }
};

template <int N>
struct SplineHelper {
  // ceres_spline_helper.h:47-83 (baseCoeffsWithTime / baseCoeffsWithTimeVar): U is double or T
  template <int D, class U>
  static void baseCoeffs(U res[N], const U& t) {
    const auto& tb = SplineTables<N>::get();
    for (int j = 0; j < N; ++j) res[j] = U(0.0);
    if (D < N) {
      res[D] = U(tb.base[D][D]);
      U _t = t;
      for (int j = D + 1; j < N; j++) { res[j] = tb.base[D][j] * _t; _t = _t * t; }
    }
  }
  // out = (sc * M) * p : Eigen evaluates `inv_dt * M * p` left to right, i.e. the matrix is scaled first.
  template <class U>
  static void matvec(double sc, const double M[N][N], const U p[N], U out[N]) {
    for (int i = 0; i < N; ++i) { U s(0.0); for (int j = 0; j < N; ++j) s = s + (sc * M[i][j]) * p[j]; out[i] = s; }
  }

  // ceres_spline_helper.h:97-180 (evaluate_lie, U=double) and :186-271 (evaluate_lie_uvar, U=T)
  template <class T, class U>
  static void evaluate_lie(T const* const* sKnots, const U& u, double inv_dt, SO3<T>* transform_out,
                           V3<T>* vel_out = nullptr, V3<T>* accel_out = nullptr, V3<T>* jerk_out = nullptr) {
    const auto& tb = SplineTables<N>::get();
    U p[N], coeff[N], dcoeff[N], ddcoeff[N], dddcoeff[N];
    baseCoeffs<0, U>(p, u);
    matvec<U>(1.0, tb.cum_blend, p, coeff);
    if (vel_out || accel_out || jerk_out) {
      baseCoeffs<1, U>(p, u);
      matvec<U>(inv_dt, tb.cum_blend, p, dcoeff);
      if (accel_out || jerk_out) {
        baseCoeffs<2, U>(p, u);
        matvec<U>(inv_dt * inv_dt, tb.cum_blend, p, ddcoeff);
        if (jerk_out) {
          baseCoeffs<3, U>(p, u);
          matvec<U>(inv_dt * inv_dt * inv_dt, tb.cum_blend, p, dddcoeff);
        }
      }
    }
    if (transform_out) *transform_out = SO3<T>::fromRaw(sKnots[0]);
    V3<T> rot_vel{T(0.0), T(0.0), T(0.0)}, rot_accel{T(0.0), T(0.0), T(0.0)}, rot_jerk{T(0.0), T(0.0), T(0.0)};
    for (int i = 0; i < N - 1; i++) {
      SO3<T> p0 = SO3<T>::fromRaw(sKnots[i]);
      SO3<T> p1 = SO3<T>::fromRaw(sKnots[i + 1]);
      SO3<T> r01 = p0.inverse() * p1;
      V3<T> delta = r01.log();
      SO3<T> exp_kdelta = SO3<T>::exp(scale(delta, T(coeff[i + 1])));
      if (transform_out) *transform_out = (*transform_out) * exp_kdelta;
      if (vel_out || accel_out || jerk_out) {
        M3<T> A = exp_kdelta.inverse().matrix();
        rot_vel = mul(A, rot_vel);
        V3<T> rot_vel_current = scale(delta, T(dcoeff[i + 1]));
        rot_vel = rot_vel + rot_vel_current;
        if (accel_out || jerk_out) {
          rot_accel = mul(A, rot_accel);
          V3<T> accel_lie_bracket = cross(rot_vel, rot_vel_current);
          rot_accel = rot_accel + scale(delta, T(ddcoeff[i + 1])) + accel_lie_bracket;
          if (jerk_out) {
            rot_jerk = mul(A, rot_jerk);
            V3<T> inner = scale(rot_vel, T(ddcoeff[i + 1])) + scale(rot_accel, T(2.0) * T(dcoeff[i + 1])) -
                          scale(accel_lie_bracket, T(dcoeff[i + 1]));
            rot_jerk = rot_jerk + scale(delta, T(dddcoeff[i + 1])) + cross(inner, delta);
          }
        }
      }
    }
    if (vel_out) *vel_out = rot_vel;
    if (accel_out) *accel_out = rot_accel;
    if (jerk_out) *jerk_out = rot_jerk;
  }

  // ceres_spline_helper.h:282-303 (evaluate, U=double) and :308-330 (evaluate_uvar, U=T), DIM=3
  template <class T, int DERIV, class U>
  static void evaluate(T const* const* sKnots, const U& u, double inv_dt, V3<T>* out) {
    const auto& tb = SplineTables<N>::get();
    U p[N], coeff[N];
    baseCoeffs<DERIV, U>(p, u);
    matvec<U>(std::pow(inv_dt, DERIV), tb.blend, p, coeff);
    V3<T> acc{T(0.0), T(0.0), T(0.0)};
    for (int i = 0; i < N; ++i) {
      T c = T(coeff[i]);
      acc.x = acc.x + c * sKnots[i][0];
      acc.y = acc.y + c * sKnots[i][1];
      acc.z = acc.z + c * sKnots[i][2];
    }
    *out = acc;
  }
};

// ---------------------------------------------------------------------------------
// Camera: include/sensor/camera.h:79-90 + camera_calib.h
struct CameraModel {
  int model = 4;  // COLMAP id: 4 rad-tan, 5 equidistant, 7 FOV (camera_calib.h:216-246)
  double fx = 1, fy = 1, cx = 0, cy = 0;
  double d[4] = {0, 0, 0, 0};
  double q_bc[4] = {0, 0, 0, 1};  // T_b_c rotation (x y z w)
  double p_bc[3] = {0, 0, 0};     // T_b_c translation
};

template <class T>
inline void project(const CameraModel& cam, const V3<T>& pc, T uv[2]) {
  T px = pc.x / pc.z, py = pc.y / pc.z;
  if (cam.model == 4) {  // camera_calib.h:52-65
    T mx2 = px * px, my2 = py * py, mxy = px * py, rho2 = mx2 + my2;
    T rad = cam.d[0] * rho2 + cam.d[1] * rho2 * rho2;
    T nx = px + (px * rad + 2.0 * cam.d[2] * mxy + cam.d[3] * (rho2 + 2.0 * mx2));
    T ny = py + (py * rad + 2.0 * cam.d[3] * mxy + cam.d[2] * (rho2 + 2.0 * my2));
    px = nx; py = ny;
  } else if (cam.model == 5) {  // camera_calib.h:131-149 (no r->0 guard)
    T r = sqrt(px * px + py * py);
    T theta = atan(r);
    T t2 = theta * theta, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
    T thetad = theta * (1.0 + cam.d[0] * t2 + cam.d[1] * t4 + cam.d[2] * t6 + cam.d[3] * t8);
    T scaling = thetad / r;
    px = px * scaling; py = py * scaling;
  } else if (cam.model == 7) {  // camera_calib.h:177-186
    T s = T(cam.d[0]);
    T tan_s_half_x2 = 2.0 * tan(s / 2.0);
    T rad = sqrt(px * px + py * py);
    T factor = (rad < 0.001) ? T(1.0) : atan(rad * tan_s_half_x2) / (s * rad);
    px = px * factor; py = py * factor;
  }
  uv[0] = cam.fx * px + cam.cx;
  uv[1] = cam.fy * py + cam.cy;
}

}  // namespace oracle
