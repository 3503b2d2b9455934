// ctba_math.h — FP64 device functions for the continuous-time bundle adjuster (sm_100a).
//
// Everything here is a from-scratch analytic formulation (values + hand-derived Jacobians in the
// right-perturbation tangent space R <- R*Exp(delta)) of what the reference obtains through Ceres
// Jets.  Reference behaviour each function reproduces:
//   so3_exp / so3_log / quat_mul / quat_inv / quat_matrix
//        thirdparty/Sophus/sophus/so3.hpp:583-619, 247-290, 325-339 (+normalising ctor 481-487), 229-231, 310-312
//   spline_coeffs (lambda, lambda', lambda'', blending tables)
//        include/spline/spline_common.h:54-120, include/optimization/continuous_time/ceres_spline_helper.h:47-83
//   so3_spline_eval   ceres_spline_helper.h:97-271 (evaluate_lie / evaluate_lie_uvar)
//   r3 coefficients   ceres_spline_helper.h:282-330 (evaluate / evaluate_uvar)
//   camera_project    include/sensor/camera.h:79-90, include/sensor/camera_calib.h:52-65, 131-149, 177-186
// The functions are CTBA_HD so that tests/ can also compile them for the host to debug formulas on a
// GPU-less machine; the product library only ever calls them from __global__ kernels.
#pragma once
#include <math.h>
#include <stdint.h>

#if defined(__CUDACC__)
#define CTBA_HD __host__ __device__ __forceinline__
#else
#define CTBA_HD inline
#endif

namespace ctba {

constexpr double kSophusEps = 1e-10;  // Sophus Constants<double>::epsilon()
constexpr double kPi = 3.141592653589793238462643383279502884;
constexpr int kMaxOrder = 6;

struct v3 { double x, y, z; };
struct m3 { double a[9]; };  // row-major
struct quat { double x, y, z, w; };

CTBA_HD v3 operator+(v3 a, v3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
CTBA_HD v3 operator-(v3 a, v3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
CTBA_HD v3 operator*(double s, v3 a) { return {s * a.x, s * a.y, s * a.z}; }
CTBA_HD v3 cross(v3 a, v3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
CTBA_HD double dot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
CTBA_HD v3 mul(const m3& A, v3 v) {
  return {A.a[0] * v.x + A.a[1] * v.y + A.a[2] * v.z, A.a[3] * v.x + A.a[4] * v.y + A.a[5] * v.z,
          A.a[6] * v.x + A.a[7] * v.y + A.a[8] * v.z};
}
CTBA_HD v3 mulT(const m3& A, v3 v) {  // A^T v
  return {A.a[0] * v.x + A.a[3] * v.y + A.a[6] * v.z, A.a[1] * v.x + A.a[4] * v.y + A.a[7] * v.z,
          A.a[2] * v.x + A.a[5] * v.y + A.a[8] * v.z};
}
CTBA_HD m3 mul(const m3& A, const m3& B) {
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[3 * i + j] = A.a[3 * i] * B.a[j] + A.a[3 * i + 1] * B.a[3 + j] + A.a[3 * i + 2] * B.a[6 + j];
  return C;
}
CTBA_HD m3 mulABt(const m3& A, const m3& B) {  // A * B^T
  m3 C;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) C.a[3 * i + j] = A.a[3 * i] * B.a[3 * j] + A.a[3 * i + 1] * B.a[3 * j + 1] + A.a[3 * i + 2] * B.a[3 * j + 2];
  return C;
}
CTBA_HD m3 transpose(const m3& A) { return {{A.a[0], A.a[3], A.a[6], A.a[1], A.a[4], A.a[7], A.a[2], A.a[5], A.a[8]}}; }
CTBA_HD m3 hat(v3 v) { return {{0.0, -v.z, v.y, v.z, 0.0, -v.x, -v.y, v.x, 0.0}}; }
CTBA_HD m3 identity3() { return {{1, 0, 0, 0, 1, 0, 0, 0, 1}}; }
CTBA_HD m3 zero3() { return {{0, 0, 0, 0, 0, 0, 0, 0, 0}}; }
CTBA_HD void axpy(m3& Y, double s, const m3& X) {
#pragma unroll
  for (int i = 0; i < 9; ++i) Y.a[i] += s * X.a[i];
}

// ---- quaternion group ops with the reference's normalisation behaviour -------------------------
CTBA_HD quat quat_normalized(quat q) {
  double len = sqrt(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
  return {q.x / len, q.y / len, q.z / len, q.w / len};
}
CTBA_HD quat quat_inv(quat q) { return quat_normalized({-q.x, -q.y, -q.z, q.w}); }
CTBA_HD quat quat_mul(quat a, quat b) {
  quat r;
  r.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
  r.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
  r.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
  r.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
  return quat_normalized(r);
}
CTBA_HD m3 quat_matrix(quat q) {
  double tx = 2.0 * q.x, ty = 2.0 * q.y, tz = 2.0 * q.z;
  double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  return {{1.0 - (tyy + tzz), txy - twz, txz + twy, txy + twz, 1.0 - (txx + tzz), tyz - twx, txz - twy, tyz + twx, 1.0 - (txx + tyy)}};
}
CTBA_HD v3 quat_act(quat q, v3 p) {
  v3 v{q.x, q.y, q.z};
  v3 uv = cross(v, p);
  uv = uv + uv;
  return p + q.w * uv + cross(v, uv);
}
CTBA_HD quat so3_exp(v3 om) {
  double theta_sq = dot(om, om);
  double imag, real;
  if (theta_sq < kSophusEps * kSophusEps) {
    double t4 = theta_sq * theta_sq;
    imag = 0.5 - (1.0 / 48.0) * theta_sq + (1.0 / 3840.0) * t4;
    real = 1.0 - (1.0 / 8.0) * theta_sq + (1.0 / 384.0) * t4;
  } else {
    double theta = sqrt(theta_sq), half = 0.5 * theta;
    imag = sin(half) / theta;
    real = cos(half);
  }
  return {imag * om.x, imag * om.y, imag * om.z, real};
}
CTBA_HD v3 so3_log(quat q) {
  double n2 = q.x * q.x + q.y * q.y + q.z * q.z;
  double f;
  if (n2 < kSophusEps * kSophusEps) {
    f = 2.0 / q.w - (2.0 / 3.0) * n2 / (q.w * (q.w * q.w));
  } else {
    double n = sqrt(n2);
    if (fabs(q.w) < kSophusEps) f = (q.w > 0.0 ? kPi : -kPi) / n;
    else f = 2.0 * atan(n / q.w) / n;
  }
  return {f * q.x, f * q.y, f * q.z};
}

// ---- SO(3) Jacobians (closed forms; Taylor below theta^2 < 1e-2 to keep 1e-10 relative accuracy) ----
// Jr(phi)      = I - a [phi]x + b [phi]x^2,  a = (1-cos t)/t^2, b = (t - sin t)/t^3
// Jr^{-1}(phi) = I + 1/2 [phi]x + c [phi]x^2, c = 1/t^2 - (1+cos t)/(2 t sin t);  Jl^{-1}(phi) = Jr^{-1}(-phi)
CTBA_HD void so3_coeffs(double t2, double* a, double* b, double* c) {
  if (t2 < 1e-2) {
    *a = 0.5 + t2 * (-1.0 / 24.0 + t2 * (1.0 / 720.0 + t2 * (-1.0 / 40320.0 + t2 * (1.0 / 3628800.0 - t2 * (1.0 / 479001600.0)))));
    *b = 1.0 / 6.0 + t2 * (-1.0 / 120.0 + t2 * (1.0 / 5040.0 + t2 * (-1.0 / 362880.0 + t2 * (1.0 / 39916800.0 - t2 * (1.0 / 6227020800.0)))));
    *c = 1.0 / 12.0 + t2 * (1.0 / 720.0 + t2 * (1.0 / 30240.0 + t2 * (1.0 / 1209600.0 + t2 * (1.0 / 47900160.0 + t2 * (691.0 / 1307674368000.0)))));
  } else {
    double t = sqrt(t2), s = sin(t), co = cos(t), sh = sin(0.5 * t);
    *a = 2.0 * sh * sh / t2;
    *b = (t - s) / (t2 * t);
    *c = 1.0 / t2 - (1.0 + co) / (2.0 * t * s);
  }
}
CTBA_HD m3 hat_sq(v3 p) {  // [p]x^2 = p p^T - |p|^2 I
  double n2 = dot(p, p);
  return {{p.x * p.x - n2, p.x * p.y, p.x * p.z, p.x * p.y, p.y * p.y - n2, p.y * p.z, p.x * p.z, p.y * p.z, p.z * p.z - n2}};
}
CTBA_HD m3 so3_Jr(v3 phi) {
  double a, b, c; so3_coeffs(dot(phi, phi), &a, &b, &c);
  m3 J = identity3(); axpy(J, -a, hat(phi)); axpy(J, b, hat_sq(phi)); return J;
}
CTBA_HD void so3_Jr_inv_Jl_inv(v3 phi, m3* Jri, m3* Jli) {
  double a, b, c; so3_coeffs(dot(phi, phi), &a, &b, &c);
  m3 H = hat(phi), H2 = hat_sq(phi);
  m3 J = identity3(); axpy(J, c, H2);
  *Jri = J; *Jli = J;
  axpy(*Jri, 0.5, H); axpy(*Jli, -0.5, H);
}

// ---- uniform B-spline tables (spline_common.h:54-120), filled on the host once per order ----------
struct SplineTables {
  int N;
  double blend[kMaxOrder][kMaxOrder];      // plain blending matrix M   (row = knot, col = power of u)
  double cum[kMaxOrder][kMaxOrder];        // cumulative blending matrix
  double base[kMaxOrder][kMaxOrder];       // derivative coefficients
};

inline uint64_t binom(uint64_t n, uint64_t k) {
  if (k > n) return 0;
  uint64_t r = 1;
  for (uint64_t d = 1; d <= k; ++d) { r *= n--; r /= d; }
  return r;
}
inline void fill_tables(int N, SplineTables* t) {
  t->N = N;
  for (int cum = 0; cum < 2; ++cum) {
    double m[kMaxOrder][kMaxOrder] = {};
    for (int i = 0; i < N; ++i)
      for (int j = 0; j < N; ++j) {
        double sum = 0;
        for (int s = j; s < N; ++s) sum += pow(-1.0, s - j) * double(binom(N, s - j)) * pow(N - s - 1.0, N - 1.0 - i);
        m[j][i] = double(binom(N - 1, N - 1 - i)) * sum;
      }
    if (cum)
      for (int i = 0; i < N; i++) for (int j = i + 1; j < N; j++) for (int c = 0; c < N; ++c) m[i][c] += m[j][c];
    uint64_t fact = 1;
    for (int i = 2; i < N; ++i) fact *= i;
    for (int i = 0; i < kMaxOrder; ++i) for (int j = 0; j < kMaxOrder; ++j) (cum ? t->cum : t->blend)[i][j] = (i < N && j < N) ? m[i][j] / double(fact) : 0.0;
  }
  for (int i = 0; i < kMaxOrder; ++i) for (int j = 0; j < kMaxOrder; ++j) t->base[i][j] = 0.0;
  for (int j = 0; j < N; ++j) t->base[0][j] = 1.0;
  int order = N - 1;
  for (int n = 1; n < N; n++) {
    for (int i = (N - 1) - order; i < N; i++) t->base[n][i] = (order - (N - 1) + i) * t->base[n - 1][i];
    order--;
  }
}

// coeff[i] = sum_j (sc*M[i][j]) * base[D][j] * u^(j-D)     (ceres_spline_helper.h:47-65 + matrix product)
template <int N>
CTBA_HD void spline_coeffs(const double (*M)[kMaxOrder], const double (*base)[kMaxOrder], int D, double sc, double u, double* coeff) {
  double p[N];
#pragma unroll
  for (int j = 0; j < N; ++j) p[j] = 0.0;
  if (D < N) {
    p[D] = base[D][D];
    double t = u;
#pragma unroll
    for (int j = 0; j < N; ++j) {
      if (j > D) { p[j] = base[D][j] * t; t = t * u; }
    }
  }
#pragma unroll
  for (int i = 0; i < N; ++i) {
    double s = 0.0;
#pragma unroll
    for (int j = 0; j < N; ++j) s = s + (sc * M[i][j]) * p[j];
    coeff[i] = s;
  }
}

// ---- cumulative SO(3) B-spline: value, body rates, and right-tangent Jacobians w.r.t. the N knots ----
template <int N>
struct So3SplineResult {
  quat q;          // R_wb(u)
  v3 w;            // body angular velocity (if VEL)
  v3 wd;           // body angular acceleration (if ACC)
  m3 dR[N];        // d(right tangent of R)/d(delta_i)       (if JAC)
  m3 dW[N];        // d(w)/d(delta_i)                        (if JAC && VEL)
};

// q4: N knots, [x y z w] each, contiguous.
template <int N, bool VEL, bool ACC, bool JAC>
CTBA_HD void so3_spline_eval(const SplineTables& tb, const double* q4, double u, double inv_dt, So3SplineResult<N>* out) {
  double lam[N], dlam[N], ddlam[N];
  spline_coeffs<N>(tb.cum, tb.base, 0, 1.0, u, lam);
  if (VEL || ACC) spline_coeffs<N>(tb.cum, tb.base, 1, inv_dt, u, dlam);
  if (ACC) spline_coeffs<N>(tb.cum, tb.base, 2, inv_dt * inv_dt, u, ddlam);

  v3 d[N];       // d[j] = Log(R_{j-1}^T R_j), j = 1..N-1
  m3 At[N];      // At[j] = matrix(inverse(Exp(lam_j d_j)))  = A_j^T
  v3 wprev[N];   // A_j^T w_{j-1}
  quat acc{q4[0], q4[1], q4[2], q4[3]};
  v3 w{0, 0, 0}, wd{0, 0, 0};
#pragma unroll
  for (int j = 1; j < N; ++j) {
    quat p0{q4[4 * (j - 1)], q4[4 * (j - 1) + 1], q4[4 * (j - 1) + 2], q4[4 * (j - 1) + 3]};
    quat p1{q4[4 * j], q4[4 * j + 1], q4[4 * j + 2], q4[4 * j + 3]};
    quat r01 = quat_mul(quat_inv(p0), p1);
    d[j] = so3_log(r01);
    quat ek = so3_exp(lam[j] * d[j]);
    acc = quat_mul(acc, ek);
    if (VEL || ACC || JAC) At[j] = quat_matrix(quat_inv(ek));
    if (VEL || ACC) {
      w = mul(At[j], w);
      wprev[j] = w;
      v3 wc = dlam[j] * d[j];
      w = w + wc;
      if (ACC) {
        wd = mul(At[j], wd);
        wd = wd + ddlam[j] * d[j] + cross(w, wc);
      }
    }
  }
  out->q = acc;
  if (VEL) out->w = w;
  if (ACC) out->wd = wd;
  if (JAC) {
#pragma unroll
    for (int i = 0; i < N; ++i) { out->dR[i] = zero3(); if (VEL) out->dW[i] = zero3(); }
    m3 P = identity3();  // P_j = A_{N-1}^T ... A_{j+1}^T
#pragma unroll
    for (int j = N - 1; j >= 1; --j) {
      m3 H = so3_Jr(lam[j] * d[j]);
#pragma unroll
      for (int e = 0; e < 9; ++e) H.a[e] *= lam[j];
      m3 Jri, Jli;
      so3_Jr_inv_Jl_inv(d[j], &Jri, &Jli);
      m3 T = mul(P, H);
      m3 TJr = mul(T, Jri), TJl = mul(T, Jli);
      axpy(out->dR[j], 1.0, TJr);
      axpy(out->dR[j - 1], -1.0, TJl);
      if (VEL) {
        m3 G = mul(hat(wprev[j]), H);
        G.a[0] += dlam[j]; G.a[4] += dlam[j]; G.a[8] += dlam[j];
        m3 U = mul(P, G);
        axpy(out->dW[j], 1.0, mul(U, Jri));
        axpy(out->dW[j - 1], -1.0, mul(U, Jli));
      }
      P = mul(P, At[j]);
    }
    axpy(out->dR[0], 1.0, P);  // P_0
  }
}

// ---- camera (COLMAP model ids 4 rad-tan, 5 equidistant, 7 FOV) --------------------------------------
struct Camera {
  int model;
  double fx, fy, cx, cy;
  double d[4];
};

// uv = project(pc);  if JAC: Jp (2x3 row-major) = d uv / d pc
template <bool JAC>
CTBA_HD void camera_project(const Camera& cam, v3 pc, double* uv, double* Jp) {
  const double iz = 1.0 / pc.z;
  const double x = pc.x / pc.z, y = pc.y / pc.z;
  double xd, yd;
  double D00 = 1, D01 = 0, D10 = 0, D11 = 1;  // d(xd,yd)/d(x,y)
  if (cam.model == 4) {
    double mx2 = x * x, my2 = y * y, mxy = x * y, rho2 = mx2 + my2;
    double rad = cam.d[0] * rho2 + cam.d[1] * rho2 * rho2;
    xd = x + (x * rad + 2.0 * cam.d[2] * mxy + cam.d[3] * (rho2 + 2.0 * mx2));
    yd = y + (y * rad + 2.0 * cam.d[3] * mxy + cam.d[2] * (rho2 + 2.0 * my2));
    if (JAC) {
      double drad = cam.d[0] + 2.0 * cam.d[1] * rho2;  // d rad / d rho2
      D00 = 1.0 + rad + 2.0 * mx2 * drad + 2.0 * cam.d[2] * y + 6.0 * cam.d[3] * x;
      D01 = 2.0 * mxy * drad + 2.0 * cam.d[2] * x + 2.0 * cam.d[3] * y;
      D10 = D01;
      D11 = 1.0 + rad + 2.0 * my2 * drad + 2.0 * cam.d[3] * x + 6.0 * cam.d[2] * y;
    }
  } else if (cam.model == 5) {
    double r2 = x * x + y * y, r = sqrt(r2);
    double th = atan(r), t2 = th * th, t4 = t2 * t2, t6 = t4 * t2, t8 = t4 * t4;
    double poly = 1.0 + cam.d[0] * t2 + cam.d[1] * t4 + cam.d[2] * t6 + cam.d[3] * t8;
    double thd = th * poly;
    double sc = thd / r;
    xd = x * sc; yd = y * sc;
    if (JAC) {
      double dthd = poly + th * (2.0 * cam.d[0] * th + 4.0 * cam.d[1] * t2 * th + 6.0 * cam.d[2] * t4 * th + 8.0 * cam.d[3] * t6 * th);
      double dsc_dr = (dthd / (1.0 + r2) - sc) / r;  // d(thd/r)/dr
      double k = dsc_dr / r;                          // d sc / d x = k x
      D00 = sc + x * x * k; D01 = x * y * k; D10 = D01; D11 = sc + y * y * k;
    }
  } else if (cam.model == 7) {
    double s = cam.d[0], T = 2.0 * tan(s / 2.0);
    double r2 = x * x + y * y, r = sqrt(r2);
    if (r < 0.001) {
      xd = x; yd = y;
    } else {
      double f = atan(r * T) / (s * r);
      xd = x * f; yd = y * f;
      if (JAC) {
        double df_dr = (T / (1.0 + r2 * T * T) / (s * r)) - f / r;
        double k = df_dr / r;
        D00 = f + x * x * k; D01 = x * y * k; D10 = D01; D11 = f + y * y * k;
      }
    }
  } else {
    xd = x; yd = y;
  }
  uv[0] = cam.fx * xd + cam.cx;
  uv[1] = cam.fy * yd + cam.cy;
  if (JAC) {
    // d(x,y)/d pc = [iz 0 -x iz; 0 iz -y iz]
    double a00 = cam.fx * D00, a01 = cam.fx * D01, a10 = cam.fy * D10, a11 = cam.fy * D11;
    Jp[0] = a00 * iz; Jp[1] = a01 * iz; Jp[2] = -(a00 * x + a01 * y) * iz;
    Jp[3] = a10 * iz; Jp[4] = a11 * iz; Jp[5] = -(a10 * x + a11 * y) * iz;
  }
}

}  // namespace ctba
