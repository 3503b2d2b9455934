// ctba_terms.h — per-element residual + analytic-Jacobian evaluators (one call = one work item of a kernel).
//
// Replaces, with closed-form Jacobians, the reference's autodiff'ed functors:
//   frame_pose_eval   : spline part of ReprojectionCostFunctor (reprojection_cost_functor.h:53-77), once per FRAME
//   reproj_obs_eval   : per-point part of the same functor (:80-104) + Camera::project, once per OBSERVATION
//   imu_sample_eval   : AccelerometerCostFunctor + GyroCostFunctor (imu_cost_functor.h:48-77, 106-139)
//   gp_sample_eval    : GlobalPositionSensorCostFunctor (global_position_cost_functor.h:45-73)
//   bias_cost_eval    : BiasCostFunctor<4> (imu_cost_functor.h:351-369)
//   gravity_eval      : GravityNormCostFunctor (gravity_norm_cost_functor.h:26-39)
//   prior_eval        : PositionCostFunctor / OrientationCostFunctor (position_cost_functor.h, orientation_cost_functor.h)
//   align_eval        : GlobalPositionSensorWithAlignmentCorrectionCostFunctor (global_position_cost_functor.h:82-150)
// Tangent conventions: SO(3) blocks use the right perturbation q <- q*exp(delta)
// (lie_local_parametrization.h:34-41); everything else is Euclidean.
#pragma once
#include "ctba_math.h"

namespace ctba {

// Problem-wide constants handed to every kernel by value (lands in the constant bank).
struct Globals {
  SplineTables tb;    // order-N tables
  SplineTables tb4;   // cubic tables for the bias spline (Nb = 4)
  Camera cam;
  double q_bc0[4];    // calibration T_b_c rotation (x y z w)
  double p_bc0[3];    // calibration T_b_c translation
  double inv_dt, dt_s, inv_dt_bias;
  double scale;       // scale_ (the correction lives in the state vector)
  double gate_sq;     // 100.0 : squared pixel-error gate (reprojection_cost_functor.h:89)
};

// Small optimisable globals, device-resident: see GLB_* offsets.
enum {
  GLB_SCALE_CORR = 0, GLB_RBC_CORR = 1 /*4: xyzw*/, GLB_PBC_CORR = 5, GLB_PBP = 8, GLB_G = 11, GLB_T_CAM_IMU_CORR = 14,
  GLB_T_GP_CORR = 15, GLB_STATE_SIZE = 16
};
// tangent (border) layout of the same globals
enum { GT_SCALE = 0, GT_RBC = 1, GT_PBC = 4, GT_PBP = 7, GT_G = 10, GT_T_CAM_IMU = 13, GT_T_GP = 14, GT_SIZE = 15 };

// Per-frame record produced by frame_pose_eval and consumed per observation.
template <int N>
struct FramePose {
  double R_wb[9], p_wb[3];
  double R_cw[9], t_cw[3];      // p_c = R_cw p_w + t_cw
  double Crot[N][9];            // world-frame rotation perturbation of R_wb per unit right-tangent of knot i: R_wb * dR[i]
  double cpos[N];               // d p_wb / d pos-knot i  (times I3)
};

template <int N, bool JAC>
CTBA_HD void frame_pose_eval(const Globals& G, const double* glb, const double* knot_q /*4N*/, const double* knot_p /*3N*/, double u,
                             FramePose<N>* out) {
  So3SplineResult<N> sr;
  so3_spline_eval<N, false, false, JAC>(G.tb, knot_q, u, G.inv_dt, &sr);
  m3 Rwb = quat_matrix(sr.q);
  double c[N];
  spline_coeffs<N>(G.tb.blend, G.tb.base, 0, 1.0, u, c);
  v3 p{0, 0, 0};
#pragma unroll
  for (int i = 0; i < N; ++i) { p.x += c[i] * knot_p[3 * i]; p.y += c[i] * knot_p[3 * i + 1]; p.z += c[i] * knot_p[3 * i + 2]; }
  const double s = G.scale + glb[GLB_SCALE_CORR];
  m3 R0 = quat_matrix(quat{G.q_bc0[0], G.q_bc0[1], G.q_bc0[2], G.q_bc0[3]});
  m3 Rcorr = quat_matrix(quat{glb[GLB_RBC_CORR], glb[GLB_RBC_CORR + 1], glb[GLB_RBC_CORR + 2], glb[GLB_RBC_CORR + 3]});
  m3 Rbc = mul(R0, Rcorr);
  v3 pbc = mul(R0, v3{glb[GLB_PBC_CORR], glb[GLB_PBC_CORR + 1], glb[GLB_PBC_CORR + 2]}) + v3{G.p_bc0[0], G.p_bc0[1], G.p_bc0[2]};
  m3 Rwc = mul(Rwb, Rbc);
  m3 sR = Rwb;
#pragma unroll
  for (int e = 0; e < 9; ++e) sR.a[e] *= s;
  v3 pwc = mul(sR, pbc) + p;
  m3 Rcw = transpose(Rwc);
  m3 mR = Rcw;
#pragma unroll
  for (int e = 0; e < 9; ++e) mR.a[e] *= -1.0;
  v3 tcw = mul(mR, pwc);
#pragma unroll
  for (int e = 0; e < 9; ++e) { out->R_wb[e] = Rwb.a[e]; out->R_cw[e] = Rcw.a[e]; }
  out->p_wb[0] = p.x; out->p_wb[1] = p.y; out->p_wb[2] = p.z;
  out->t_cw[0] = tcw.x; out->t_cw[1] = tcw.y; out->t_cw[2] = tcw.z;
  if (JAC) {
#pragma unroll
    for (int i = 0; i < N; ++i) {
      m3 C = mul(Rwb, sr.dR[i]);
#pragma unroll
      for (int e = 0; e < 9; ++e) out->Crot[i][e] = C.a[e];
      out->cpos[i] = c[i];
    }
  }
}

// One observation: r (2) and Jl = d r / d p_w (2x3, includes inv_std; zero when gated).
template <bool JAC>
CTBA_HD void reproj_obs_eval(const Globals& G, const double* R_cw, const double* t_cw, v3 pw, double u_meas, double v_meas, double inv_std,
                             double* r, double* Jl) {
  v3 pc{R_cw[0] * pw.x + R_cw[1] * pw.y + R_cw[2] * pw.z + t_cw[0], R_cw[3] * pw.x + R_cw[4] * pw.y + R_cw[5] * pw.z + t_cw[1],
        R_cw[6] * pw.x + R_cw[7] * pw.y + R_cw[8] * pw.z + t_cw[2]};
  double uv[2], Jp[6];
  camera_project<JAC>(G.cam, pc, uv, Jp);
  const double e0 = uv[0] - u_meas, e1 = uv[1] - v_meas;
  const double sq = e0 * e0 + e1 * e1;
  if (sq > G.gate_sq) {
    r[0] = 0.0; r[1] = 0.0;
    if (JAC) {
#pragma unroll
      for (int e = 0; e < 6; ++e) Jl[e] = 0.0;
    }
    return;
  }
  r[0] = inv_std * e0; r[1] = inv_std * e1;
  if (JAC) {
#pragma unroll
    for (int rr = 0; rr < 2; ++rr)
#pragma unroll
      for (int cc = 0; cc < 3; ++cc)
        Jl[3 * rr + cc] = inv_std * (Jp[3 * rr] * R_cw[cc] + Jp[3 * rr + 1] * R_cw[3 + cc] + Jp[3 * rr + 2] * R_cw[6 + cc]);
  }
}

// Expand the compact (Jl, frame pose, landmark) form to the dense per-observation local Jacobian row pair
// in the layout of ctba.h: [3N rot | 3N pos | scale 1 | R_bc 3 | p_bc 3 | landmark 3], 2 x (6N+10) row-major.
template <int N>
CTBA_HD void reproj_expand(const Globals& G, const double* glb, const FramePose<N>& fp, v3 pw, const double* Jl, double* J) {
  const int W = 6 * N + 10;
  v3 d{pw.x - fp.p_wb[0], pw.y - fp.p_wb[1], pw.z - fp.p_wb[2]};
  m3 Rwb; for (int e = 0; e < 9; ++e) Rwb.a[e] = fp.R_wb[e];
  m3 Rcw; for (int e = 0; e < 9; ++e) Rcw.a[e] = fp.R_cw[e];
  m3 Rwc = transpose(Rcw);
  const double s = G.scale + glb[GLB_SCALE_CORR];
  m3 R0 = quat_matrix(quat{G.q_bc0[0], G.q_bc0[1], G.q_bc0[2], G.q_bc0[3]});
  v3 pbc = mul(R0, v3{glb[GLB_PBC_CORR], glb[GLB_PBC_CORR + 1], glb[GLB_PBC_CORR + 2]}) + v3{G.p_bc0[0], G.p_bc0[1], G.p_bc0[2]};
  v3 Rp = mul(Rwb, pbc);                       // R_wb p_bc
  v3 pwc = s * Rp + v3{fp.p_wb[0], fp.p_wb[1], fp.p_wb[2]};
  m3 hd = hat(d);
  m3 hc = mul(hat(pw - pwc), Rwc);             // [p_w - p_wc]x R_wc
  m3 RR0 = mul(Rwb, R0);
  for (int rr = 0; rr < 2; ++rr) {
    const double* jl = Jl + 3 * rr;
    double* row = J + rr * W;
    // J_phi = jl * [d]x
    double jphi[3] = {jl[0] * hd.a[0] + jl[1] * hd.a[3] + jl[2] * hd.a[6], jl[0] * hd.a[1] + jl[1] * hd.a[4] + jl[2] * hd.a[7],
                      jl[0] * hd.a[2] + jl[1] * hd.a[5] + jl[2] * hd.a[8]};
    for (int i = 0; i < N; ++i) {
      const double* C = fp.Crot[i];
      for (int cc = 0; cc < 3; ++cc) row[3 * i + cc] = jphi[0] * C[cc] + jphi[1] * C[3 + cc] + jphi[2] * C[6 + cc];
      for (int cc = 0; cc < 3; ++cc) row[3 * N + 3 * i + cc] = -fp.cpos[i] * jl[cc];
    }
    row[6 * N] = -(jl[0] * Rp.x + jl[1] * Rp.y + jl[2] * Rp.z);
    for (int cc = 0; cc < 3; ++cc) row[6 * N + 1 + cc] = jl[0] * hc.a[cc] + jl[1] * hc.a[3 + cc] + jl[2] * hc.a[6 + cc];
    for (int cc = 0; cc < 3; ++cc) row[6 * N + 4 + cc] = -s * (jl[0] * RR0.a[cc] + jl[1] * RR0.a[3 + cc] + jl[2] * RR0.a[6 + cc]);
    for (int cc = 0; cc < 3; ++cc) row[6 * N + 7 + cc] = jl[cc];
  }
}

// IMU sample: gyro + accel share the SO(3) window.
//   gyro  J layout: 3 x (3N rot | 12 gyro-bias | 1 t_corr)              = 3 x (3N+13)
//   accel J layout: 3 x (3N rot | 3N pos | 12 acc-bias | 3 g | 1 t_corr) = 3 x (6N+16)
template <int N, bool JAC>
CTBA_HD void imu_sample_eval(const Globals& G, const double* glb, const double* knot_q, const double* knot_p, const double* bias_a /*12*/,
                             const double* bias_g /*12*/, double st_s, double k, double u_bias, const double* acc_meas, const double* gyr_meas,
                             double inv_std_a, double inv_std_g, double* r_a, double* r_g, double* J_a, double* J_g, size_t ld = 1) {
  // ld: distance between consecutive Jacobian entries of this sample (1: row-major per sample; n: planar [entry][sample], coalesced stores)
  const double t_corr = glb[GLB_T_CAM_IMU_CORR];
  const double st_corr = st_s - t_corr;
  const double u = (st_corr - k * G.dt_s) / G.dt_s;
  So3SplineResult<N> sr;
  so3_spline_eval<N, true, JAC, JAC>(G.tb, knot_q, u, G.inv_dt, &sr);
  double cb[4];
  spline_coeffs<4>(G.tb4.blend, G.tb4.base, 0, 1.0, u_bias, cb);
  v3 ba{0, 0, 0}, bg{0, 0, 0};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    ba.x += cb[i] * bias_a[3 * i]; ba.y += cb[i] * bias_a[3 * i + 1]; ba.z += cb[i] * bias_a[3 * i + 2];
    bg.x += cb[i] * bias_g[3 * i]; bg.y += cb[i] * bias_g[3 * i + 1]; bg.z += cb[i] * bias_g[3 * i + 2];
  }
  // gyro
  r_g[0] = inv_std_g * (sr.w.x - gyr_meas[0] + bg.x);
  r_g[1] = inv_std_g * (sr.w.y - gyr_meas[1] + bg.y);
  r_g[2] = inv_std_g * (sr.w.z - gyr_meas[2] + bg.z);
  // accel
  double c2[N];
  spline_coeffs<N>(G.tb.blend, G.tb.base, 2, G.inv_dt * G.inv_dt, u, c2);
  v3 aw{0, 0, 0};
#pragma unroll
  for (int i = 0; i < N; ++i) { aw.x += c2[i] * knot_p[3 * i]; aw.y += c2[i] * knot_p[3 * i + 1]; aw.z += c2[i] * knot_p[3 * i + 2]; }
  v3 ag = aw + v3{glb[GLB_G], glb[GLB_G + 1], glb[GLB_G + 2]};
  v3 ab = quat_act(quat_inv(sr.q), ag);
  r_a[0] = inv_std_a * (ab.x - acc_meas[0] + ba.x);
  r_a[1] = inv_std_a * (ab.y - acc_meas[1] + ba.y);
  r_a[2] = inv_std_a * (ab.z - acc_meas[2] + ba.z);
  if (JAC) {
    const int WG = 3 * N + 13, WA = 6 * N + 16;
    m3 Rt = transpose(quat_matrix(sr.q));  // R_wb^T
    m3 hab = hat(ab);
    double c3[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, 3, G.inv_dt * G.inv_dt * G.inv_dt, u, c3);
    v3 jw{0, 0, 0};
#pragma unroll
    for (int i = 0; i < N; ++i) { jw.x += c3[i] * knot_p[3 * i]; jw.y += c3[i] * knot_p[3 * i + 1]; jw.z += c3[i] * knot_p[3 * i + 2]; }
    // d/dt [R^T (a+g)] = [R^T(a+g)]x w + R^T jerk ;  d/d t_corr = - d/dt
    v3 dadt = cross(ab, sr.w) + mul(Rt, jw);
#pragma unroll
    for (int i = 0; i < N; ++i) {
      m3 Ja = mul(hab, sr.dR[i]);
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
          J_g[(rr * WG + 3 * i + cc) * ld] = inv_std_g * sr.dW[i].a[3 * rr + cc];
          J_a[(rr * WA + 3 * i + cc) * ld] = inv_std_a * Ja.a[3 * rr + cc];
          J_a[(rr * WA + 3 * N + 3 * i + cc) * ld] = inv_std_a * c2[i] * Rt.a[3 * rr + cc];
        }
    }
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
          J_g[(rr * WG + 3 * N + 3 * i + cc) * ld] = (rr == cc) ? inv_std_g * cb[i] : 0.0;
          J_a[(rr * WA + 6 * N + 3 * i + cc) * ld] = (rr == cc) ? inv_std_a * cb[i] : 0.0;
        }
    const double wd[3] = {sr.wd.x, sr.wd.y, sr.wd.z};
    const double da[3] = {dadt.x, dadt.y, dadt.z};
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) J_a[(rr * WA + 6 * N + 12 + cc) * ld] = inv_std_a * Rt.a[3 * rr + cc];
      J_g[(rr * WG + 3 * N + 12) * ld] = -inv_std_g * wd[rr];
      J_a[(rr * WA + 6 * N + 15) * ld] = -inv_std_a * da[rr];
    }
  }
}

// GP sample. J layout: 3 x (3N rot | 3N pos | p_bp 3 | t_corr 1 | scale_corr 1) = 3 x (6N+5)
template <int N, bool JAC>
CTBA_HD void gp_sample_eval(const Globals& G, const double* glb, const double* knot_q, const double* knot_p, double st_s, double k,
                            const double* meas, double inv_std, double* r, double* J) {
  const double t_corr = glb[GLB_T_GP_CORR];
  const double st_corr = st_s - t_corr;
  const double u = (st_corr - k * G.dt_s) / G.dt_s;
  So3SplineResult<N> sr;
  so3_spline_eval<N, JAC, false, JAC>(G.tb, knot_q, u, G.inv_dt, &sr);
  double c[N];
  spline_coeffs<N>(G.tb.blend, G.tb.base, 0, 1.0, u, c);
  v3 p{0, 0, 0};
#pragma unroll
  for (int i = 0; i < N; ++i) { p.x += c[i] * knot_p[3 * i]; p.y += c[i] * knot_p[3 * i + 1]; p.z += c[i] * knot_p[3 * i + 2]; }
  const double s = G.scale + glb[GLB_SCALE_CORR];
  m3 R = quat_matrix(sr.q);
  v3 pbp{glb[GLB_PBP], glb[GLB_PBP + 1], glb[GLB_PBP + 2]};
  m3 sR = R;
#pragma unroll
  for (int e = 0; e < 9; ++e) sR.a[e] *= s;
  v3 pwp = mul(sR, pbp) + p;
  r[0] = inv_std * (pwp.x - meas[0]); r[1] = inv_std * (pwp.y - meas[1]); r[2] = inv_std * (pwp.z - meas[2]);
  if (JAC) {
    const int W = 6 * N + 5;
    double c1[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, 1, G.inv_dt, u, c1);
    v3 vel{0, 0, 0};
#pragma unroll
    for (int i = 0; i < N; ++i) { vel.x += c1[i] * knot_p[3 * i]; vel.y += c1[i] * knot_p[3 * i + 1]; vel.z += c1[i] * knot_p[3 * i + 2]; }
    // d(s R p_bp)/d(right tangent) = -s R [p_bp]x
    m3 B = mul(sR, hat(pbp));
    v3 Rp = mul(R, pbp);
    v3 ddt = mul(sR, cross(sr.w, pbp)) + vel;  // time derivative of s R p_bp + p
    const double rp[3] = {Rp.x, Rp.y, Rp.z}, dt3[3] = {ddt.x, ddt.y, ddt.z};
#pragma unroll
    for (int i = 0; i < N; ++i) {
      m3 Jr = mul(B, sr.dR[i]);
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
          J[rr * W + 3 * i + cc] = -inv_std * Jr.a[3 * rr + cc];
          J[rr * W + 3 * N + 3 * i + cc] = (rr == cc) ? inv_std * c[i] : 0.0;
        }
    }
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) J[rr * W + 6 * N + cc] = inv_std * sR.a[3 * rr + cc];
      J[rr * W + 6 * N + 3] = -inv_std * dt3[rr];
      J[rr * W + 6 * N + 4] = inv_std * rp[rr];
    }
  }
}

// Pose priors of the spline-fitting tools.  which = 0: PositionCostFunctor (position_cost_functor.h:39-53), r = inv_std (p(u) - meas),
// J 3 x 3N over the position knots;  which = 1: OrientationCostFunctor (orientation_cost_functor.h:39-57),
// r = inv_std * 2 vec(R(u) * meas^-1), J 3 x 3N over the right tangents of the rotation knots.
template <int N, bool JAC>
CTBA_HD void prior_eval(const Globals& G, const double* knot_q, const double* knot_p, int which, double u, const double* meas /*4*/,
                        double inv_std, double* r, double* J) {
  constexpr int W = 3 * N;
  if (which == 0) {
    double c[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, 0, 1.0, u, c);
    v3 p{0, 0, 0};
#pragma unroll
    for (int i = 0; i < N; ++i) { p.x += c[i] * knot_p[3 * i]; p.y += c[i] * knot_p[3 * i + 1]; p.z += c[i] * knot_p[3 * i + 2]; }
    r[0] = inv_std * (p.x - meas[0]); r[1] = inv_std * (p.y - meas[1]); r[2] = inv_std * (p.z - meas[2]);
    if (JAC) {
#pragma unroll
      for (int i = 0; i < N; ++i)
#pragma unroll
        for (int rr = 0; rr < 3; ++rr)
#pragma unroll
          for (int cc = 0; cc < 3; ++cc) J[rr * W + 3 * i + cc] = (rr == cc) ? inv_std * c[i] : 0.0;
    }
    return;
  }
  So3SplineResult<N> sr;
  so3_spline_eval<N, false, false, JAC>(G.tb, knot_q, u, G.inv_dt, &sr);
  const quat mi = quat_inv(quat{meas[0], meas[1], meas[2], meas[3]});
  const quat dr = quat_mul(sr.q, mi);
  r[0] = inv_std * 2.0 * dr.x; r[1] = inv_std * 2.0 * dr.y; r[2] = inv_std * 2.0 * dr.z;
  if (JAC) {
    // d[2 vec(q (e/2, 1) m^-1)]/de = vec(q (e_k, 0) m^-1): plain Hamilton products (a unit-norm product stays unit to first order)
    auto ham = [](quat a, quat b) {
      quat o;
      o.w = a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z;
      o.x = a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y;
      o.y = a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z;
      o.z = a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x;
      return o;
    };
    m3 Je;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const quat e{k == 0 ? 1.0 : 0.0, k == 1 ? 1.0 : 0.0, k == 2 ? 1.0 : 0.0, 0.0};
      const quat o = ham(ham(sr.q, e), mi);
      Je.a[k] = o.x; Je.a[3 + k] = o.y; Je.a[6 + k] = o.z;
    }
#pragma unroll
    for (int i = 0; i < N; ++i) {
      const m3 Ji = mul(Je, sr.dR[i]);
#pragma unroll
      for (int rr = 0; rr < 3; ++rr)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) J[rr * W + 3 * i + cc] = inv_std * Ji.a[3 * rr + cc];
    }
  }
}

// Alignment of a FROZEN spline (expressed in a gauge frame g, e.g. COLMAP's) to the world frame of the global position sensor:
// GlobalPositionSensorWithAlignmentCorrectionCostFunctor (global_position_cost_functor.h:82-150) as wired by
// addWorldColmapAlignmentMeasurement (bspline_so3r3.h:1121-1211: every knot block is set constant).
//   p_w_p = s R_wg (R_gb(u) p_bp + p_gb(u)) + p_wg,  R_wg = R(T_wg) R_corr,  p_wg = t(T_wg) + p_corr,  s = scale + scale_corr,
//   u = (st_s - t_corr - k dt) / dt.
// State x (12): [t_corr | scale_corr | R_corr x y z w | p_corr | p_bp];  tangent (11): [t | scale | R_corr right tangent | p_corr | p_bp].
// e = p_w_p - meas and J = de/d(tangent) are returned UNWEIGHTED (the caller applies inv_std and the loss).
enum { AL_T = 0, AL_SCALE = 1, AL_R = 2, AL_P = 6, AL_PBP = 9, AL_STATE = 12 };
enum { AT_T = 0, AT_SCALE = 1, AT_R = 2, AT_P = 5, AT_PBP = 8, AT_SIZE = 11 };
template <int N, bool JAC>
CTBA_HD void align_eval(const Globals& G, const double* T_wg /*q x y z w | p*/, const double* x, const double* knot_q, const double* knot_p,
                        double st_s, double k, const double* meas, double* e, double* J) {
  const double st_corr = st_s - x[AL_T];
  const double u = (st_corr - k * G.dt_s) / G.dt_s;
  So3SplineResult<N> sr;
  so3_spline_eval<N, JAC, false, false>(G.tb, knot_q, u, G.inv_dt, &sr);
  double c[N];
  spline_coeffs<N>(G.tb.blend, G.tb.base, 0, 1.0, u, c);
  v3 p{0, 0, 0};
#pragma unroll
  for (int i = 0; i < N; ++i) { p.x += c[i] * knot_p[3 * i]; p.y += c[i] * knot_p[3 * i + 1]; p.z += c[i] * knot_p[3 * i + 2]; }
  const double s = G.scale + x[AL_SCALE];
  const m3 R0 = quat_matrix(quat{T_wg[0], T_wg[1], T_wg[2], T_wg[3]});
  const m3 Rc = quat_matrix(quat{x[AL_R], x[AL_R + 1], x[AL_R + 2], x[AL_R + 3]});
  const m3 Rwg = mul(R0, Rc);
  const m3 Rgb = quat_matrix(sr.q);
  const v3 pbp{x[AL_PBP], x[AL_PBP + 1], x[AL_PBP + 2]};
  const v3 xg = mul(Rgb, pbp) + p;             // antenna in the gauge frame
  const v3 y = mul(Rwg, xg);
  e[0] = s * y.x + (T_wg[4] + x[AL_P]) - meas[0];
  e[1] = s * y.y + (T_wg[5] + x[AL_P + 1]) - meas[1];
  e[2] = s * y.z + (T_wg[6] + x[AL_P + 2]) - meas[2];
  if (JAC) {
    double c1[N];
    spline_coeffs<N>(G.tb.blend, G.tb.base, 1, G.inv_dt, u, c1);
    v3 vel{0, 0, 0};
#pragma unroll
    for (int i = 0; i < N; ++i) { vel.x += c1[i] * knot_p[3 * i]; vel.y += c1[i] * knot_p[3 * i + 1]; vel.z += c1[i] * knot_p[3 * i + 2]; }
    const v3 xdot = mul(Rgb, cross(sr.w, pbp)) + vel;   // d xg / d time
    const v3 dt = mul(Rwg, xdot);
    const m3 B = mul(Rwg, hat(xg));                     // d(R_wg Exp(d) xg)/dd = -R_wg [xg]x
    const m3 RR = mul(Rwg, Rgb);
    const double dt3[3] = {dt.x, dt.y, dt.z}, y3[3] = {y.x, y.y, y.z};
#pragma unroll
    for (int rr = 0; rr < 3; ++rr) {
      J[rr * AT_SIZE + AT_T] = -s * dt3[rr];
      J[rr * AT_SIZE + AT_SCALE] = y3[rr];
#pragma unroll
      for (int cc = 0; cc < 3; ++cc) {
        J[rr * AT_SIZE + AT_R + cc] = -s * B.a[3 * rr + cc];
        J[rr * AT_SIZE + AT_P + cc] = (rr == cc) ? 1.0 : 0.0;
        J[rr * AT_SIZE + AT_PBP + cc] = s * RR.a[3 * rr + cc];
      }
    }
  }
}

// Bias smoothness. J layout 3 x 24: [4 knots at s0 | 4 knots at s1]
template <bool JAC>
CTBA_HD void bias_cost_eval(const Globals& G, const double* b0 /*12*/, const double* b1 /*12*/, double u0, double u1, double inv_std, double* r,
                            double* J) {
  double c0[4], c1[4];
  spline_coeffs<4>(G.tb4.blend, G.tb4.base, 0, 1.0, u0, c0);
  spline_coeffs<4>(G.tb4.blend, G.tb4.base, 0, 1.0, u1, c1);
  double v0[3] = {0, 0, 0}, v1[3] = {0, 0, 0};
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int cc = 0; cc < 3; ++cc) { v0[cc] += c0[i] * b0[3 * i + cc]; v1[cc] += c1[i] * b1[3 * i + cc]; }
#pragma unroll
  for (int cc = 0; cc < 3; ++cc) r[cc] = inv_std * (v0[cc] - v1[cc]);
  if (JAC) {
#pragma unroll
    for (int rr = 0; rr < 3; ++rr)
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int cc = 0; cc < 3; ++cc) {
          J[rr * 24 + 3 * i + cc] = (rr == cc) ? inv_std * c0[i] : 0.0;
          J[rr * 24 + 12 + 3 * i + cc] = (rr == cc) ? -inv_std * c1[i] : 0.0;
        }
  }
}

CTBA_HD void gravity_eval(const double* g, double* r, double* J) {
  const double g_norm = 9.80665, weight = 1e6;
  const double n2 = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
  r[0] = weight * (n2 - (g_norm * g_norm));
  if (J) { J[0] = weight * 2.0 * g[0]; J[1] = weight * 2.0 * g[1]; J[2] = weight * 2.0 * g[2]; }
}

}  // namespace ctba
