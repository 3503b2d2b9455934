// ctba_align.h — alignment of a FROZEN spline to the world frame of the global position sensor: the problem that
// src/align_spline_to_global_frame.cpp:153-226 builds with addWorldColmapAlignmentMeasurement (include/spline/bspline_so3r3.h:1121-1211)
// and hands to BSplineSO3R3::optimize.  Every knot block is constant, so the whole problem has at most 11 unknowns
//   [ t_s_gp_corr | scale_corr | R_w_g_corr (right tangent, 3) | p_w_g_corr (3) | p_bp (3) ]
// (t_s_gp_corr and p_bp only when their optimise flags are set, :1201-1209) and n x 3 residuals, each 3-block under CauchyLoss(0.10).
//
// The pieces here are CTBA_HD so that the kernels of ctba_align.cu and the tests' host harness share them:
//   align_sample_accumulate : one measurement -> its loss-corrected contribution to the packed 11x11 normal equations
//   align_lm_step           : one trust-region step on the reduced (<= 11) system: Jacobi scaling, LM diagonal, Cholesky, candidate state,
//                             model cost change, |step|, |x|, max-norm of the gradient  (same constants/ordering as ctba_solve.cu)
#pragma once
#include <cmath>
#include "../../include/ctba.h"
#include "ctba_terms.h"

namespace ctba {

// packed system of one evaluation: upper triangle of J^T J (66, row-major i <= j), J^T r (11), cost (1)
enum { AS_H = 0, AS_G = 66, AS_COST = 77, AS_SIZE = 78 };
CTBA_HD int as_h(int i, int j) { return i * AT_SIZE - (i * (i - 1)) / 2 + (j - i); }   // i <= j

// scalars of one LM step (device -> host)
enum { AK_OK = 0, AK_MODEL = 1, AK_STEP = 2, AK_XNORM = 3, AK_GMAX = 4, AK_SIZE = 8 };

template <int N>
CTBA_HD void align_sample_accumulate(const Globals& G, const double* T_wg, const double* x, const double* kq, const double* kp, double st_s,
                                     double k, const double* meas, double inv_std, double cauchy_b, bool jac, double* acc /*AS_SIZE*/) {
  double e[3], J[3 * AT_SIZE];
  if (jac) align_eval<N, true>(G, T_wg, x, kq, kp, st_s, k, meas, e, J);
  else align_eval<N, false>(G, T_wg, x, kq, kp, st_s, k, meas, e, J);
  const double r0 = inv_std * e[0], r1 = inv_std * e[1], r2 = inv_std * e[2];
  const double s = r0 * r0 + r1 * r1 + r2 * r2;
  double w = inv_std;
  if (cauchy_b > 0.0) {                       // ceres::CauchyLoss: rho = b log(1 + s/b), rho' = 1 / (1 + s/b); rho'' < 0 => r, J scaled by sqrt(rho')
    const double sum = 1.0 + s / cauchy_b;
    acc[AS_COST] += 0.5 * cauchy_b * log(sum);
    w *= sqrt(1.0 / sum);
  } else {
    acc[AS_COST] += 0.5 * s;
  }
  if (!jac) return;
#pragma unroll
  for (int rr = 0; rr < 3; ++rr) {
    const double rw = w * e[rr];
    double row[AT_SIZE];
#pragma unroll
    for (int c = 0; c < AT_SIZE; ++c) row[c] = w * J[rr * AT_SIZE + c];
#pragma unroll
    for (int i = 0; i < AT_SIZE; ++i) {
      acc[AS_G + i] += row[i] * rw;
#pragma unroll
      for (int j = i; j < AT_SIZE; ++j) acc[AS_H + as_h(i, j)] += row[i] * row[j];
    }
  }
}

// max | x - Plus(x, -g) | over the ambient coordinates of a rotation block (what Ceres' gradient max-norm sees through the manifold)
CTBA_HD double align_gmax_quat(quat q, v3 g) {
  quat qn = quat_mul(q, so3_exp(v3{-g.x, -g.y, -g.z}));
  return fmax(fmax(fabs(qn.x - q.x), fabs(qn.y - q.y)), fmax(fabs(qn.z - q.z), fabs(qn.w - q.w)));
}

// One LM step on the packed system `sys` at state x (AL_STATE doubles).  free_mask: bit c set = tangent direction c is optimised.
//   radius < 0 : no step; only jscale (if init_jscale), |x| and the gradient max-norm are produced.
// jscale[AT_SIZE]: Jacobi scaling 1 / (1 + sqrt(H_cc)), fixed at iteration 0 of a Solve (written when init_jscale).
// Outputs: cand[AL_STATE] = Plus(x, step);  sc[AK_*].
CTBA_HD void align_lm_step(const double* sys, const double* x, unsigned free_mask, double radius, bool init_jscale, double* jscale,
                           double* cand, double* sc) {
  const double min_lm_diagonal = 1e-6, max_lm_diagonal = 1e32;   // ceres::Solver::Options defaults [Ceres, recalled]
  int col[AT_SIZE], n = 0;
  for (int c = 0; c < AT_SIZE; ++c) if (free_mask >> c & 1u) col[n++] = c;
  if (init_jscale)
    for (int c = 0; c < AT_SIZE; ++c) jscale[c] = 1.0 / (1.0 + sqrt(sys[AS_H + as_h(c, c)]));
  // |x| over the ambient coordinates of the free blocks, gradient max-norm
  double xn2 = 0.0, gmx = 0.0;
  const quat qc{x[AL_R], x[AL_R + 1], x[AL_R + 2], x[AL_R + 3]};
  if (free_mask >> AT_T & 1u) { xn2 += x[AL_T] * x[AL_T]; gmx = fmax(gmx, fabs(sys[AS_G + AT_T])); }
  if (free_mask >> AT_SCALE & 1u) { xn2 += x[AL_SCALE] * x[AL_SCALE]; gmx = fmax(gmx, fabs(sys[AS_G + AT_SCALE])); }
  if (free_mask >> AT_R & 1u) {
    xn2 += qc.x * qc.x + qc.y * qc.y + qc.z * qc.z + qc.w * qc.w;
    gmx = fmax(gmx, align_gmax_quat(qc, v3{sys[AS_G + AT_R], sys[AS_G + AT_R + 1], sys[AS_G + AT_R + 2]}));
  }
  for (int c = 0; c < 3; ++c) {
    if (free_mask >> (AT_P + c) & 1u) { xn2 += x[AL_P + c] * x[AL_P + c]; gmx = fmax(gmx, fabs(sys[AS_G + AT_P + c])); }
    if (free_mask >> (AT_PBP + c) & 1u) { xn2 += x[AL_PBP + c] * x[AL_PBP + c]; gmx = fmax(gmx, fabs(sys[AS_G + AT_PBP + c])); }
  }
  sc[AK_XNORM] = sqrt(xn2); sc[AK_GMAX] = gmx;
  sc[AK_OK] = 1.0; sc[AK_MODEL] = 0.0; sc[AK_STEP] = 0.0;
  for (int i = 0; i < AL_STATE; ++i) cand[i] = x[i];
  if (radius < 0.0) return;

  // scaled system Hs = S H S, rhs = -S g;  A = Hs + diag(clamp(diag Hs)) / radius
  double A[AT_SIZE * AT_SIZE], Hs[AT_SIZE * AT_SIZE], rhs[AT_SIZE], delta[AT_SIZE];
  for (int i = 0; i < n; ++i) {
    for (int j = 0; j < n; ++j) {
      const int a = col[i] < col[j] ? col[i] : col[j], b = col[i] < col[j] ? col[j] : col[i];
      Hs[i * AT_SIZE + j] = sys[AS_H + as_h(a, b)] * jscale[col[i]] * jscale[col[j]];
      A[i * AT_SIZE + j] = Hs[i * AT_SIZE + j];
    }
    rhs[i] = -sys[AS_G + col[i]] * jscale[col[i]];
  }
  for (int i = 0; i < n; ++i) {
    const double d = fmin(fmax(Hs[i * AT_SIZE + i], min_lm_diagonal), max_lm_diagonal);
    A[i * AT_SIZE + i] += d / radius;
  }
  // in-place lower Cholesky
  bool ok = true;
  for (int j = 0; j < n && ok; ++j) {
    double d = A[j * AT_SIZE + j];
    for (int k = 0; k < j; ++k) d -= A[j * AT_SIZE + k] * A[j * AT_SIZE + k];
    if (!(d > 0.0)) { ok = false; break; }
    d = sqrt(d);
    A[j * AT_SIZE + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = A[i * AT_SIZE + j];
      for (int k = 0; k < j; ++k) s -= A[i * AT_SIZE + k] * A[j * AT_SIZE + k];
      A[i * AT_SIZE + j] = s / d;
    }
  }
  if (!ok) { sc[AK_OK] = 0.0; return; }
  for (int i = 0; i < n; ++i) { double s = rhs[i]; for (int k = 0; k < i; ++k) s -= A[i * AT_SIZE + k] * delta[k]; delta[i] = s / A[i * AT_SIZE + i]; }
  for (int i = n - 1; i >= 0; --i) { double s = delta[i]; for (int k = i + 1; k < n; ++k) s -= A[k * AT_SIZE + i] * delta[k]; delta[i] = s / A[i * AT_SIZE + i]; }
  // model cost change = -delta^T (gs + 1/2 Hs delta)
  double accm = 0.0;
  for (int i = 0; i < n; ++i) {
    double hd = 0.0;
    for (int j = 0; j < n; ++j) hd += Hs[i * AT_SIZE + j] * delta[j];
    accm += delta[i] * (-rhs[i] + 0.5 * hd);
    if (!(fabs(delta[i]) <= 1.7976931348623157e308)) ok = false;   // NaN or Inf
  }
  sc[AK_MODEL] = -accm;
  if (!ok) { sc[AK_OK] = 0.0; return; }
  // candidate = Plus(x, S delta)
  double du[AT_SIZE];
  for (int c = 0; c < AT_SIZE; ++c) du[c] = 0.0;
  for (int i = 0; i < n; ++i) du[col[i]] = delta[i] * jscale[col[i]];
  cand[AL_T] = x[AL_T] + du[AT_T];
  cand[AL_SCALE] = x[AL_SCALE] + du[AT_SCALE];
  if (free_mask >> AT_R & 1u) {
    const quat qn = quat_mul(qc, so3_exp(v3{du[AT_R], du[AT_R + 1], du[AT_R + 2]}));
    cand[AL_R] = qn.x; cand[AL_R + 1] = qn.y; cand[AL_R + 2] = qn.z; cand[AL_R + 3] = qn.w;
  }
  for (int c = 0; c < 3; ++c) { cand[AL_P + c] = x[AL_P + c] + du[AT_P + c]; cand[AL_PBP + c] = x[AL_PBP + c] + du[AT_PBP + c]; }
  double st2 = 0.0;
  for (int i = 0; i < AL_STATE; ++i) { const double d = cand[i] - x[i]; st2 += d * d; }
  sc[AK_STEP] = sqrt(st2);
}

// The trust-region loop of ceres::Solve() for this problem: same constants, tests and ordering as solver_solve (ctba_solve.cu), over a
// backend that owns the state and the packed systems (the device in ctba_align.cu; a serial host loop in the tests' harness):
//   int  evaluate(bool candidate, double* cost)   packed system (with Jacobians) at the current or at the candidate state
//   int  step(double radius, bool init_jscale, double* sc /*AK_SIZE*/)   align_lm_step at the current state -> candidate
//   void accept()                                  candidate state and its system become current
template <class Backend>
int align_lm_loop(Backend& be, int max_num_iterations, bool no_termination, ctba_summary* out) {
  const double initial_radius = 1e4, max_radius = 1e16, min_radius = 1e-32, min_relative_decrease = 1e-3;
  const double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  const int max_consecutive_invalid = 5;
  int nit = 0, rc;
  auto log_it = [&](double cost, double cc, double gmax, double sn, double ratio, double radius, int ok) {
    if (out->iterations && nit < out->iterations_capacity) out->iterations[nit] = {cost, cc, gmax, sn, ratio, radius, ok};
    ++nit;
  };
  double sc[AK_SIZE];
  double x_cost = 0.0;
  rc = be.evaluate(false, &x_cost); if (rc) return rc;
  rc = be.step(-1.0, true, sc); if (rc) return rc;
  double x_norm = sc[AK_XNORM], gmax = sc[AK_GMAX];
  out->initial_cost = x_cost;
  out->num_successful_steps = out->num_unsuccessful_steps = 0;
  out->termination_type = 1;
  log_it(x_cost, 0.0, gmax, 0.0, 0.0, initial_radius, 1);
  double radius = initial_radius, decrease_factor = 2.0;
  int invalid = 0, iter = 0;
  bool stop = false;
  if (!std::isfinite(x_cost)) { out->termination_type = 2; stop = true; }
  else if (!no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; stop = true; }
  while (!stop) {
    if (iter >= max_num_iterations) { out->termination_type = 1; break; }
    if (radius < min_radius) { if (no_termination) radius = initial_radius; else { out->termination_type = 0; break; } }
    ++iter;
    rc = be.step(radius, false, sc); if (rc) return rc;
    const double model_cost_change = sc[AK_MODEL], step_norm = sc[AK_STEP];
    const bool ok = sc[AK_OK] != 0.0 && std::isfinite(model_cost_change) && std::isfinite(step_norm);
    if (!ok || model_cost_change <= 0.0) {
      if (++invalid >= max_consecutive_invalid && !no_termination) { out->termination_type = 2; break; }   // HandleInvalidStep [Ceres, recalled]
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, 0.0, gmax, 0.0, 0.0, radius, 0);
      continue;
    }
    invalid = 0;
    double cand_cost = 0.0;
    rc = be.evaluate(true, &cand_cost); if (rc) return rc;
    const double cost_change = x_cost - cand_cost;
    const double ratio = cost_change / model_cost_change;
    if (!no_termination && step_norm <= parameter_tolerance * (x_norm + parameter_tolerance)) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (!no_termination && std::fabs(cost_change) <= function_tolerance * x_cost) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (std::isfinite(cand_cost) && ratio > min_relative_decrease) {
      be.accept();
      x_cost = cand_cost;
      rc = be.step(-1.0, false, sc); if (rc) return rc;
      x_norm = sc[AK_XNORM]; gmax = sc[AK_GMAX];
      radius = std::fmin(max_radius, radius / std::fmax(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3)));
      decrease_factor = 2.0;
      out->num_successful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 1);
      if (!no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; break; }
    } else {
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
    }
  }
  out->final_cost = x_cost;
  out->num_iterations = nit;
  return CTBA_OK;
}

}  // namespace ctba
