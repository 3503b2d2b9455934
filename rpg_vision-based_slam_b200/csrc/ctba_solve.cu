// ctba_solve.cu — bordered block-banded Cholesky on the device + the trust-region LM loop.
//
// Replaces ceres::Solve() as configured by the reference (include/spline/bspline_so3r3.h:1768-1782:
// trust-region Levenberg-Marquardt, SPARSE_NORMAL_CHOLESKY, all other options default).  The LM constants and the
// accept/reject/termination logic restate Ceres' defaults [Ceres, recalled — SURVEY.md Appendix B]; the linear system
// is solved exactly (Schur complement on the landmarks + bordered banded Cholesky), so the step equals the step of
// the reference's sparse Cholesky up to rounding.
//
// The factorisation / substitution kernels live in ctba_band.cu (band_factor_solve).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <cstdlib>
#include <vector>
#include "ctba_solver.h"

namespace ctba {

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

void solver_free(SolverData* sd) {
  if (!sd) return;
  band_plan_release(sd);
  pcg_release(sd);
  DBuf<double>* dd[] = {&sd->band0, &sd->bandw, &sd->border0, &sd->borderw, &sd->corner0, &sd->cornerw, &sd->grad, &sd->tot, &sd->V0, &sd->gl, &sd->Sp, &sd->Uk, &sd->Sb,
                        &sd->rp, &sd->Vi, &sd->vig, &sd->ViYg, &sd->Yo, &sd->Zo, &sd->lm_corner_partial, &sd->jscale, &sd->jscale_lm, &sd->delta, &sd->delta_lm, &sd->dpose, &sd->scal, &sd->cand.knot_q, &sd->cand.knot_p,
                        &sd->cand.bias_a, &sd->cand.bias_g, &sd->cand.lm, &sd->cand.glb};
  for (auto* d : dd) d->release();
  DBuf<int32_t>* di[] = {&sd->d_node_of_knot, &sd->d_node_of_bias, &sd->d_lm_obs_begin, &sd->d_lm_obs, &sd->d_o_pos, &sd->d_o_end, &sd->d_imu_seg, &sd->d_gp_seg, &sd->d_prior_seg, &sd->d_lm_frame, &sd->d_kf_begin, &sd->d_kf_list, &sd->flag, &sd->d_lm_owner};
  for (auto* d : di) d->release();
  sd->d_touch_knot.release(); sd->d_touch_bias.release(); sd->d_touch_lm.release(); sd->d_o_mask.release();
  delete sd;
}

int solver_num_unknowns(ctba_problem* p) {
  const bool has_imu = !p->i_s.empty() || !p->b_s0.empty();
  return 6 * ((int)p->numKnots() + (has_imu ? (int)p->numBiasKnots() : 0));
}

static void swap_state(DeviceState& a, DeviceState& b) {
  std::swap(a.knot_q, b.knot_q); std::swap(a.knot_p, b.knot_p); std::swap(a.bias_a, b.bias_a); std::swap(a.bias_g, b.bias_g);
  std::swap(a.lm, b.lm); std::swap(a.glb, b.glb);
}

// Phase timers that never stall the stream: a scope records two events from a small pool and the elapsed time is read back at the next
// point where the LM loop synchronises anyway (EvTimers::flush after the read-back of the step scalars / of the candidate cost).  The
// loop's only host round trips are the ones its decisions need: [factorisation flag + step scalars], [candidate cost], [gradient norm of
// the accepted state].
struct EvTimers {
  static constexpr int kPairs = 16;
  cudaEvent_t a[kPairs], b[kPairs]; double* acc[kPairs]; int n = 0, made = 0; cudaStream_t st;
  explicit EvTimers(cudaStream_t s) : st(s) {}
  ~EvTimers() { for (int i = 0; i < made; ++i) { cudaEventDestroy(a[i]); cudaEventDestroy(b[i]); } }
  int begin(double* t) {
    if (n == kPairs) { cudaEventSynchronize(b[n - 1]); flush(); }
    if (n == made) { cudaEventCreate(&a[made]); cudaEventCreate(&b[made]); ++made; }
    acc[n] = t; cudaEventRecord(a[n], st); return n++;
  }
  void end(int i) { cudaEventRecord(b[i], st); }
  void flush() {   // call after a stream synchronisation: every recorded pair has completed
    for (int i = 0; i < n; ++i) { float ms = 0; if (cudaEventElapsedTime(&ms, a[i], b[i]) == cudaSuccess) *acc[i] += ms * 1e-3; }
    n = 0;
  }
};

int solver_solve(ctba_problem* p, int max_num_iterations, ctba_summary* out) {
  // ceres::Solver::Options defaults [Ceres, recalled]
  const double initial_radius = 1e4, max_radius = 1e16, min_radius = 1e-32, min_relative_decrease = 1e-3;
  const double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  const int max_consecutive_invalid = 5;
  auto t_begin = std::chrono::steady_clock::now();
  const int64_t launches0 = p->launches;
  out->time_eval_s = out->time_build_s = out->time_solve_s = 0.0;
  if (!p->solver || !p->solver->built || p->struct_dirty) {
    int rc0 = assemble_structure(p); if (rc0) return rc0;
    p->struct_dirty = false;
  }
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  int rc;
  double host_cost[8], host_sc[SC_COUNT];
  int32_t host_flag[4];
  int nit = 0;
  auto log_it = [&](double cost, double cc, double gmax, double sn, double ratio, double radius, int ok) {
    if (out->iterations && nit < out->iterations_capacity) out->iterations[nit] = {cost, cc, gmax, sn, ratio, radius, ok};
    ++nit;
  };
  EvTimers timers(st);
  auto eval_at = [&](EvalBuffers& ev, const StatePtrs& S, bool jac, double* cost) -> int {
    const int t = timers.begin(&out->time_eval_s);
    int r = launch_evaluate(p, ev, S, jac); if (r) return r;
    r = launch_cost(p, ev); if (r) return r;
    timers.end(t);
    r = shard_reduce_cost(p, ev, host_cost); if (r) return r;   // single GPU: plain read-back of the 7 cost terms (synchronises the stream)
    timers.flush();
    *cost = host_cost[6];
    return CTBA_OK;
  };
  auto build_at = [&]() -> int {  // direct normal equations + gradient at the current state (evaluation buffers hold its Jacobians)
    const int t = timers.begin(&out->time_build_s);
    int r = assemble_direct(p);
    timers.end(t);
    return r;
  };
  auto scalars = [&](const StatePtrs& S, const StatePtrs& C, double radius) -> int {
    int r = step_scalars(p, S, C, radius); if (r) return r;
    if (!p->sharded) {
      CTBA_CUDA(p, cudaMemcpyAsync(host_sc, sd->scal.p, SC_COUNT * sizeof(double), cudaMemcpyDeviceToHost, st));
      CTBA_CUDA(p, cudaStreamSynchronize(st));
      timers.flush();
      return CTBA_OK;
    }
    // sharded: every rank holds the sums over the nodes / landmarks it owns; combined in rank order so that all ranks take the same decisions
    const int W = p->comm->world;
    CTBA_CUDA(p, p->d_gather.resize((size_t)W * SC_COUNT));
    if (p->comm->allgather(sd->scal.p, p->d_gather.p, SC_COUNT, st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
    std::vector<double> h((size_t)W * SC_COUNT);
    CTBA_CUDA(p, cudaMemcpyAsync(h.data(), p->d_gather.p, h.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));
    timers.flush();
    for (int c = 0; c < SC_COUNT; ++c) host_sc[c] = 0.0;
    for (int rk = 0; rk < W; ++rk) {
      host_sc[SC_MODEL] += h[(size_t)rk * SC_COUNT + SC_MODEL]; host_sc[SC_STEP2] += h[(size_t)rk * SC_COUNT + SC_STEP2];
      host_sc[SC_XNORM2] += h[(size_t)rk * SC_COUNT + SC_XNORM2]; host_sc[SC_GMAX] = std::max(host_sc[SC_GMAX], h[(size_t)rk * SC_COUNT + SC_GMAX]);
    }
    return CTBA_OK;
  };

  double x_cost = 0.0;
  rc = eval_at(p->ev, p->dstate.ptrs(), true, &x_cost); if (rc) return rc;
  rc = build_at(); if (rc) return rc;
  rc = compute_jscale(p); if (rc) return rc;   // Jacobi scaling is fixed at iteration 0 of each Solve
  rc = scalars(p->dstate.ptrs(), p->dstate.ptrs(), -1.0); if (rc) return rc;
  double x_norm = std::sqrt(host_sc[SC_XNORM2]), gmax = host_sc[SC_GMAX];
  out->initial_cost = x_cost;
  out->num_successful_steps = out->num_unsuccessful_steps = 0;
  out->termination_type = 1;
  log_it(x_cost, 0.0, gmax, 0.0, 0.0, initial_radius, 1);
  double radius = initial_radius, decrease_factor = 2.0;
  int invalid = 0, iter = 0;
  if (!std::isfinite(x_cost)) { out->termination_type = 2; goto done; }
  if (!p->bench_no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; goto done; }
  while (true) {
    if (iter >= max_num_iterations) { out->termination_type = 1; break; }
    if (radius < min_radius) { if (p->bench_no_termination) radius = initial_radius; else { out->termination_type = 0; break; } }
    ++iter;
    {
      const int tb = timers.begin(&out->time_build_s);
      rc = assemble_schur(p, radius);
      timers.end(tb);
      if (rc) return rc;
    }
    {
      const int ts = timers.begin(&out->time_solve_s);
      rc = p->linear_solver ? band_pcg_solve(p) : band_factor_solve(p);
      if (rc) return rc;
      CTBA_CUDA(p, cudaMemcpyAsync(host_flag, sd->flag.p, sizeof(host_flag), cudaMemcpyDeviceToHost, st));
      timers.end(ts);
    }
    // The candidate and the step scalars are enqueued without waiting for the factorisation flag: one round trip reads both.  After a
    // failed factorisation they are garbage that is never used (the candidate buffers are only swapped in by an accepted step).
    rc = backsub_and_candidate(p, p->dstate.ptrs(), sd->cand.ptrs(), radius); if (rc) return rc;
    rc = scalars(p->dstate.ptrs(), sd->cand.ptrs(), radius); if (rc) return rc;
    bool ok = host_flag[0] == 0;
    double model_cost_change = 0.0, step_norm = 0.0;
    if (ok) {
      model_cost_change = host_sc[SC_MODEL];
      step_norm = std::sqrt(host_sc[SC_STEP2]);
      if (!std::isfinite(model_cost_change) || !std::isfinite(step_norm)) ok = false;
    }
    if (!ok || model_cost_change <= 0.0) {
      if (getenv("CTBA_DEBUG")) fprintf(stderr, "[ctba %p] iter %d invalid step: flag %d %d model %.6g step %.6g radius %.3g\n", (void*)p, iter, host_flag[0], host_flag[1], model_cost_change, step_norm, radius);
      // HandleInvalidStep [Ceres, recalled]: the max_num_consecutive_invalid_steps-th invalid step in a row ends the solve with FAILURE
      if (++invalid >= max_consecutive_invalid && !p->bench_no_termination) { out->termination_type = 2; break; }
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, 0.0, gmax, 0.0, 0.0, radius, 0);
      continue;
    }
    invalid = 0;
    double cand_cost = 0.0;
    // residuals AND Jacobians at the candidate go to the alternate buffers: an accepted step only swaps them in
    rc = eval_at(p->ev_alt, sd->cand.ptrs(), true, &cand_cost); if (rc) return rc;
    const double cost_change = x_cost - cand_cost;
    const double ratio = cost_change / model_cost_change;
    // ParameterToleranceReached / FunctionToleranceReached are tested before accept/reject (Ceres 2.x order; same in the oracle)
    if (!p->bench_no_termination && step_norm <= parameter_tolerance * (x_norm + parameter_tolerance)) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (!p->bench_no_termination && std::fabs(cost_change) <= function_tolerance * x_cost) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (std::isfinite(cand_cost) && (ratio > min_relative_decrease || p->bench_accept_all)) {
      swap_state(p->dstate, sd->cand);
      std::swap(p->ev, p->ev_alt);
      p->host_state_stale = true;
      x_cost = cand_cost;
      rc = build_at(); if (rc) return rc;
      rc = scalars(p->dstate.ptrs(), p->dstate.ptrs(), -1.0); if (rc) return rc;
      x_norm = std::sqrt(host_sc[SC_XNORM2]); gmax = host_sc[SC_GMAX];
      if (ratio > min_relative_decrease) { radius = std::min(max_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3))); decrease_factor = 2.0; }
      else { radius /= decrease_factor; decrease_factor *= 2.0; }   // bench_accept_all only
      out->num_successful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 1);
      if (!p->bench_no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; break; }
    } else {
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
    }
  }
done:
  rc = shard_gather_state(p); if (rc) return rc;   // sharded: the owners' knots / biases / landmarks reach every rank
  out->final_cost = x_cost;
  out->num_iterations = nit;
  out->kernel_launches = p->launches - launches0;
  CTBA_CUDA(p, cudaStreamSynchronize(st));
  timers.flush();
  out->time_total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  return CTBA_OK;
}

}  // namespace ctba
