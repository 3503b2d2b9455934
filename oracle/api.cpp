// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
// C entry points (ctypes) over the CPU restatement.  Used only by tests/, __graft_entry__.smoke()
// and bench.py's cpu_baseline / --impl reference legs.  "parity unpinned": the reference repo holds
// no tests or golden vectors for this path and Ceres/Eigen are absent, so the only
// external anchors are the Sophus SO3 fixtures and the closed-form spline constants (tests/).
#include <chrono>
#include <cstdint>
#include <cstring>
#include "lm.h"
#include "sparse_lm.h"
#include "dt_problem.h"
#include "select.h"
#include "sample.h"

using namespace oracle;

extern "C" {

void* orc_create(int order, int64_t dt_ns, int64_t start_t_ns) {
  if (order != 4 && order != 5 && order != 6) return nullptr;
  return new SplineProblem(order, dt_ns, start_t_ns);
}
void orc_destroy(void* h) { delete (SplineProblem*)h; }

int orc_set_knots(void* h, int n, const double* quat, const double* pos) {
  auto* p = (SplineProblem*)h;
  p->so3.assign(quat, quat + 4 * (size_t)n); p->r3.assign(pos, pos + 3 * (size_t)n); p->built = false; return 0;
}
int orc_num_knots(void* h) { return (int)((SplineProblem*)h)->numKnots(); }
int orc_get_knots(void* h, double* quat, double* pos) {
  auto* p = (SplineProblem*)h;
  std::memcpy(quat, p->so3.data(), sizeof(double) * p->so3.size()); std::memcpy(pos, p->r3.data(), sizeof(double) * p->r3.size()); return 0;
}
int orc_initialize_biases(void* h, double dt_s_bias, const double* a0, const double* g0) {
  auto* p = (SplineProblem*)h; p->initializeBiases(dt_s_bias, a0, g0); p->built = false; return (int)p->numBiasKnots();
}
int orc_num_bias_knots(void* h) { return (int)((SplineProblem*)h)->numBiasKnots(); }
int orc_set_biases(void* h, const double* acc, const double* gyr) {
  auto* p = (SplineProblem*)h;
  std::memcpy(p->acc_bias.data(), acc, sizeof(double) * p->acc_bias.size()); std::memcpy(p->gyr_bias.data(), gyr, sizeof(double) * p->gyr_bias.size()); return 0;
}
int orc_get_biases(void* h, double* acc, double* gyr) {
  auto* p = (SplineProblem*)h;
  std::memcpy(acc, p->acc_bias.data(), sizeof(double) * p->acc_bias.size()); std::memcpy(gyr, p->gyr_bias.data(), sizeof(double) * p->gyr_bias.size()); return 0;
}
int orc_set_landmarks(void* h, int n, const double* xyz) {
  auto* p = (SplineProblem*)h; p->lm.assign(xyz, xyz + 3 * (size_t)n); p->built = false; return 0;
}
int orc_get_landmarks(void* h, double* xyz) {
  auto* p = (SplineProblem*)h; std::memcpy(xyz, p->lm.data(), sizeof(double) * p->lm.size()); return 0;
}
int orc_set_camera(void* h, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc, double t_offset_cam_imu) {
  auto* p = (SplineProblem*)h;
  p->cam.model = model; p->cam.fx = intr[0]; p->cam.fy = intr[1]; p->cam.cx = intr[2]; p->cam.cy = intr[3];
  for (int i = 0; i < 4; ++i) { p->cam.d[i] = dist[i]; p->cam.q_bc[i] = q_bc[i]; }
  for (int i = 0; i < 3; ++i) p->cam.p_bc[i] = p_bc[i];
  p->cam_t_offset_cam_imu = t_offset_cam_imu;
  p->t_offset_cam_imu_s = t_offset_cam_imu;  // setCameras, bspline_so3r3.h:1899-1903
  return 0;
}
int orc_set_globals(void* h, double scale, const double* p_bp, const double* g, double t_s_gp_s) {
  auto* p = (SplineProblem*)h;
  p->scale = scale; for (int i = 0; i < 3; ++i) { p->p_bp[i] = p_bp[i]; p->g[i] = g[i]; }
  p->t_s_gp_init_s = t_s_gp_s; p->t_s_gp_s = t_s_gp_s; return 0;
}
// corr[0]=scale_corr, [1..4]=R_bc_corr xyzw, [5..7]=p_bc_corr, [8]=t_offset_cam_imu_corr, [9]=t_s_gp_corr
int orc_set_corrections(void* h, const double* c) {
  auto* p = (SplineProblem*)h;
  p->scale_corr = c[0]; for (int i = 0; i < 4; ++i) p->R_bc_corr[i] = c[1 + i]; for (int i = 0; i < 3; ++i) p->p_bc_corr[i] = c[5 + i];
  p->t_offset_cam_imu_corr_s = c[8]; p->t_s_gp_corr_s = c[9]; return 0;
}
// out: corr[0..9] as above, [10..12]=p_bp, [13..15]=g, [16]=t_offset_cam_imu_s, [17]=t_s_gp_s, [18]=scale
int orc_get_globals(void* h, double* o) {
  auto* p = (SplineProblem*)h;
  o[0] = p->scale_corr; for (int i = 0; i < 4; ++i) o[1 + i] = p->R_bc_corr[i]; for (int i = 0; i < 3; ++i) o[5 + i] = p->p_bc_corr[i];
  o[8] = p->t_offset_cam_imu_corr_s; o[9] = p->t_s_gp_corr_s;
  for (int i = 0; i < 3; ++i) { o[10 + i] = p->p_bp[i]; o[13 + i] = p->g[i]; }
  o[16] = p->t_offset_cam_imu_s; o[17] = p->t_s_gp_s; o[18] = p->scale; return 0;
}
int orc_set_flags(void* h, int opt_scale, int opt_pbp, int opt_gravity, int opt_camimu, int opt_t_gs) {
  auto* p = (SplineProblem*)h;
  p->optimize_scale = opt_scale; p->optimize_p_bp = opt_pbp; p->optimize_gravity = opt_gravity;
  p->optimize_camimu_extrinsics = opt_camimu; p->optimize_t_offset_s_gs = opt_t_gs; p->built = false; return 0;
}
int orc_add_frames(void* h, int F, const int64_t* t_ns, const int32_t* obs_begin, const double* uv, const int32_t* lm, double noise) {
  return ((SplineProblem*)h)->addFrames(F, t_ns, obs_begin, uv, lm, noise);
}
int orc_add_imu(void* h, int n, const int64_t* t_ns, const double* acc, const double* gyr, double acc_nd, double gyr_nd) {
  return ((SplineProblem*)h)->addImu(n, t_ns, acc, gyr, acc_nd, gyr_nd);
}
int orc_add_gp(void* h, int n, const int64_t* t_ns, const double* pos, double noise) { return ((SplineProblem*)h)->addGp(n, t_ns, pos, noise); }
int orc_add_priors(void* h, int which, int n, const int64_t* t_ns, const double* meas, double noise) { return ((SplineProblem*)h)->addPriors(which, n, t_ns, meas, noise); }
int orc_num_priors(void* h) { return (int)((SplineProblem*)h)->priors.size(); }
// world alignment of a frozen spline (align_spline_to_global_frame.cpp:153-226)
int orc_add_align(void* h, int n, const int64_t* t_ns, const double* pos, double noise) { return ((SplineProblem*)h)->addAlign(n, t_ns, pos, noise); }
int orc_num_align(void* h) { return (int)((SplineProblem*)h)->align.size(); }
int orc_set_alignment(void* h, const double* T_wg, const double* corr) {   // setWorldColmapTransformation :1825 (+ the corrections, for tests)
  auto* p = (SplineProblem*)h;
  for (int i = 0; i < 7; ++i) p->T_wg[i] = T_wg[i];
  if (corr) { for (int i = 0; i < 4; ++i) p->R_wg_corr[i] = corr[i]; for (int i = 0; i < 3; ++i) p->p_wg_corr[i] = corr[4 + i]; }
  p->built = false; return 0;
}
// T (7): getWorldColmapTransformation :1827-1834 = (R(T_wg) R_corr, t(T_wg) + p_corr);  corr (7): getWorldColmapTransformationCorr :1836-1840
int orc_get_alignment(void* h, double* T, double* corr) {
  auto* p = (SplineProblem*)h;
  SO3<double> q0{p->T_wg[0], p->T_wg[1], p->T_wg[2], p->T_wg[3]}, qc{p->R_wg_corr[0], p->R_wg_corr[1], p->R_wg_corr[2], p->R_wg_corr[3]};
  SO3<double> q = q0 * qc;
  T[0] = q.x; T[1] = q.y; T[2] = q.z; T[3] = q.w;
  for (int i = 0; i < 3; ++i) T[4 + i] = p->T_wg[4 + i] + p->p_wg_corr[i];
  for (int i = 0; i < 4; ++i) corr[i] = p->R_wg_corr[i];
  for (int i = 0; i < 3; ++i) corr[4 + i] = p->p_wg_corr[i];
  return 0;
}
int orc_add_bias_cost(void* h, double freq_hz, double dt_s, double acc_rw, double gyr_rw) { return ((SplineProblem*)h)->addBiasCost(freq_hz, dt_s, acc_rw, gyr_rw); }
int orc_add_gravity_norm(void* h) { auto* p = (SplineProblem*)h; p->gravity_norm = true; p->built = false; return 0; }
int orc_relinearize(void* h) { ((SplineProblem*)h)->relinearize(); return 0; }
int orc_relinearize_mask(void* h, int mask) { ((SplineProblem*)h)->relinearize(mask); return 0; }

// counts: [frames, obs, accel, gyro, gp, bias, gravity, tangent columns]
int orc_counts(void* h, int64_t* c) {
  auto* p = (SplineProblem*)h; p->build();
  int64_t obs = 0; for (auto& f : p->frames) obs += (int64_t)f.lm.size();
  c[0] = (int64_t)p->frames.size(); c[1] = obs; c[2] = (int64_t)p->accel.size(); c[3] = (int64_t)p->gyro.size();
  c[4] = (int64_t)p->gp.size(); c[5] = (int64_t)p->bias.size(); c[6] = p->gravity_norm ? 1 : 0; c[7] = p->mp.num_cols; return 0;
}

// Per-kind raw residuals and LOCAL Jacobians (no loss correction), row-major, in the documented
// layouts of include/ctba.h.  kind: 0 reproj (per observation, 2 x (6N+10)), 1 accel (3 x (6N+16)),
// 2 gyro (3 x (3N+13)), 3 gp (3 x (6N+5)), 4 bias (3 x 24), 5 gravity (1 x 3).
// aux[0] receives the max |entry| of every Jacobian column the reference block owns but the layout
// drops (cross-landmark columns; bias blocks 4..N-1) — must be exactly 0.
int orc_evaluate(void* h, int kind, double* r_out, double* j_out, double* aux) {
  auto* p = (SplineProblem*)h; p->build();
  const int N = p->N;
  double dropped = 0.0;
  size_t ro = 0, jo = 0;
  for (const auto& b : p->mp.blocks) {
    if (b.kind != kind) continue;
    EvalBlockOut o;
    p->mp.evalBlock(b, j_out != nullptr, &o);
    const int nc = o.ncols_local;
    if (kind == 0) {
      const int n = b.nres / 2, W = 6 * N + 10, base = 6 * N + 7;
      for (int i = 0; i < n; ++i)
        for (int rr = 0; rr < 2; ++rr) {
          if (r_out) r_out[ro++] = o.r[2 * i + rr];
          if (j_out) {
            const double* row = &o.jl[(size_t)(2 * i + rr) * nc];
            for (int c = 0; c < base; ++c) j_out[jo++] = row[c];
            for (int c = 0; c < 3; ++c) j_out[jo++] = row[base + 3 * i + c];
            for (int c = base; c < nc; ++c) if (c < base + 3 * i || c >= base + 3 * i + 3) dropped = std::max(dropped, std::fabs(row[c]));
            (void)W;
          }
        }
    } else if (kind == 1 || kind == 2) {
      // accel block: [3N rot | 3N pos | 3N bias | 3 g | 1 t]; gyro: [3N rot | 3N bias | 1 t]
      const int pre = kind == 1 ? 6 * N : 3 * N;
      for (int rr = 0; rr < 3; ++rr) {
        if (r_out) r_out[ro++] = o.r[rr];
        if (j_out) {
          const double* row = &o.jl[(size_t)rr * nc];
          for (int c = 0; c < pre + 12; ++c) j_out[jo++] = row[c];
          for (int c = pre + 12; c < pre + 3 * N; ++c) dropped = std::max(dropped, std::fabs(row[c]));
          for (int c = pre + 3 * N; c < nc; ++c) j_out[jo++] = row[c];
        }
      }
    } else {
      for (int rr = 0; rr < b.nres; ++rr) {
        if (r_out) r_out[ro++] = o.r[rr];
        if (j_out) for (int c = 0; c < nc; ++c) j_out[jo++] = o.jl[(size_t)rr * nc + c];
      }
    }
  }
  if (aux) aux[0] = dropped;
  return (int)ro;
}

// cost[k] = 1/2 sum over blocks of kind k of rho(||r||^2); cost[6] = total
int orc_cost(void* h, double* cost) {
  auto* p = (SplineProblem*)h; p->build();
  for (int k = 0; k < 8; ++k) cost[k] = 0;     // cost[7] = pose priors (block kind 6)
  for (const auto& b : p->mp.blocks) {
    EvalBlockOut o; p->mp.evalBlock(b, false, &o);
    double c = 0.5 * o.sq_norm;
    if (b.cauchy_a > 0) { double rho[3]; cauchy(b.cauchy_a, o.sq_norm, rho); c = 0.5 * rho[0]; }
    cost[b.kind == 6 ? 7 : b.kind] += c; cost[6] += c;
  }
  return 0;
}

// summary: [termination_type, initial_cost, final_cost, n_successful, n_unsuccessful, n_iterations_logged]
// iter_log: n x 7 (cost, cost_change, gradient_max_norm, step_norm, tr_ratio, tr_radius, successful)
int orc_solve(void* h, int max_iters, int num_threads, double* summary, double* iter_log, int iter_cap) {
  auto* p = (SplineProblem*)h; p->build();
  SolveSummary s = solveLM(p->mp, max_iters, num_threads);
  summary[0] = s.termination_type; summary[1] = s.initial_cost; summary[2] = s.final_cost;
  summary[3] = s.num_successful_steps; summary[4] = s.num_unsuccessful_steps; summary[5] = (double)s.iterations.size();
  for (size_t i = 0; i < s.iterations.size() && (int)i < iter_cap; ++i) {
    const auto& it = s.iterations[i];
    double* o = iter_log + 7 * i;
    o[0] = it.cost; o[1] = it.cost_change; o[2] = it.gradient_max_norm; o[3] = it.step_norm; o[4] = it.tr_ratio; o[5] = it.tr_radius; o[6] = it.successful;
  }
  return 0;
}

// Sparse LM (sparse_lm.h): the whole problem on `num_threads` host threads.  summary / iter_log as orc_solve; timings (12):
// [eval_jac_s, eval_res_s, build_s, factor_s, total_s, jac_evals, res_evals, solves, n_band, half_bandwidth, n_globals, n_landmarks]
int orc_solve_sparse(void* h, int max_iters, int num_threads, int no_termination, double* summary, double* iter_log, int iter_cap, double* timings) {
  auto* p = (SplineProblem*)h; p->build();
  SparseTimings tm;
  SolveSummary s = solveLMSparse(*p, max_iters, num_threads, &tm, no_termination != 0);
  summary[0] = s.termination_type; summary[1] = s.initial_cost; summary[2] = s.final_cost;
  summary[3] = s.num_successful_steps; summary[4] = s.num_unsuccessful_steps; summary[5] = (double)s.iterations.size();
  for (size_t i = 0; i < s.iterations.size() && (int)i < iter_cap; ++i) {
    const auto& it = s.iterations[i];
    double* o = iter_log + 7 * i;
    o[0] = it.cost; o[1] = it.cost_change; o[2] = it.gradient_max_norm; o[3] = it.step_norm; o[4] = it.tr_ratio; o[5] = it.tr_radius; o[6] = it.successful;
  }
  if (timings) {
    const double t[12] = {tm.eval_jac_s, tm.eval_res_s, tm.build_s, tm.factor_s, tm.total_s, (double)tm.jac_evals, (double)tm.res_evals, (double)tm.solves,
                          (double)tm.n_band, (double)tm.half_bandwidth, (double)tm.n_globals, (double)tm.n_landmarks};
    std::memcpy(timings, t, sizeof(t));
  }
  return 0;
}

// Wall time (seconds) of `reps` residual+Jacobian evaluations of every block with the Ceres
// DynamicAutoDiff schedule (stride-4 Jets, one block per frame) on `num_threads` threads.
// frame_limit > 0 restricts the run to the first `frame_limit` frame blocks (bounded sample);
// obs_done receives the reprojection observations evaluated per repetition.
double orc_time_evaluate(void* h, int num_threads, int reps, int frame_limit, int64_t* obs_done) {
  auto* p = (SplineProblem*)h; p->build();
  std::vector<const ResidualBlock*> sel;
  int nf = 0; int64_t obs = 0;
  for (const auto& b : p->mp.blocks) {
    if (b.kind == 0) { if (frame_limit > 0 && nf >= frame_limit) continue; ++nf; obs += b.nres / 2; }
    sel.push_back(&b);
  }
  if (obs_done) *obs_done = obs;
  auto t0 = std::chrono::steady_clock::now();
  for (int rep = 0; rep < reps; ++rep) {
    auto work = [&](int t) {
      EvalBlockOut o;
      for (size_t i = t; i < sel.size(); i += num_threads) p->mp.evalBlock(*sel[i], true, &o);
    };
    std::vector<std::thread> th;
    for (int t = 0; t < num_threads; ++t) th.emplace_back(work, t);
    for (auto& t : th) t.join();
  }
  return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// ---- primitives exposed for known-answer tests (Sophus fixtures, Appendix E constants) ----
void orc_so3_exp(const double* w, double* q) { auto r = SO3<double>::exp(V3<double>{w[0], w[1], w[2]}); q[0] = r.x; q[1] = r.y; q[2] = r.z; q[3] = r.w; }
void orc_so3_log(const double* q, double* w) { auto r = SO3<double>{q[0], q[1], q[2], q[3]}.log(); w[0] = r.x; w[1] = r.y; w[2] = r.z; }
void orc_so3_mul(const double* a, const double* b, double* c) { auto r = SO3<double>{a[0], a[1], a[2], a[3]} * SO3<double>{b[0], b[1], b[2], b[3]}; c[0] = r.x; c[1] = r.y; c[2] = r.z; c[3] = r.w; }
void orc_so3_matrix(const double* q, double* R) { auto m = SO3<double>{q[0], q[1], q[2], q[3]}.matrix(); for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R[3 * i + j] = m.m[i][j]; }
void orc_so3_act(const double* q, const double* p, double* o) { auto r = SO3<double>{q[0], q[1], q[2], q[3]}.act(V3<double>{p[0], p[1], p[2]}); o[0] = r.x; o[1] = r.y; o[2] = r.z; }
void orc_so3_plus_jacobian(const double* q, double* J) { double PJ[4][3]; SO3<double>{q[0], q[1], q[2], q[3]}.Dx_this_mul_exp_x_at_0(PJ); for (int i = 0; i < 4; ++i) for (int j = 0; j < 3; ++j) J[3 * i + j] = PJ[i][j]; }
// tables: which 0 blend, 1 cumulative blend, 2 base coefficients; out N*N row-major
int orc_spline_table(int N, int which, double* out) {
  auto cp = [&](auto& tb) { const double(*src)[sizeof(tb.blend[0]) / sizeof(double)] = which == 0 ? tb.blend : which == 1 ? tb.cum_blend : tb.base;
    const int n = (int)(sizeof(tb.blend[0]) / sizeof(double)); for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) out[i * n + j] = src[i][j]; };
  if (N == 4) cp(SplineTables<4>::get()); else if (N == 5) cp(SplineTables<5>::get()); else if (N == 6) cp(SplineTables<6>::get()); else return -1;
  return 0;
}
// spline sample: value R (quat), p, body omega, body omega_dot, world velocity, world accel at (s,u) — fixed-u path
int orc_sample(void* h, int64_t s, double u, double* out19) {
  auto* p = (SplineProblem*)h;
  std::vector<const double*> kn;
  for (int j = 0; j < p->N; ++j) kn.push_back(&p->so3[4 * (s + j)]);
  for (int j = 0; j < p->N; ++j) kn.push_back(&p->r3[3 * (s + j)]);
  SO3<double> R; V3<double> w, wd, jk, pos, vel, acc;
  auto run = [&](auto tag) {
    constexpr int NN = decltype(tag)::value;
    SplineHelper<NN>::template evaluate_lie<double, double>(kn.data(), u, p->inv_dt, &R, &w, &wd, &jk);
    SplineHelper<NN>::template evaluate<double, 0, double>(kn.data() + NN, u, p->inv_dt, &pos);
    SplineHelper<NN>::template evaluate<double, 1, double>(kn.data() + NN, u, p->inv_dt, &vel);
    SplineHelper<NN>::template evaluate<double, 2, double>(kn.data() + NN, u, p->inv_dt, &acc);
  };
  if (p->N == 4) run(std::integral_constant<int, 4>()); else if (p->N == 5) run(std::integral_constant<int, 5>()); else run(std::integral_constant<int, 6>());
  double o[19] = {R.x, R.y, R.z, R.w, pos.x, pos.y, pos.z, w.x, w.y, w.z, wd.x, wd.y, wd.z, vel.x, vel.y, vel.z, acc.x, acc.y, acc.z};
  std::memcpy(out19, o, sizeof(o));
  return 0;
}

// getPose / getAccel / getGyro / get*Bias at timestamps (sample.h)
int orc_sample_at(void* h, int what, int64_t n, const int64_t* t_ns, double* out) { return orc_getters::sample(*(SplineProblem*)h, what, n, t_ns, out); }

// computeValidReprojections (select.h): frames as CSR over the 2-D points that have a 3-D point
int orc_compute_valid_reprojections(void* h, int F, const int64_t* t_ns, const int32_t* pt_begin, const double* uv, const int32_t* lm_index,
                                    int width, int height, int cell_size_px, uint8_t* valid, int32_t* n_valid) {
  auto* p = (SplineProblem*)h;
  std::vector<orc_select::Frame> frames(F);
  for (int f = 0; f < F; ++f) {
    frames[f].time_ns = t_ns[f];
    frames[f].setGrid(width, height, cell_size_px);
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) frames[f].addPoint2D(uv[2 * q], uv[2 * q + 1], (uint64_t)lm_index[q]);
  }
  orc_select::computeValidReprojections(*p, frames);
  for (int f = 0; f < F; ++f) {
    n_valid[f] = frames[f].n_valid_reprojections;
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) valid[q] = frames[f].points2D[q - pt_begin[f]].is_reproj_valid ? 1 : 0;
  }
  return 0;
}


// ======================= discrete-time estimator (dt_problem.h) =======================
void* orc_dt_create() { return new DtProblem(); }
void orc_dt_destroy(void* h) { delete (DtProblem*)h; }
int orc_dt_set_camera(void* h, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc, double t_off, double w, double hh) {
  auto* p = (DtProblem*)h;
  p->camera.cam.model = model; p->camera.cam.fx = intr[0]; p->camera.cam.fy = intr[1]; p->camera.cam.cx = intr[2]; p->camera.cam.cy = intr[3];
  for (int i = 0; i < 4; ++i) { p->camera.cam.d[i] = dist[i]; p->camera.cam.q_bc[i] = q_bc[i]; }
  for (int i = 0; i < 3; ++i) p->camera.cam.p_bc[i] = p_bc[i];
  p->camera.t_offset_cam_imu = t_off; p->camera.w = w; p->camera.h = hh; p->built = false;
  return 0;
}
int orc_dt_set_imu_params(void* h, double gnd, double and_, double grw, double arw, double g_max, double a_max, const double* gW) {
  auto* p = (DtProblem*)h;
  p->imu.gyro_noise_density = gnd; p->imu.acc_noise_density = and_; p->imu.gyro_random_walk = grw; p->imu.acc_random_walk = arw; p->imu.g_max = g_max; p->imu.a_max = a_max;
  for (int i = 0; i < 3; ++i) p->imu.g_W[i] = gW[i];
  return 0;
}
int orc_dt_set_states(void* h, int F, const double* poses, const double* ts_cam) { ((DtProblem*)h)->initializeStates(F, poses, ts_cam); return 0; }
int orc_dt_init_velocity(void* h) { ((DtProblem*)h)->initializeLinearVelocity(); return 0; }
int orc_dt_set_speed_bias(void* h, const double* sb) { auto* p = (DtProblem*)h; std::memcpy(p->sb.data(), sb, p->sb.size() * sizeof(double)); return 0; }
int orc_dt_overwrite_states(void* h, const double* poses, const double* sb) { auto* p = (DtProblem*)h; std::memcpy(p->pose.data(), poses, p->pose.size() * 8); std::memcpy(p->sb.data(), sb, p->sb.size() * 8); return 0; }
int orc_dt_set_pbp(void* h, const double* v) { auto* p = (DtProblem*)h; for (int i = 0; i < 3; ++i) p->p_bp[i] = v[i]; return 0; }
int orc_dt_get_states(void* h, double* poses, double* sb) { auto* p = (DtProblem*)h; std::memcpy(poses, p->pose.data(), p->pose.size() * 8); std::memcpy(sb, p->sb.data(), p->sb.size() * 8); return 0; }
int orc_dt_set_landmarks(void* h, int L, const double* xyz) { auto* p = (DtProblem*)h; p->lm.assign(xyz, xyz + 3 * (size_t)L); p->built = false; return 0; }
int orc_dt_get_landmarks(void* h, double* xyz) { auto* p = (DtProblem*)h; std::memcpy(xyz, p->lm.data(), p->lm.size() * 8); return 0; }
int orc_dt_set_globals(void* h, double scale, const double* p_bp, double t_gs_imu) { auto* p = (DtProblem*)h; p->scale = scale; for (int i = 0; i < 3; ++i) p->p_bp[i] = p_bp[i]; p->t_offset_gs_imu = t_gs_imu; p->built = false; return 0; }
// corr9: scale_corr, R_bc_corr (x y z w), p_bc_corr (3), t_offset_cam_imu_corr ; t_gs_corr separately
int orc_dt_set_corrections(void* h, const double* c9, double t_gs_corr) {
  auto* p = (DtProblem*)h;
  p->scale_corr = c9[0]; for (int i = 0; i < 4; ++i) p->R_bc_corr[i] = c9[1 + i]; for (int i = 0; i < 3; ++i) p->p_bc_corr[i] = c9[5 + i]; p->t_offset_cam_imu_corr = c9[8]; p->t_offset_gs_imu_corr = t_gs_corr;
  return 0;
}
// out13: scale_corr, R_bc_corr(4), p_bc_corr(3), t_offset_cam_imu_corr, t_offset_gs_imu_corr, p_bp(3)
int orc_dt_get_globals(void* h, double* o) {
  auto* p = (DtProblem*)h;
  o[0] = p->scale_corr; for (int i = 0; i < 4; ++i) o[1 + i] = p->R_bc_corr[i]; for (int i = 0; i < 3; ++i) o[5 + i] = p->p_bc_corr[i]; o[8] = p->t_offset_cam_imu_corr; o[9] = p->t_offset_gs_imu_corr;
  for (int i = 0; i < 3; ++i) o[10 + i] = p->p_bp[i];
  return 0;
}
int orc_dt_set_flags(void* h, int scale, int pbp, int camimu, int tgs) { auto* p = (DtProblem*)h; p->optimize_scale = scale; p->optimize_p_bp = pbp; p->optimize_camimu_extrinsics = camimu; p->optimize_t_offset_gs_imu = tgs; p->built = false; return 0; }
// valid reprojections only: CSR over frames; vel = Point2D::velocity (estimator.h:293-325)
int orc_dt_add_frames(void* h, int nF, const int32_t* state_id, const int32_t* obs_begin, const double* uv, const double* vel, const int32_t* lm, double sigma) {
  auto* p = (DtProblem*)h;
  for (int f = 0; f < nF; ++f) {
    DtFrame fr; fr.state_id = state_id[f]; fr.inv_std = 1.0 / sigma;
    for (int q = obs_begin[f]; q < obs_begin[f + 1]; ++q) { fr.uv.push_back(uv[2 * q]); fr.uv.push_back(uv[2 * q + 1]); fr.vel.push_back(vel[2 * q]); fr.vel.push_back(vel[2 * q + 1]); fr.lm.push_back(lm[q]); }
    p->frames.push_back(std::move(fr));
  }
  p->built = false;
  return nF;
}
int orc_dt_add_imu_factor(void* h, int frame0, int n, const double* t, const double* gyr, const double* acc) {
  auto* p = (DtProblem*)h;
  auto f = std::make_shared<DtImuFactor>();
  f->par = p->imu; f->t0 = p->ts_imu[frame0]; f->t1 = p->ts_imu[frame0 + 1];
  for (int i = 0; i < n; ++i) { DtImuSample s; s.t = t[i]; for (int c = 0; c < 3; ++c) { s.gyr[c] = gyr[3 * i + c]; s.acc[c] = acc[3 * i + c]; } f->meas.push_back(s); }
  p->imu_factors.push_back(f); p->imu_frame0.push_back(frame0); p->built = false;
  return 0;
}
int orc_dt_add_gp(void* h, int n, const double* t, const double* pos, const int32_t* frame0, double sigma) {
  auto* p = (DtProblem*)h;
  for (int i = 0; i < n; ++i) p->gps.push_back({t[i], {pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]}, frame0[i], 1.0 / sigma});
  p->built = false;
  return n;
}
int orc_dt_counts(void* h, int64_t* c) {
  auto* p = (DtProblem*)h; p->build();
  int64_t obs = 0, nf = 0; for (auto& f : p->frames) { obs += (int64_t)f.lm.size(); nf += !f.lm.empty(); }
  c[0] = nf; c[1] = obs; c[2] = (int64_t)p->imu_factors.size(); c[3] = (int64_t)p->gps.size(); c[4] = p->mp.num_cols;
  return 0;
}
// raw residuals + LOCAL Jacobians, row-major.  kind 0: per observation 2 x 17 [pose 6 | scale 1 | R_bc 3 | p_bc 3 | t_cam_imu 1 | landmark 3];
// kind 1: per IMU factor 15 x 30 [pose k 6 | sb k 9 | pose k+1 6 | sb k+1 9]; kind 2: per GP 3 x 17 [pose0 6 | pose1 6 | scale 1 | t 1 | p_bp 3].
// EVERY call evaluates the IMU factors (their pre-integration cache may be redone: imu_cost_functor.h:693-704).
int orc_dt_evaluate(void* h, int kind, double* r_out, double* j_out) {
  auto* p = (DtProblem*)h; p->build();
  size_t ro = 0, jo = 0;
  for (const auto& b : p->mp.blocks) {
    if (b.kind != kind) continue;
    EvalBlockOut o;
    p->mp.evalBlock(b, j_out != nullptr, &o);
    const int nc = o.ncols_local;
    if (kind == 0) {
      const int n = b.nres / 2, base = 14;
      for (int i = 0; i < n; ++i)
        for (int rr = 0; rr < 2; ++rr) {
          if (r_out) r_out[ro++] = o.r[2 * i + rr];
          if (j_out) {
            const double* row = &o.jl[(size_t)(2 * i + rr) * nc];
            for (int c = 0; c < base; ++c) j_out[jo++] = row[c];
            for (int c = 0; c < 3; ++c) j_out[jo++] = row[base + 3 * i + c];
          }
        }
    } else {
      for (int rr = 0; rr < b.nres; ++rr) {
        if (r_out) r_out[ro++] = o.r[rr];
        if (j_out) for (int c = 0; c < nc; ++c) j_out[jo++] = o.jl[(size_t)rr * nc + c];
      }
    }
  }
  return (int)ro;
}
int orc_dt_cost(void* h, double* cost4) {
  auto* p = (DtProblem*)h; p->build();
  for (int k = 0; k < 4; ++k) cost4[k] = 0;
  for (const auto& b : p->mp.blocks) {
    EvalBlockOut o; p->mp.evalBlock(b, false, &o);
    double c = 0.5 * o.sq_norm;
    if (b.cauchy_a > 0) { double rho[3]; cauchy(b.cauchy_a, o.sq_norm, rho); c = 0.5 * rho[0]; }
    cost4[b.kind] += c; cost4[3] += c;
  }
  return 0;
}
int orc_dt_solve(void* h, int max_iters, int num_threads, double* summary, double* iter_log, int iter_cap) {
  auto* p = (DtProblem*)h; p->build();
  SolveSummary s = solveLM(p->mp, max_iters, num_threads);
  summary[0] = s.termination_type; summary[1] = s.initial_cost; summary[2] = s.final_cost;
  summary[3] = s.num_successful_steps; summary[4] = s.num_unsuccessful_steps; summary[5] = (double)s.iterations.size();
  for (size_t i = 0; i < s.iterations.size() && (int)i < iter_cap; ++i) {
    const auto& it = s.iterations[i];
    double* o = iter_log + 7 * i;
    o[0] = it.cost; o[1] = it.cost_change; o[2] = it.gradient_max_norm; o[3] = it.step_norm; o[4] = it.tr_ratio; o[5] = it.tr_radius; o[6] = it.successful;
  }
  return 0;
}
// DiscreteTimeEstimator::computeValidReprojections + computeFeatureVelocity (select.h), frames as CSR; vel (2 per point) may be NULL
int orc_dt_compute_valid_reprojections(void* h, int F, const double* ts, const int32_t* pt_begin, const double* uv, const int32_t* lm_index, int width, int height,
                                       int cell_size_px, uint8_t* valid, int32_t* n_valid, double* vel) {
  auto* p = (DtProblem*)h;
  std::vector<orc_select::Frame> frames(F);
  for (int f = 0; f < F; ++f) {
    frames[f].setGrid(width, height, cell_size_px);
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) frames[f].addPoint2D(uv[2 * q], uv[2 * q + 1], (uint64_t)lm_index[q]);
  }
  orc_select::computeValidReprojectionsDt(*p, frames);
  if (vel) orc_select::computeFeatureVelocity(frames, std::vector<double>(ts, ts + F));
  for (int f = 0; f < F; ++f) {
    n_valid[f] = frames[f].n_valid_reprojections;
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) {
      const auto& pt = frames[f].points2D[q - pt_begin[f]];
      valid[q] = pt.is_reproj_valid ? 1 : 0;
      if (vel) { vel[2 * q] = pt.vel_u; vel[2 * q + 1] = pt.vel_v; }
    }
  }
  return 0;
}
// Sparse LM on the discrete-time problem (sparse_lm.h): band = 15 scalars per state [pose tangent 6 | speed-and-bias 9] in time order.
int orc_dt_solve_sparse(void* h, int max_iters, int num_threads, int no_termination, double* summary, double* iter_log, int iter_cap, double* timings) {
  auto* p = (DtProblem*)h; p->build();
  // band indices in time order over the blocks that are in the problem (a state without IMU factor has no speed-and-bias block)
  const size_t F = p->numStates();
  auto in_pose = [p](const double* ptr) { return !p->pose.empty() && ptr >= p->pose.data() && ptr < p->pose.data() + p->pose.size(); };
  auto in_sb = [p](const double* ptr) { return !p->sb.empty() && ptr >= p->sb.data() && ptr < p->sb.data() + p->sb.size(); };
  std::vector<char> has_pose(F, 0), has_sb(F, 0);
  for (const auto& q : p->mp.params) {
    if (q.col < 0) continue;
    if (in_pose(q.ptr)) has_pose[(q.ptr - p->pose.data()) / 7] = 1;
    else if (in_sb(q.ptr)) has_sb[(q.ptr - p->sb.data()) / 9] = 1;
  }
  std::vector<int> pose_base(F, -1), sb_base(F, -1);
  int nb = 0;
  for (size_t i = 0; i < F; ++i) { if (has_pose[i]) { pose_base[i] = nb; nb += 6; } if (has_sb[i]) { sb_base[i] = nb; nb += 9; } }
  SparseSpec spec;
  spec.n_band = nb;
  spec.lm = &p->lm;
  spec.band_base = [&](const double* ptr) -> int {
    if (in_pose(ptr)) return pose_base[(ptr - p->pose.data()) / 7];
    if (in_sb(ptr)) return sb_base[(ptr - p->sb.data()) / 9];
    return -1;
  };
  SparseTimings tm;
  SolveSummary s = solveLMSparseSpec(p->mp, spec, max_iters, num_threads, &tm, no_termination != 0);
  summary[0] = s.termination_type; summary[1] = s.initial_cost; summary[2] = s.final_cost;
  summary[3] = s.num_successful_steps; summary[4] = s.num_unsuccessful_steps; summary[5] = (double)s.iterations.size();
  for (size_t i = 0; i < s.iterations.size() && (int)i < iter_cap; ++i) {
    const auto& it = s.iterations[i];
    double* o = iter_log + 7 * i;
    o[0] = it.cost; o[1] = it.cost_change; o[2] = it.gradient_max_norm; o[3] = it.step_norm; o[4] = it.tr_ratio; o[5] = it.tr_radius; o[6] = it.successful;
  }
  if (timings) {
    const double t[12] = {tm.eval_jac_s, tm.eval_res_s, tm.build_s, tm.factor_s, tm.total_s, (double)tm.jac_evals, (double)tm.res_evals, (double)tm.solves,
                          (double)tm.n_band, (double)tm.half_bandwidth, (double)tm.n_globals, (double)tm.n_landmarks};
    std::memcpy(timings, t, sizeof(t));
  }
  return 0;
}
// one IMU factor at the current state: weighted residual (15) and the reference's MINIMAL Jacobians (15 x 30, jacobians_minimal)
int orc_dt_imu_minimal(void* h, int k, double* r15, double* jmin) {
  auto* p = (DtProblem*)h;
  const int f0 = p->imu_frame0[k];
  std::vector<const double*> ptrs = {&p->pose[7 * f0], &p->sb[9 * f0], &p->pose[7 * (f0 + 1)], &p->sb[9 * (f0 + 1)]};
  double jac[15 * 32];
  p->imu_factors[k]->evaluate(ptrs, r15, jac, jmin);
  return 0;
}
int orc_dt_imu_redo_count(void* h, int k) { return ((DtProblem*)h)->imu_factors[k]->redo_counter; }
void orc_se3_plus(const double* T, const double* d, double* out) { se3_plus(T, d, out); }
void orc_se3_plus_jacobian(const double* T, double* J) { double Jm[7][6]; se3_plus_jacobian(T, Jm); for (int i = 0; i < 7; ++i) for (int j = 0; j < 6; ++j) J[6 * i + j] = Jm[i][j]; }

}  // extern "C"
