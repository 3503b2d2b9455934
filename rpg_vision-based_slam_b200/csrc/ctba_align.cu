// ctba_align.cu — alignment of a frozen spline to the world frame on the device.
//
// Replaces what the reference runs for src/align_spline_to_global_frame.cpp:153-226: n residual blocks
// GlobalPositionSensorWithAlignmentCorrectionCostFunctor (global_position_cost_functor.h:82-150) under CauchyLoss(0.10), every knot constant
// (bspline_so3r3.h:1121-1211), then ceres::Solve (:1768-1782).  At most 11 unknowns, so the Jacobian is never materialised:
//   k_align_accumulate  one thread per measurement (grid-stride): residual, closed-form 3x11 Jacobian, loss correction and the packed
//                       upper triangle of J^T J / J^T r / cost in registers; warp-shuffle + shared-memory reduction per CTA, fixed order
//   k_align_reduce      one CTA sums the per-CTA partials in CTA order (deterministic: the same run gives the same bits)
//   k_align_step        one thread: Jacobi scaling, LM diagonal, 11x11 Cholesky, candidate state, model cost change (align_lm_step)
//   k_align_eval        residuals / Jacobians / unit-weight error norms for ctba_evaluate and evaluateWorldColmapAlignmentErrors (:1682-1790)
// The trust-region loop (align_lm_loop, ctba_align.h) runs on the host exactly like solver_solve; per iteration it reads back 1 + 8 doubles.
#include <chrono>
#include <cstring>
#include "ctba_internal.h"
#include "ctba_align.h"

namespace ctba {

// device work buffer (doubles)
enum { AW_X = 0, AW_CAND = 12, AW_SYS0 = 24, AW_SYS1 = 104, AW_JSCALE = 184, AW_SC = 196, AW_TWG = 204, AW_SIZE = 212 };
constexpr int kAlThreads = 128, kAlWarps = kAlThreads / 32;

struct AlignData {
  DBuf<int32_t> s;
  DBuf<double> st_s, k, inv_std, meas;
  DBuf<double> partial;   // grid x AS_SIZE
  DBuf<double> work;      // AW_SIZE
  int grid = 0;
};

void align_free(AlignData* a) {
  if (!a) return;
  a->s.release(); a->st_s.release(); a->k.release(); a->inv_std.release(); a->meas.release(); a->partial.release(); a->work.release();
  delete a;
}

template <int N>
__global__ void __launch_bounds__(kAlThreads) k_align_accumulate(const __grid_constant__ Globals G, const double* __restrict__ T_wg,
                                                                 const double* __restrict__ x, const double* __restrict__ knot_q,
                                                                 const double* __restrict__ knot_p, int n, const int32_t* __restrict__ a_s,
                                                                 const double* __restrict__ st_s, const double* __restrict__ kk,
                                                                 const double* __restrict__ meas, const double* __restrict__ inv_std,
                                                                 double cauchy_b, double* __restrict__ partial) {
  __shared__ double sh[kAlWarps][AS_SIZE];
  double acc[AS_SIZE];
#pragma unroll
  for (int v = 0; v < AS_SIZE; ++v) acc[v] = 0.0;
  double T[7], xs[AL_STATE];
#pragma unroll
  for (int i = 0; i < 7; ++i) T[i] = T_wg[i];
#pragma unroll
  for (int i = 0; i < AL_STATE; ++i) xs[i] = x[i];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const int s = a_s[i];
    double kq[4 * N], kp[3 * N];
#pragma unroll
    for (int e = 0; e < 4 * N; ++e) kq[e] = knot_q[4 * (size_t)s + e];
#pragma unroll
    for (int e = 0; e < 3 * N; ++e) kp[e] = knot_p[3 * (size_t)s + e];
    const double m[3] = {meas[3 * (size_t)i], meas[3 * (size_t)i + 1], meas[3 * (size_t)i + 2]};
    align_sample_accumulate<N>(G, T, xs, kq, kp, st_s[i], kk[i], m, inv_std[i], cauchy_b, true, acc);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int v = 0; v < AS_SIZE; ++v) {
    double val = acc[v];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) val += __shfl_down_sync(0xffffffffu, val, o);
    if (lane == 0) sh[warp][v] = val;
  }
  __syncthreads();
  if (threadIdx.x < AS_SIZE) {
    double sum = 0.0;
#pragma unroll
    for (int w = 0; w < kAlWarps; ++w) sum += sh[w][threadIdx.x];
    partial[(size_t)blockIdx.x * AS_SIZE + threadIdx.x] = sum;
  }
}

__global__ void __launch_bounds__(kAlThreads) k_align_reduce(int nblocks, const double* __restrict__ partial, double* __restrict__ sys) {
  if (threadIdx.x < AS_SIZE) {
    double sum = 0.0;
    for (int b = 0; b < nblocks; ++b) sum += partial[(size_t)b * AS_SIZE + threadIdx.x];
    sys[threadIdx.x] = sum;
  }
}

__global__ void k_align_step(const double* __restrict__ sys, const double* __restrict__ x, unsigned free_mask, double radius, int init_jscale,
                             double* __restrict__ jscale, double* __restrict__ cand, double* __restrict__ sc) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  double xs[AL_STATE], js[AT_SIZE], cd[AL_STATE], out[AK_SIZE];
  for (int i = 0; i < AL_STATE; ++i) xs[i] = x[i];
  for (int i = 0; i < AT_SIZE; ++i) js[i] = jscale[i];
  for (int i = 0; i < AK_SIZE; ++i) out[i] = 0.0;
  align_lm_step(sys, xs, free_mask, radius, init_jscale != 0, js, cd, out);
  if (init_jscale) for (int i = 0; i < AT_SIZE; ++i) jscale[i] = js[i];
  for (int i = 0; i < AL_STATE; ++i) cand[i] = cd[i];
  for (int i = 0; i < AK_SIZE; ++i) sc[i] = out[i];
}

// r (3n, weighted by 1/noise, no loss), J (3n x 11, same weight), err (n: |e| with unit weight); any of them may be null
template <int N>
__global__ void __launch_bounds__(64) k_align_eval(const __grid_constant__ Globals G, const double* __restrict__ T_wg, const double* __restrict__ x,
                                                   const double* __restrict__ knot_q, const double* __restrict__ knot_p, int n,
                                                   const int32_t* __restrict__ a_s, const double* __restrict__ st_s, const double* __restrict__ kk,
                                                   const double* __restrict__ meas, const double* __restrict__ inv_std, double* __restrict__ r,
                                                   double* __restrict__ J, double* __restrict__ err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double T[7], xs[AL_STATE];
  for (int c = 0; c < 7; ++c) T[c] = T_wg[c];
  for (int c = 0; c < AL_STATE; ++c) xs[c] = x[c];
  const int s = a_s[i];
  double kq[4 * N], kp[3 * N];
#pragma unroll
  for (int e = 0; e < 4 * N; ++e) kq[e] = knot_q[4 * (size_t)s + e];
#pragma unroll
  for (int e = 0; e < 3 * N; ++e) kp[e] = knot_p[3 * (size_t)s + e];
  const double m[3] = {meas[3 * (size_t)i], meas[3 * (size_t)i + 1], meas[3 * (size_t)i + 2]};
  double e3[3], Jl[3 * AT_SIZE];
  if (J) align_eval<N, true>(G, T, xs, kq, kp, st_s[i], kk[i], m, e3, Jl);
  else align_eval<N, false>(G, T, xs, kq, kp, st_s[i], kk[i], m, e3, Jl);
  const double w = inv_std[i];
  if (r) for (int c = 0; c < 3; ++c) r[3 * (size_t)i + c] = w * e3[c];
  if (J) for (int c = 0; c < 3 * AT_SIZE; ++c) J[(size_t)i * 3 * AT_SIZE + c] = w * Jl[c];
  if (err) err[i] = sqrt(e3[0] * e3[0] + e3[1] * e3[1] + e3[2] * e3[2]);
}

static unsigned align_free_mask(const ctba_problem* p) {
  unsigned m = (1u << AT_SCALE) | (7u << AT_R) | (7u << AT_P);          // scale_corr, R_w_g_corr, p_w_g_corr are always free (:1183-1209)
  if (p->optimize_t_offset_s_gs) m |= 1u << AT_T;
  if (p->optimize_p_bp) m |= 7u << AT_PBP;
  return m;
}

// measurements, T_w_g and the packed state -> device.  Knots come from the handle's device mirror (sync_to_device).
static int align_prepare(ctba_problem* p) {
  if (p->al_s.empty()) { p->err = "alignment: no measurements (ctba_add_alignment_measurements)"; return CTBA_E_STATE; }
  if (!p->f_s.empty() || !p->i_s.empty() || !p->g_s.empty() || !p->q_s.empty() || !p->b_s0.empty() || p->gravity_norm) {
    p->err = "alignment measurements cannot be mixed with other terms: the alignment problem freezes every knot (bspline_so3r3.h:1191-1199)";
    return CTBA_E_STATE;
  }
  if (p->comm && p->comm->world > 1) { p->err = "alignment problems run on one GPU"; return CTBA_E_STATE; }
  int rc = sync_to_device(p); if (rc) return rc;
  rc = download_state(p); if (rc) return rc;
  if (!p->align) p->align = new AlignData();
  AlignData* a = p->align;
  cudaStream_t st = p->stream;
  if (p->align_dirty) {
    CTBA_CUDA(p, a->s.upload(p->al_s, st)); CTBA_CUDA(p, a->st_s.upload(p->al_st_s, st)); CTBA_CUDA(p, a->k.upload(p->al_k, st));
    CTBA_CUDA(p, a->inv_std.upload(p->al_inv_std, st)); CTBA_CUDA(p, a->meas.upload(p->al_meas, st));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, p->device);
    const int n = (int)p->al_s.size();
    a->grid = std::max(1, std::min((n + kAlThreads - 1) / kAlThreads, 2 * sms));
    CTBA_CUDA(p, a->partial.resize((size_t)a->grid * AS_SIZE));
    CTBA_CUDA(p, a->work.resize(AW_SIZE));
    p->align_dirty = false;
  }
  double h[AW_SIZE];
  for (int i = 0; i < AW_SIZE; ++i) h[i] = 0.0;
  h[AW_X + AL_T] = p->glb[GLB_T_GP_CORR]; h[AW_X + AL_SCALE] = p->glb[GLB_SCALE_CORR];
  for (int c = 0; c < 4; ++c) h[AW_X + AL_R + c] = p->R_wg_corr[c];
  for (int c = 0; c < 3; ++c) { h[AW_X + AL_P + c] = p->p_wg_corr[c]; h[AW_X + AL_PBP + c] = p->glb[GLB_PBP + c]; }
  for (int c = 0; c < AL_STATE; ++c) h[AW_CAND + c] = h[AW_X + c];
  for (int c = 0; c < 7; ++c) h[AW_TWG + c] = p->T_wg[c];
  CTBA_CUDA(p, cudaMemcpyAsync(a->work.p, h, sizeof(h), cudaMemcpyHostToDevice, st));
  CTBA_CUDA(p, cudaStreamSynchronize(st));   // h is a stack temporary
  return CTBA_OK;
}

namespace {
struct DeviceAlignBackend {
  ctba_problem* p; AlignData* a; Globals G; unsigned mask; double cauchy_b; int cur = 0;
  double t_eval = 0, t_solve = 0;
  double* sys(int which) const { return a->work.p + (which ? AW_SYS1 : AW_SYS0); }
  int evaluate(bool candidate, double* cost) {
    auto t0 = std::chrono::steady_clock::now();
    cudaStream_t st = p->stream;
    const int n = (int)p->al_s.size();
    const double* x = a->work.p + (candidate ? AW_CAND : AW_X);
    double* S = sys(candidate ? 1 - cur : cur);
    const double* T = a->work.p + AW_TWG;
    if (p->N == 4)
      k_align_accumulate<4><<<a->grid, kAlThreads, 0, st>>>(G, T, x, p->dstate.knot_q.p, p->dstate.knot_p.p, n, a->s.p, a->st_s.p, a->k.p, a->meas.p,
                                                            a->inv_std.p, cauchy_b, a->partial.p);
    else
      k_align_accumulate<6><<<a->grid, kAlThreads, 0, st>>>(G, T, x, p->dstate.knot_q.p, p->dstate.knot_p.p, n, a->s.p, a->st_s.p, a->k.p, a->meas.p,
                                                            a->inv_std.p, cauchy_b, a->partial.p);
    k_align_reduce<<<1, kAlThreads, 0, st>>>(a->grid, a->partial.p, S);
    p->launches += 2;
    CTBA_CUDA(p, cudaGetLastError());
    CTBA_CUDA(p, cudaMemcpyAsync(cost, S + AS_COST, sizeof(double), cudaMemcpyDeviceToHost, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));
    t_eval += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return CTBA_OK;
  }
  int step(double radius, bool init_jscale, double* sc) {
    auto t0 = std::chrono::steady_clock::now();
    cudaStream_t st = p->stream;
    k_align_step<<<1, 32, 0, st>>>(sys(cur), a->work.p + AW_X, mask, radius, init_jscale ? 1 : 0, a->work.p + AW_JSCALE, a->work.p + AW_CAND,
                                   a->work.p + AW_SC);
    p->launches += 1;
    CTBA_CUDA(p, cudaGetLastError());
    CTBA_CUDA(p, cudaMemcpyAsync(sc, a->work.p + AW_SC, AK_SIZE * sizeof(double), cudaMemcpyDeviceToHost, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));
    t_solve += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return CTBA_OK;
  }
  void accept() {
    cudaMemcpyAsync(a->work.p + AW_X, a->work.p + AW_CAND, AL_STATE * sizeof(double), cudaMemcpyDeviceToDevice, p->stream);
    cur = 1 - cur;
  }
};
}  // namespace

int align_solve(ctba_problem* p, int max_iters, ctba_summary* out) {
  auto t_begin = std::chrono::steady_clock::now();
  const int64_t launches0 = p->launches;
  int rc = align_prepare(p); if (rc) return rc;
  DeviceAlignBackend be;
  be.p = p; be.a = p->align; be.G = make_globals(p); be.mask = align_free_mask(p); be.cauchy_b = p->align_cauchy_a * p->align_cauchy_a;
  rc = align_lm_loop(be, max_iters, p->bench_no_termination, out); if (rc) return rc;
  // optimised blocks back into the caller-visible state (Ceres updates scale_corr_, R_w_g_corr_, p_w_g_corr_, p_bp_, t_s_gp_corr_s_ in place)
  double x[AL_STATE];
  CTBA_CUDA(p, cudaMemcpyAsync(x, p->align->work.p + AW_X, sizeof(x), cudaMemcpyDeviceToHost, p->stream));
  CTBA_CUDA(p, cudaStreamSynchronize(p->stream));
  p->glb[GLB_T_GP_CORR] = x[AL_T]; p->glb[GLB_SCALE_CORR] = x[AL_SCALE];
  for (int c = 0; c < 4; ++c) p->R_wg_corr[c] = x[AL_R + c];
  for (int c = 0; c < 3; ++c) { p->p_wg_corr[c] = x[AL_P + c]; p->glb[GLB_PBP + c] = x[AL_PBP + c]; }
  p->state_dirty = true;   // the device mirror of the globals is refreshed by the next evaluate / sample
  out->time_eval_s = be.t_eval; out->time_solve_s = be.t_solve; out->time_build_s = 0.0;
  out->kernel_launches = p->launches - launches0;
  out->time_total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  return CTBA_OK;
}

int align_evaluate(ctba_problem* p, double* r_out, double* j_out, double* err_out) {
  int rc = align_prepare(p); if (rc) return rc;
  AlignData* a = p->align;
  cudaStream_t st = p->stream;
  const int n = (int)p->al_s.size();
  DBuf<double> d_r, d_J, d_e;
  if (r_out) CTBA_CUDA(p, d_r.resize(3 * (size_t)n));
  if (j_out) CTBA_CUDA(p, d_J.resize(3 * (size_t)n * AT_SIZE));
  if (err_out) CTBA_CUDA(p, d_e.resize((size_t)n));
  const Globals G = make_globals(p);
  const double *T = a->work.p + AW_TWG, *x = a->work.p + AW_X;
  if (p->N == 4)
    k_align_eval<4><<<(n + 63) / 64, 64, 0, st>>>(G, T, x, p->dstate.knot_q.p, p->dstate.knot_p.p, n, a->s.p, a->st_s.p, a->k.p, a->meas.p, a->inv_std.p,
                                                 d_r.p, d_J.p, d_e.p);
  else
    k_align_eval<6><<<(n + 63) / 64, 64, 0, st>>>(G, T, x, p->dstate.knot_q.p, p->dstate.knot_p.p, n, a->s.p, a->st_s.p, a->k.p, a->meas.p, a->inv_std.p,
                                                 d_r.p, d_J.p, d_e.p);
  p->launches += 1;
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess && r_out) e = cudaMemcpyAsync(r_out, d_r.p, 3 * (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess && j_out) e = cudaMemcpyAsync(j_out, d_J.p, 3 * (size_t)n * AT_SIZE * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess && err_out) e = cudaMemcpyAsync(err_out, d_e.p, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st);
  if (e == cudaSuccess) e = cudaStreamSynchronize(st);
  d_r.release(); d_J.release(); d_e.release();
  if (e != cudaSuccess) { p->err = cudaGetErrorString(e); return CTBA_E_CUDA; }
  return n;
}

}  // namespace ctba
