// ctba_eval.cu — residual + analytic-Jacobian kernels (sm_100a, FP64).
//
//   k_frame_pose   one thread per frame      : spline pose + pose->knot chain (FramePose<N>)
//   k_reproj       one thread per observation: r (double2) and Jl = dr/dp_w (6 planar doubles); HBM-bound
//   k_frame_robust one warp per frame        : |r_frame|^2, Cauchy weight sqrt(rho'), 1/2 rho   (CauchyLoss(3.0) per frame block)
//   k_imu / k_gp / k_bias / k_gravity        : dense local Jacobian rows of the small terms
//   k_sum_sq + k_final_cost                  : deterministic two-stage cost reduction
// Replaces Ceres' Evaluator over the reference's DynamicAutoDiffCostFunctions.
#include "ctba_internal.h"

namespace ctba {

template <int N, bool JAC>
__global__ void __launch_bounds__(64) k_frame_pose(const __grid_constant__ Globals G, StatePtrs S, int F, const int32_t* __restrict__ f_s,
                                                   const double* __restrict__ f_u, FramePose<N>* __restrict__ out) {
  int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const int s = f_s[f];
  double kq[4 * N], kp[3 * N];
#pragma unroll
  for (int i = 0; i < 4 * N; ++i) kq[i] = S.knot_q[4 * (size_t)s + i];
#pragma unroll
  for (int i = 0; i < 3 * N; ++i) kp[i] = S.knot_p[3 * (size_t)s + i];
  FramePose<N> fp;
  frame_pose_eval<N, JAC>(G, S.glb, kq, kp, f_u[f], &fp);
  out[f] = fp;
}

template <int N, bool JAC>
__global__ void __launch_bounds__(256) k_reproj(const __grid_constant__ Globals G, const FramePose<N>* __restrict__ fp,
                                                const double* __restrict__ lm, int M, const int32_t* __restrict__ o_frame,
                                                const int32_t* __restrict__ o_lm, const double2* __restrict__ o_uv,
                                                const double* __restrict__ f_inv_std, double2* __restrict__ r_out, double* __restrict__ Jl_out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  const int f = o_frame[i];
  const int l = o_lm[i];
  const double2 uv = o_uv[i];
  const FramePose<N>& P = fp[f];
  double Rcw[9], tcw[3];
#pragma unroll
  for (int e = 0; e < 9; ++e) Rcw[e] = P.R_cw[e];
#pragma unroll
  for (int e = 0; e < 3; ++e) tcw[e] = P.t_cw[e];
  v3 pw{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
  double r[2], Jl[6];
  reproj_obs_eval<JAC>(G, Rcw, tcw, pw, uv.x, uv.y, f_inv_std[f], r, Jl);
  r_out[i] = make_double2(r[0], r[1]);
  if (JAC) {
#pragma unroll
    for (int c = 0; c < 6; ++c) Jl_out[(size_t)c * M + i] = Jl[c];
  }
}

// CauchyLoss(a) [Ceres]: rho = b log(1+s/b), rho' = 1/(1+s/b);  corrector scales r and J by sqrt(rho') (rho''<0)
__global__ void k_frame_robust(int F, const int32_t* __restrict__ obs_begin, const double2* __restrict__ r, double cauchy_a,
                               double* __restrict__ f_sq, double* __restrict__ f_w, double* __restrict__ f_cost) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= F) return;
  const int b = obs_begin[warp], e = obs_begin[warp + 1];
  double s = 0.0;
  for (int i = b + lane; i < e; i += 32) { double2 v = r[i]; s += v.x * v.x + v.y * v.y; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    const double bb = cauchy_a * cauchy_a;
    const double sum = 1.0 + s / bb;
    f_sq[warp] = s;
    f_w[warp] = sqrt(1.0 / sum);
    f_cost[warp] = 0.5 * bb * log(sum);
  }
}

// Sharded problems: a frame in the overlap of two time segments has observations on two ranks, and CauchyLoss acts on the whole
// frame block (bspline_so3r3.h:1032).  k_frame_sq writes this rank's partial |r_frame|^2 at the frame's GLOBAL index, the array is
// all-reduced, k_frame_weight turns the total into the corrector weight; the cost of a shared frame is counted by the lowest rank.
__global__ void k_frame_sq(int F, const int32_t* __restrict__ obs_begin, const double2* __restrict__ r, int f_first, double* __restrict__ fsq_glob) {
  int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= F) return;
  const int b = obs_begin[warp], e = obs_begin[warp + 1];
  double s = 0.0;
  for (int i = b + lane; i < e; i += 32) { double2 v = r[i]; s += v.x * v.x + v.y * v.y; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) fsq_glob[f_first + warp] = s;
}
__global__ void k_frame_weight(int F, int f_first, int f_cost_begin, const double* __restrict__ fsq_glob, double cauchy_a, double* __restrict__ f_sq,
                               double* __restrict__ f_w, double* __restrict__ f_cost) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f >= F) return;
  const double s = fsq_glob[f_first + f];
  const double bb = cauchy_a * cauchy_a;
  const double sum = 1.0 + s / bb;
  f_sq[f] = s;
  f_w[f] = sqrt(1.0 / sum);
  f_cost[f] = f >= f_cost_begin ? 0.5 * bb * log(sum) : 0.0;
}

template <int N, bool JAC>
__global__ void __launch_bounds__(64) k_imu(const __grid_constant__ Globals G, StatePtrs S, int n, const int32_t* __restrict__ i_s,
                                            const int32_t* __restrict__ i_sb, const double* __restrict__ st_s, const double* __restrict__ kk,
                                            const double* __restrict__ ub, const double* __restrict__ acc, const double* __restrict__ gyr,
                                            const double* __restrict__ inv_a, const double* __restrict__ inv_g, double* __restrict__ r_a,
                                            double* __restrict__ r_g, double* __restrict__ J_a, double* __restrict__ J_g) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = i_s[i], sb = i_sb[i];
  double kq[4 * N], kp[3 * N], ba[12], bg[12];
#pragma unroll
  for (int e = 0; e < 4 * N; ++e) kq[e] = S.knot_q[4 * (size_t)s + e];
#pragma unroll
  for (int e = 0; e < 3 * N; ++e) kp[e] = S.knot_p[3 * (size_t)s + e];
#pragma unroll
  for (int e = 0; e < 12; ++e) { ba[e] = S.bias_a[3 * (size_t)sb + e]; bg[e] = S.bias_g[3 * (size_t)sb + e]; }
  const double am[3] = {acc[3 * (size_t)i], acc[3 * (size_t)i + 1], acc[3 * (size_t)i + 2]};
  const double gm[3] = {gyr[3 * (size_t)i], gyr[3 * (size_t)i + 1], gyr[3 * (size_t)i + 2]};
  double ra[3], rg[3];
  constexpr int WA = 6 * N + 16, WG = 3 * N + 13;
  // planar Jacobians J[entry][sample]: the 32 samples of a warp store 256 contiguous bytes per entry (row-major per sample made every
  // lane write its own 1.2 KB region: 8-byte stores into 32 different sectors per instruction, 110 MB of DRAM writes for 72 MB of data)
  double* Ja = JAC ? J_a + i : nullptr;
  double* Jg = JAC ? J_g + i : nullptr;
  (void)WA; (void)WG;
  imu_sample_eval<N, JAC>(G, S.glb, kq, kp, ba, bg, st_s[i], kk[i], ub[i], am, gm, inv_a[i], inv_g[i], ra, rg, Ja, Jg, (size_t)n);
#pragma unroll
  for (int c = 0; c < 3; ++c) { r_a[3 * (size_t)i + c] = ra[c]; r_g[3 * (size_t)i + c] = rg[c]; }
}

template <int N, bool JAC>
__global__ void __launch_bounds__(64) k_gp(const __grid_constant__ Globals G, StatePtrs S, int n, const int32_t* __restrict__ g_s,
                                           const double* __restrict__ st_s, const double* __restrict__ kk, const double* __restrict__ meas,
                                           const double* __restrict__ inv_std, double* __restrict__ r, double* __restrict__ J) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = g_s[i];
  double kq[4 * N], kp[3 * N];
#pragma unroll
  for (int e = 0; e < 4 * N; ++e) kq[e] = S.knot_q[4 * (size_t)s + e];
#pragma unroll
  for (int e = 0; e < 3 * N; ++e) kp[e] = S.knot_p[3 * (size_t)s + e];
  const double m[3] = {meas[3 * (size_t)i], meas[3 * (size_t)i + 1], meas[3 * (size_t)i + 2]};
  double rr[3];
  constexpr int W = 6 * N + 5;
  gp_sample_eval<N, JAC>(G, S.glb, kq, kp, st_s[i], kk[i], m, inv_std[i], rr, JAC ? J + (size_t)i * 3 * W : nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) r[3 * (size_t)i + c] = rr[c];
}

template <int N, bool JAC>
__global__ void __launch_bounds__(64) k_prior(const __grid_constant__ Globals G, StatePtrs S, int n, const int32_t* __restrict__ q_s,
                                              const int32_t* __restrict__ q_which, const double* __restrict__ q_u,
                                              const double* __restrict__ meas, const double* __restrict__ inv_std, double* __restrict__ r,
                                              double* __restrict__ J) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = q_s[i];
  double kq[4 * N], kp[3 * N];
#pragma unroll
  for (int e = 0; e < 4 * N; ++e) kq[e] = S.knot_q[4 * (size_t)s + e];
#pragma unroll
  for (int e = 0; e < 3 * N; ++e) kp[e] = S.knot_p[3 * (size_t)s + e];
  const double m[4] = {meas[4 * (size_t)i], meas[4 * (size_t)i + 1], meas[4 * (size_t)i + 2], meas[4 * (size_t)i + 3]};
  double rr[3];
  prior_eval<N, JAC>(G, kq, kp, q_which[i], q_u[i], m, inv_std[i], rr, JAC ? J + (size_t)i * 9 * N : nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) r[3 * (size_t)i + c] = rr[c];
}

template <bool JAC>
__global__ void k_bias(const __grid_constant__ Globals G, StatePtrs S, int n, const int32_t* __restrict__ s0, const int32_t* __restrict__ s1,
                       const int32_t* __restrict__ which, const double* __restrict__ u0, const double* __restrict__ u1,
                       const double* __restrict__ inv_std, double* __restrict__ r, double* __restrict__ J) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double* B = which[i] == 0 ? S.bias_a : S.bias_g;
  double b0[12], b1[12];
#pragma unroll
  for (int e = 0; e < 12; ++e) { b0[e] = B[3 * (size_t)s0[i] + e]; b1[e] = B[3 * (size_t)s1[i] + e]; }
  double rr[3];
  bias_cost_eval<JAC>(G, b0, b1, u0[i], u1[i], inv_std[i], rr, JAC ? J + (size_t)i * 72 : nullptr);
#pragma unroll
  for (int c = 0; c < 3; ++c) r[3 * (size_t)i + c] = rr[c];
}

__global__ void k_gravity(StatePtrs S, double* r, double* J) {
  if (threadIdx.x == 0 && blockIdx.x == 0) gravity_eval(S.glb + GLB_G, r, J);
}

// ---- deterministic reductions ---------------------------------------------------------------------
constexpr int kRedBlocks = 256, kRedThreads = 256;
// partial[b] = sum over a grid-stride slice of x[i]^2 (sq=true) or x[i]
__global__ void __launch_bounds__(kRedThreads) k_partial_sum(const double* __restrict__ x, size_t n, bool sq, double* __restrict__ partial) {
  __shared__ double sh[kRedThreads];
  double s = 0.0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    double v = x[i];
    s += sq ? v * v : v;
  }
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = kRedThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) partial[blockIdx.x] = sh[0];
}
// cost[k] = factor * sum(partial[k*kRedBlocks ..]) for k<6 and k = 7 (pose priors) ; cost[6] = total
__global__ void __launch_bounds__(kRedThreads) k_final_cost(const double* __restrict__ partial, double* __restrict__ cost) {
  __shared__ double sh[kRedThreads];
  __shared__ double tot[8];
  for (int k = 0; k < 8; ++k) {
    if (k == 6) continue;
    sh[threadIdx.x] = partial[k * kRedBlocks + threadIdx.x];
    __syncthreads();
    for (int o = kRedThreads / 2; o > 0; o >>= 1) {
      if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
      __syncthreads();
    }
    if (threadIdx.x == 0) tot[k] = (k == 0 ? 1.0 : 0.5) * sh[0];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double t = 0;
    for (int k = 0; k < 6; ++k) { cost[k] = tot[k]; t += tot[k]; }
    cost[7] = tot[7];
    cost[6] = t + tot[7];
  }
}

template <int N>
__global__ void k_expand_reproj(const __grid_constant__ Globals G, StatePtrs S, const FramePose<N>* __restrict__ fp, int M,
                                const int32_t* __restrict__ o_frame, const int32_t* __restrict__ o_lm, const double* __restrict__ Jl,
                                double* __restrict__ out) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  const int l = o_lm[i];
  v3 pw{S.lm[3 * (size_t)l], S.lm[3 * (size_t)l + 1], S.lm[3 * (size_t)l + 2]};
  double jl[6];
  for (int c = 0; c < 6; ++c) jl[c] = Jl[(size_t)c * M + i];
  reproj_expand<N>(G, S.glb, fp[o_frame[i]], pw, jl, out + (size_t)i * 2 * (6 * N + 10));
}

// ---------------------------------------------------------------------------------------------------
Globals make_globals(const ctba_problem* p) {
  Globals G;
  fill_tables(p->N, &G.tb);
  fill_tables(4, &G.tb4);
  G.cam = p->cam;
  for (int i = 0; i < 4; ++i) G.q_bc0[i] = p->q_bc0[i];
  for (int i = 0; i < 3; ++i) G.p_bc0[i] = p->p_bc0[i];
  G.inv_dt = p->inv_dt; G.dt_s = p->dt_s; G.inv_dt_bias = p->inv_dt_bias;
  G.scale = p->scale; G.gate_sq = 100.0;
  return G;
}

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

template <int N>
static int launch_evaluate_T(ctba_problem* p, EvalBuffers& ev, const StatePtrs& S, bool jac) {
  const Globals G = make_globals(p);
  cudaStream_t st = p->stream;
  const int F = (int)p->numFrames(), M = (int)p->numObs();
  const int nI = (int)p->i_s.size(), nG = (int)p->g_s.size(), nB = (int)p->b_s0.size();
  if (F > 0) {
    CTBA_CUDA(p, ev.frame_pose.resize((size_t)F * sizeof(FramePose<N>) / 8));
    CTBA_CUDA(p, ev.r_reproj.resize(2 * (size_t)M));
    CTBA_CUDA(p, ev.f_sq.resize(F)); CTBA_CUDA(p, ev.f_w.resize(F)); CTBA_CUDA(p, ev.f_cost.resize(F));
    if (jac) CTBA_CUDA(p, ev.Jl.resize(6 * (size_t)M));
    FramePose<N>* fp = (FramePose<N>*)ev.frame_pose.p;
    if (jac) {
      k_frame_pose<N, true><<<cdiv(F, 64), 64, 0, st>>>(G, S, F, p->d_f_s.p, p->d_f_u.p, fp);
      k_reproj<N, true><<<cdiv(M, 256), 256, 0, st>>>(G, fp, S.lm, M, p->d_o_frame.p, p->d_o_lm.p, (const double2*)p->d_o_uv.p,
                                                       p->d_f_inv_std.p, (double2*)ev.r_reproj.p, ev.Jl.p);
    } else {
      k_frame_pose<N, false><<<cdiv(F, 64), 64, 0, st>>>(G, S, F, p->d_f_s.p, p->d_f_u.p, fp);
      k_reproj<N, false><<<cdiv(M, 256), 256, 0, st>>>(G, fp, S.lm, M, p->d_o_frame.p, p->d_o_lm.p, (const double2*)p->d_o_uv.p,
                                                        p->d_f_inv_std.p, (double2*)ev.r_reproj.p, nullptr);
    }
    if (!p->sharded) {
      k_frame_robust<<<cdiv((size_t)F * 32, 256), 256, 0, st>>>(F, p->d_f_obs_begin.p, (const double2*)ev.r_reproj.p, p->cauchy_a, ev.f_sq.p,
                                                               ev.f_w.p, ev.f_cost.p);
      p->launches += 3;
    }
  }
  if (p->sharded && p->shard->F_global > 0) {   // collective: every rank takes part, with or without local frames
    const ShardInfo& sh = *p->shard;
    CTBA_CUDA(p, p->d_fsq_glob.resize(sh.F_global));
    CTBA_CUDA(p, cudaMemsetAsync(p->d_fsq_glob.p, 0, sh.F_global * sizeof(double), st));
    if (F > 0) k_frame_sq<<<cdiv((size_t)F * 32, 256), 256, 0, st>>>(F, p->d_f_obs_begin.p, (const double2*)ev.r_reproj.p, sh.f_first, p->d_fsq_glob.p);
    if (p->comm->allreduce_sum(p->d_fsq_glob.p, sh.F_global, st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
    if (F > 0) {
      k_frame_weight<<<cdiv(F, 256), 256, 0, st>>>(F, sh.f_first, sh.f_cost_begin, p->d_fsq_glob.p, p->cauchy_a, ev.f_sq.p, ev.f_w.p, ev.f_cost.p);
      p->launches += 4;
    }
  }
  if (nI > 0) {
    CTBA_CUDA(p, ev.r_acc.resize(3 * (size_t)nI)); CTBA_CUDA(p, ev.r_gyr.resize(3 * (size_t)nI));
    if (jac) { CTBA_CUDA(p, ev.J_acc.resize((size_t)nI * 3 * (6 * N + 16))); CTBA_CUDA(p, ev.J_gyr.resize((size_t)nI * 3 * (3 * N + 13))); }
    if (jac)
      k_imu<N, true><<<cdiv(nI, 64), 64, 0, st>>>(G, S, nI, p->d_i_s.p, p->d_i_sb.p, p->d_i_st_s.p, p->d_i_k.p, p->d_i_ub.p, p->d_i_acc.p,
                                                   p->d_i_gyr.p, p->d_i_inv_a.p, p->d_i_inv_g.p, ev.r_acc.p, ev.r_gyr.p, ev.J_acc.p, ev.J_gyr.p);
    else
      k_imu<N, false><<<cdiv(nI, 64), 64, 0, st>>>(G, S, nI, p->d_i_s.p, p->d_i_sb.p, p->d_i_st_s.p, p->d_i_k.p, p->d_i_ub.p, p->d_i_acc.p,
                                                    p->d_i_gyr.p, p->d_i_inv_a.p, p->d_i_inv_g.p, ev.r_acc.p, ev.r_gyr.p, nullptr, nullptr);
    p->launches += 1;
  }
  if (nG > 0) {
    CTBA_CUDA(p, ev.r_gp.resize(3 * (size_t)nG));
    if (jac) CTBA_CUDA(p, ev.J_gp.resize((size_t)nG * 3 * (6 * N + 5)));
    if (jac)
      k_gp<N, true><<<cdiv(nG, 64), 64, 0, st>>>(G, S, nG, p->d_g_s.p, p->d_g_st_s.p, p->d_g_k.p, p->d_g_meas.p, p->d_g_inv_std.p, ev.r_gp.p, ev.J_gp.p);
    else
      k_gp<N, false><<<cdiv(nG, 64), 64, 0, st>>>(G, S, nG, p->d_g_s.p, p->d_g_st_s.p, p->d_g_k.p, p->d_g_meas.p, p->d_g_inv_std.p, ev.r_gp.p, nullptr);
    p->launches += 1;
  }
  if (nB > 0) {
    CTBA_CUDA(p, ev.r_bias.resize(3 * (size_t)nB));
    if (jac) CTBA_CUDA(p, ev.J_bias.resize((size_t)nB * 72));
    if (jac)
      k_bias<true><<<cdiv(nB, 128), 128, 0, st>>>(G, S, nB, p->d_b_s0.p, p->d_b_s1.p, p->d_b_which.p, p->d_b_u0.p, p->d_b_u1.p, p->d_b_inv_std.p,
                                                   ev.r_bias.p, ev.J_bias.p);
    else
      k_bias<false><<<cdiv(nB, 128), 128, 0, st>>>(G, S, nB, p->d_b_s0.p, p->d_b_s1.p, p->d_b_which.p, p->d_b_u0.p, p->d_b_u1.p, p->d_b_inv_std.p,
                                                    ev.r_bias.p, nullptr);
    p->launches += 1;
  }
  const int nQ = (int)p->q_s.size();
  if (nQ > 0) {
    CTBA_CUDA(p, ev.r_prior.resize(3 * (size_t)nQ));
    if (jac) CTBA_CUDA(p, ev.J_prior.resize((size_t)nQ * 9 * N));
    if (jac)
      k_prior<N, true><<<cdiv(nQ, 64), 64, 0, st>>>(G, S, nQ, p->d_q_s.p, p->d_q_which.p, p->d_q_u.p, p->d_q_meas.p, p->d_q_inv_std.p, ev.r_prior.p, ev.J_prior.p);
    else
      k_prior<N, false><<<cdiv(nQ, 64), 64, 0, st>>>(G, S, nQ, p->d_q_s.p, p->d_q_which.p, p->d_q_u.p, p->d_q_meas.p, p->d_q_inv_std.p, ev.r_prior.p, nullptr);
    p->launches += 1;
  }
  if (p->gravity_norm) {
    CTBA_CUDA(p, ev.r_grav.resize(1)); CTBA_CUDA(p, ev.J_grav.resize(3));
    k_gravity<<<1, 32, 0, st>>>(S, ev.r_grav.p, jac ? ev.J_grav.p : nullptr);
    p->launches += 1;
  }
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

// k_reproj alone (the HBM-bound evaluation kernel), timed launch by launch with CUDA events on the handle's stream; between launches
// a buffer larger than L2 is overwritten so that every launch reads its inputs from HBM.  Returns the average launch time in ms.
int bench_reproj(ctba_problem* p, int reps, int flush_l2, double* ms_avg) {
  const Globals G = make_globals(p);
  cudaStream_t st = p->stream;
  const int F = (int)p->numFrames(), M = (int)p->numObs();
  *ms_avg = 0.0;
  if (F == 0 || M == 0 || reps <= 0) return CTBA_OK;
  int rc = launch_evaluate(p, p->ev, p->dstate.ptrs(), true); if (rc) return rc;   // frame records + output buffers exist
  DBuf<double> scratch;
  const size_t flush_n = (size_t)256 << 17;   // 256 MiB of doubles
  if (flush_l2) CTBA_CUDA(p, scratch.resize(flush_n));
  cudaEvent_t a, b;
  CTBA_CUDA(p, cudaEventCreate(&a)); CTBA_CUDA(p, cudaEventCreate(&b));
  double total = 0.0;
  const StatePtrs S = p->dstate.ptrs();
  for (int i = 0; i < reps; ++i) {
    if (flush_l2) CTBA_CUDA(p, cudaMemsetAsync(scratch.p, i & 1, flush_n * sizeof(double), st));
    CTBA_CUDA(p, cudaEventRecord(a, st));
    if (p->N == 4)
      k_reproj<4, true><<<cdiv(M, 256), 256, 0, st>>>(G, (const FramePose<4>*)p->ev.frame_pose.p, S.lm, M, p->d_o_frame.p, p->d_o_lm.p,
                                                       (const double2*)p->d_o_uv.p, p->d_f_inv_std.p, (double2*)p->ev.r_reproj.p, p->ev.Jl.p);
    else
      k_reproj<6, true><<<cdiv(M, 256), 256, 0, st>>>(G, (const FramePose<6>*)p->ev.frame_pose.p, S.lm, M, p->d_o_frame.p, p->d_o_lm.p,
                                                       (const double2*)p->d_o_uv.p, p->d_f_inv_std.p, (double2*)p->ev.r_reproj.p, p->ev.Jl.p);
    CTBA_CUDA(p, cudaEventRecord(b, st));
    CTBA_CUDA(p, cudaEventSynchronize(b));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, a, b);
    total += ms;
    p->launches++;
  }
  cudaEventDestroy(a); cudaEventDestroy(b);
  scratch.release();
  *ms_avg = total / reps;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

int launch_evaluate(ctba_problem* p, EvalBuffers& ev, const StatePtrs& S, bool jac) {
  return p->N == 4 ? launch_evaluate_T<4>(p, ev, S, jac) : launch_evaluate_T<6>(p, ev, S, jac);
}

int launch_cost(ctba_problem* p, EvalBuffers& ev) {
  cudaStream_t st = p->stream;
  CTBA_CUDA(p, ev.partial.resize(8 * kRedBlocks));
  CTBA_CUDA(p, ev.cost.resize(8));
  CTBA_CUDA(p, cudaMemsetAsync(ev.partial.p, 0, 8 * kRedBlocks * sizeof(double), st));
  const size_t F = p->numFrames(), nI = p->i_s.size(), nG = p->g_s.size(), nB = p->b_s0.size();
  if (F) { k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.f_cost.p, F, false, ev.partial.p + 0 * kRedBlocks); p->launches++; }
  if (nI) {
    k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_acc.p, 3 * nI, true, ev.partial.p + 1 * kRedBlocks);
    k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_gyr.p, 3 * nI, true, ev.partial.p + 2 * kRedBlocks);
    p->launches += 2;
  }
  if (nG) { k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_gp.p, 3 * nG, true, ev.partial.p + 3 * kRedBlocks); p->launches++; }
  if (nB) { k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_bias.p, 3 * nB, true, ev.partial.p + 4 * kRedBlocks); p->launches++; }
  if (p->gravity_norm) { k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_grav.p, 1, true, ev.partial.p + 5 * kRedBlocks); p->launches++; }
  if (!p->q_s.empty()) { k_partial_sum<<<kRedBlocks, kRedThreads, 0, st>>>(ev.r_prior.p, 3 * p->q_s.size(), true, ev.partial.p + 7 * kRedBlocks); p->launches++; }
  k_final_cost<<<1, kRedThreads, 0, st>>>(ev.partial.p, ev.cost.p);
  p->launches++;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

int expand_reproj_dense(ctba_problem* p, const StatePtrs& S, double* d_out) {
  const Globals G = make_globals(p);
  const int M = (int)p->numObs();
  if (M == 0) return CTBA_OK;
  if (p->N == 4)
    k_expand_reproj<4><<<cdiv(M, 128), 128, 0, p->stream>>>(G, S, (const FramePose<4>*)p->ev.frame_pose.p, M, p->d_o_frame.p, p->d_o_lm.p, p->ev.Jl.p, d_out);
  else
    k_expand_reproj<6><<<cdiv(M, 128), 128, 0, p->stream>>>(G, S, (const FramePose<6>*)p->ev.frame_pose.p, M, p->d_o_frame.p, p->d_o_lm.p, p->ev.Jl.p, d_out);
  p->launches++;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

}  // namespace ctba
