// ctba_dt.cu — the discrete-time estimator on the device (include/ctba_dt.h; BASELINE configs[3]).
//
// Replaces DiscreteTimeEstimator (include/discrete_time_estimator/estimator.h:47-824) behind the C ABI: per-frame SE3 + [v, b_g, b_a]
// states, reprojection / pre-integrated IMU / interpolated global-position terms (ctba_dt_math.h), and ceres::Solve as the reference
// configures it (estimator.h:699-713: trust-region LM, SPARSE_NORMAL_CHOLESKY, defaults) restated like ctba_solve.cu.
//
// Unknown ordering: 15 scalars per frame [pose tangent 6 | v 3 | b_g 3 | b_a 3] -> band index 15 f + k ; the <= 12 global tangents
// (scale, R_BC, p_BC, t_cam_imu, t_gs, p_BP) are border rows; landmarks are eliminated by a Schur complement (exact, like the
// reference's sparse Cholesky).  The reduced system lives in the BandLayout of ctba_solver.h and is factorised by band_factor_solve
// (ctba_band.cu) through an embedded solver context.
//
// HBM layout: per valid reprojection one 36-double record [r 2 | J 2 x 17] (288 B, sector aligned); per IMU factor r 15 + J 15 x 30 and a
// persistent DtImuCache (the reference's mutable pre-integration members); per landmark V (6), gradient (3), V^-1 (6), the summed
// landmark-global coupling (24); per reprojection the landmark-pose coupling W (18).
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <numeric>
#include <string>
#include <vector>
#include "../../include/ctba_dt.h"
#include "ctba_dt_math.h"
#include "ctba_solver.h"
#include "ctba_select_core.h"

using namespace ctba;

namespace {

template <class T>
struct Dev : DBuf<T> {
  Dev() = default;
  Dev(const Dev&) = delete;
  Dev& operator=(const Dev&) = delete;
  ~Dev() { this->release(); }
  void swap(Dev& o) { std::swap(this->p, o.p); std::swap(this->n, o.n); }
};

struct DtEval {
  Dev<double> rec;     // 36 M : [r 2 | J 34] per valid reprojection
  Dev<double> f_sq;    // per measurement frame: |r|^2
  Dev<double> r_imu, J_imu, r_gp, J_gp;
  Dev<double> cost;    // 4
  void swap(DtEval& o) { rec.swap(o.rec); f_sq.swap(o.f_sq); r_imu.swap(o.r_imu); J_imu.swap(o.J_imu); r_gp.swap(o.r_gp); J_gp.swap(o.J_gp); cost.swap(o.cost); }
};
struct DtState {
  Dev<double> pose, sb, lm, glb;
  void swap(DtState& o) { pose.swap(o.pose); sb.swap(o.sb); lm.swap(o.lm); glb.swap(o.glb); }
};
struct DtStateP { double *pose, *sb, *lm, *glb; };
static DtStateP ptrs(const DtState& s) { return {s.pose.p, s.sb.p, s.lm.p, s.glb.p}; }

enum { DSC_MODEL = 0, DSC_STEP2 = 1, DSC_XNORM2 = 2, DSC_GMAX = 3, DSC_COUNT = 4 };

}  // namespace

struct ctba_dt_problem {
  int device = 0;
  ctba_problem* ctx = nullptr;   // embedded context of the band solver (stream, SolverData with Hw / delta / flag)
  cudaStream_t stream = nullptr;
  std::string err;
  DtConst K{};
  bool have_camera = false, have_imu = false;
  // ---- host state ----
  std::vector<double> pose, sb, ts_cam, ts_imu, lm;
  double glb[DG_SIZE] = {0, 0, 0, 0, 1, 0, 0, 0, 0, 0, 0, 0, 0};
  bool optimize_p_bp = true, optimize_scale = false, optimize_camimu = true, optimize_t_gs = true;   // estimator.h:59-66
  bool bench_no_termination = false, bench_accept_all = false;
  // ---- host measurements ----
  std::vector<int32_t> f_state, f_obs_begin{0}, o_lm;
  std::vector<double> f_inv_std, o_uv, o_vel;
  std::vector<int32_t> i_frame0, i_begin{0};
  std::vector<double> i_t, i_gyr, i_acc;
  std::vector<int32_t> g_frame0;
  std::vector<double> g_t, g_pos, g_inv_std;
  // ---- device ----
  bool meas_dirty = true, state_dirty = true, host_stale = false, struct_dirty = true;
  DtState st, cand;
  Dev<int32_t> d_f_state, d_f_obs_begin, d_o_lm, d_o_frame, d_i_frame0, d_i_begin, d_g_frame0;
  Dev<double> d_f_inv_std, d_o_uv, d_o_vel, d_ts_imu, d_i_t, d_i_gyr, d_i_acc, d_g_t, d_g_pos, d_g_inv_std;
  Dev<double> d_imu_scratch;     // 675 doubles per factor
  Dev<char> d_imu_cache;         // sizeof(DtImuCache) per factor: the factors' mutable pre-integration state
  Dev<int32_t> d_imu_rc;
  size_t n_cache = 0;
  DtEval ev, ev_alt;
  // ---- structure + solver ----
  int n_band = 0, n_pad = 0, nbw = 0, Lm = 0;
  int gactive[GT_SIZE] = {0};
  Dev<int32_t> d_lm_begin, d_lm_obs;       // landmark -> its reprojections, sorted by state
  Dev<uint8_t> d_sb_active, d_lm_active;
  Dev<double> f_w;                          // per measurement frame sqrt(rho')
  Dev<double> grad, diag, jscale, dvec;     // n_pad + 16 each
  Dev<double> V0, gl, jscale_lm, Vi, Vg, WgS, Wp, Yp, delta_lm, rhs_acc;   // 6L, 3L, 3L, 6L, 3L, 24L, 18M, 18M, 3L, n_pad+16
  Dev<double> cp_frame, cp_lm, cp_gp;       // partials of the global corner / gradient: 44 per frame, 44 per landmark, 20 per GP run
  Dev<int64_t> d_key_begin;                 // (state, distance) key -> its landmark pairs (CSR)
  Dev<int2> d_pairs;                        // (reprojection at the later state, reprojection at the earlier state)
  Dev<int32_t> d_run_begin, d_gp_order;     // GP samples grouped by first state
  int T1 = 1;                               // largest state distance coupled through a landmark, + 1
  Dev<double> scal, scal_part;
  int64_t launches = 0;
  size_t F() const { return ts_cam.size(); }
  size_t M() const { return o_lm.size(); }
  size_t nI() const { return i_frame0.size(); }
  size_t nG() const { return g_t.size(); }
};

namespace {

#define DT_CUDA(p, call)                                                                    \
  do {                                                                                      \
    cudaError_t e__ = (call);                                                               \
    if (e__ != cudaSuccess) { (p)->err = std::string(#call) + ": " + cudaGetErrorString(e__); return CTBA_E_CUDA; } \
  } while (0)
#define DT_CHECK(p) do { if (!(p)) return CTBA_E_ARG; } while (0)

inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

struct GAct { int a[GT_SIZE]; };

// 256-bit accesses (32-byte aligned): one sector per request instead of four 8-byte requests
__device__ __forceinline__ void ld256(const double* p, double* v) {
  asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
}
__device__ __forceinline__ void ld256c(const double* p, double* v) {   // coherent: data written earlier by another kernel of the same graph / stream is fine either way; kept for buffers rewritten every iteration
  asm volatile("ld.global.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p) : "memory");
}
__device__ __forceinline__ void st256(double* p, const double* v) {
  asm volatile("st.global.v4.f64 [%0], {%1,%2,%3,%4};" ::"l"(p), "d"(v[0]), "d"(v[1]), "d"(v[2]), "d"(v[3]) : "memory");
}
constexpr int kWS = 20;   // stride of the per-reprojection W / Y records: 18 doubles padded to 160 B = 5 sectors

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ void atomic_max_nonneg(double* addr, double v) {
  atomicMax(reinterpret_cast<unsigned long long*>(addr), (unsigned long long)__double_as_longlong(v));
}

// ================= evaluation =================
// one warp per measurement frame; lanes over its valid reprojections.  rec = [r 2 | J 34] ; f_sq = sum of squares (fixed order)
template <bool JAC>
__global__ void __launch_bounds__(128) k_dt_reproj(DtConst K, int nF, const int32_t* __restrict__ f_state, const int32_t* __restrict__ f_obs_begin,
                                                   const double* __restrict__ f_inv_std, const double* __restrict__ o_uv, const double* __restrict__ o_vel,
                                                   const int32_t* __restrict__ o_lm, DtStateP S, double* __restrict__ rec, double* __restrict__ f_sq) {
  const int f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (f >= nF) return;
  DtFrameCam fc;
  dt_reproj_frame(K, S.pose + 7 * (size_t)f_state[f], S.glb, &fc);
  const double inv_std = f_inv_std[f];
  double sq = 0.0;
  for (int o = f_obs_begin[f] + lane; o < f_obs_begin[f + 1]; o += 32) {
    const double* pw = S.lm + 3 * (size_t)o_lm[o];
    double R[36];
    dt_reproj_point<JAC>(K, fc, v3{pw[0], pw[1], pw[2]}, o_uv + 2 * (size_t)o, o_vel + 2 * (size_t)o, inv_std, R, R + 2);
    double* out = rec + 36 * (size_t)o;
    if (JAC) {
#pragma unroll
      for (int k = 0; k < 9; ++k) st256(out + 4 * k, R + 4 * k);
    } else {
      out[0] = R[0]; out[1] = R[1];
    }
    sq += R[0] * R[0] + R[1] * R[1];
  }
  sq = warp_sum(sq);
  if (lane == 0) f_sq[f] = sq;
}

// one thread per IMU factor (the pre-integration is a serial recursion over the samples)
__global__ void __launch_bounds__(64) k_dt_imu(DtConst K, int nI, const int32_t* __restrict__ i_frame0, const int32_t* __restrict__ i_begin, const double* __restrict__ i_t,
                                               const double* __restrict__ i_gyr, const double* __restrict__ i_acc, const double* __restrict__ ts_imu, DtStateP S,
                                               DtImuCache* __restrict__ cache, double* __restrict__ scratch, double* __restrict__ r_out, double* __restrict__ J_out,
                                               int32_t* __restrict__ rc_out) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nI) return;
  const int f0 = i_frame0[k], b = i_begin[k], n = i_begin[k + 1] - b;
  const int rc = dt_imu_eval(K, cache + k, i_t + b, i_gyr + 3 * (size_t)b, i_acc + 3 * (size_t)b, n, ts_imu[f0], ts_imu[f0 + 1], S.pose + 7 * (size_t)f0, S.sb + 9 * (size_t)f0,
                             S.pose + 7 * (size_t)(f0 + 1), S.sb + 9 * (size_t)(f0 + 1), r_out + 15 * (size_t)k, J_out ? J_out + 450 * (size_t)k : nullptr,
                             scratch + 675 * (size_t)k);
  if (rc < 0) atomicMin(rc_out, rc);
}

template <bool JAC>
__global__ void k_dt_gp(DtConst K, int nG, const int32_t* __restrict__ g_frame0, const double* __restrict__ g_t, const double* __restrict__ g_pos,
                        const double* __restrict__ g_inv_std, const double* __restrict__ ts_imu, DtStateP S, double* __restrict__ r_out, double* __restrict__ J_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= nG) return;
  const int f0 = g_frame0[i];
  double r[3], J[51];
  dt_gp_eval<JAC>(K, S.pose + 7 * (size_t)f0, S.pose + 7 * (size_t)(f0 + 1), S.glb, g_t[i], g_pos + 3 * (size_t)i, ts_imu[f0], ts_imu[f0 + 1], g_inv_std[i], r, J);
  for (int c = 0; c < 3; ++c) r_out[3 * (size_t)i + c] = r[c];
  if (JAC) for (int c = 0; c < 51; ++c) J_out[51 * (size_t)i + c] = J[c];
}

// cost4 = 1/2 sum rho: one block, fixed summation order
__global__ void __launch_bounds__(1024) k_dt_cost(double cauchy_a, int nF, const double* __restrict__ f_sq, size_t n_imu, const double* __restrict__ r_imu, size_t n_gp,
                                                  const double* __restrict__ r_gp, double* __restrict__ cost) {
  __shared__ double sh[3][32];
  double c[3] = {0, 0, 0};
  const double b = cauchy_a * cauchy_a;
  for (int f = threadIdx.x; f < nF; f += blockDim.x) c[0] += 0.5 * b * log1p(f_sq[f] / b);   // CauchyLoss: rho(s) = b log(1 + s / b)
  for (size_t i = threadIdx.x; i < n_imu; i += blockDim.x) c[1] += 0.5 * r_imu[i] * r_imu[i];
  for (size_t i = threadIdx.x; i < n_gp; i += blockDim.x) c[2] += 0.5 * r_gp[i] * r_gp[i];
  for (int k = 0; k < 3; ++k) { c[k] = warp_sum(c[k]); if ((threadIdx.x & 31) == 0) sh[k][threadIdx.x >> 5] = c[k]; }
  __syncthreads();
  if (threadIdx.x < 32) {
    double t[3];
    for (int k = 0; k < 3; ++k) t[k] = warp_sum(sh[k][threadIdx.x]);
    if (threadIdx.x == 0) { cost[0] = t[0]; cost[1] = t[1]; cost[2] = t[2]; cost[3] = t[0] + t[1] + t[2]; }
  }
}

// ================= direct normal equations of an accepted state: H0, gradient =================
// Determinism rule of this file: within one kernel pass every output entry has exactly ONE writer, and the passes run in stream order —
// so the floating-point sums have a fixed order although the writes use atomicAdd (IMU factors and global-position runs of even / odd
// first state go in separate passes; many-to-one sums are tree reductions over partial buffers).
__global__ void k_dt_frame_w(double cauchy_a, int nF, const double* __restrict__ f_sq, double* __restrict__ f_w) {
  const int f = blockIdx.x * blockDim.x + threadIdx.x;
  if (f < nF) f_w[f] = sqrt(1.0 / (1.0 + f_sq[f] / (cauchy_a * cauchy_a)));   // rho'' < 0: residual and Jacobian scaled by sqrt(rho')
}
// one block per measurement frame, one thread per output entry: pose block (21), pose x global (48), gradient (6 + 8), global x global
// partial (36).  All threads walk the frame's records together (broadcast loads).
constexpr int kCpFrame = 44;   // per-frame / per-landmark partial: 36 corner entries (8 x 8 lower) + 8 gradient / rhs entries
__global__ void __launch_bounds__(128) k_dt_direct_frame(BandLayout H0, int nF, const int32_t* __restrict__ f_state, const int32_t* __restrict__ f_obs_begin,
                                                         const double* __restrict__ f_w, const double* __restrict__ rec, GAct ga, int any_global,
                                                         double* __restrict__ grad, double* __restrict__ cp_frame) {
  const int f = blockIdx.x, e = threadIdx.x;
  constexpr int kTile = 64;
  __shared__ double tile[kTile * 36];
  int ci = 0, cj = -1, kind = 0;   // kind 0: H entry (ci, cj) ; 1: gradient of column ci
  bool live = true;
  if (e < 21) { int i = 0; while ((i + 1) * (i + 2) / 2 <= e) ++i; ci = i; cj = e - i * (i + 1) / 2; kind = 0; }
  else if (e < 69) { ci = 6 + (e - 21) % 8; cj = (e - 21) / 8; kind = 0; }
  else if (e < 75) { ci = e - 69; kind = 1; }
  else if (e < 83) { ci = 6 + (e - 75); kind = 1; }
  else if (e < 119) { const int q = e - 83; int i = 0; while ((i + 1) * (i + 2) / 2 <= q) ++i; ci = 6 + i; cj = 6 + q - i * (i + 1) / 2; kind = 0; }
  else live = false;
  const bool glob = ci >= 6;
  if (live && glob && (!any_global || !ga.a[ci - 6] || (cj >= 6 && !ga.a[cj - 6]))) {
    if (e >= 83 && any_global) cp_frame[(size_t)f * kCpFrame + (e - 83)] = 0.0;
    if (e >= 75 && e < 83 && any_global) cp_frame[(size_t)f * kCpFrame + 36 + (e - 75)] = 0.0;
    live = false;
  }
  double v0 = 0.0, v1 = 0.0;
  const int o0 = f_obs_begin[f], o1 = f_obs_begin[f + 1];
  for (int t0 = o0; t0 < o1; t0 += kTile) {
    const int nt = min(kTile, o1 - t0);
    __syncthreads();
    // the tile's records, copied by the whole block with 256-bit accesses (contiguous in memory)
    for (int q = threadIdx.x; q < nt * 9; q += 128) {
      double v[4];
      ld256c(rec + 36 * (size_t)t0 + 4 * q, v);
      tile[4 * q] = v[0]; tile[4 * q + 1] = v[1]; tile[4 * q + 2] = v[2]; tile[4 * q + 3] = v[3];
    }
    __syncthreads();
    if (live) {
      if (kind == 0) {
        int o = 0;
        for (; o + 1 < nt; o += 2) {
          const double* R = tile + 36 * o;
          v0 += R[2 + ci] * R[2 + cj] + R[19 + ci] * R[19 + cj];
          v1 += R[38 + ci] * R[38 + cj] + R[55 + ci] * R[55 + cj];
        }
        if (o < nt) { const double* R = tile + 36 * o; v0 += R[2 + ci] * R[2 + cj] + R[19 + ci] * R[19 + cj]; }
      } else {
        for (int o = 0; o < nt; ++o) { const double* R = tile + 36 * o; v0 += R[2 + ci] * R[0] + R[19 + ci] * R[1]; }
      }
    }
  }
  if (!live) return;
  const double w2 = f_w[f] * f_w[f];
  const double v = (v0 + v1) * w2;
  const int base = 15 * f_state[f];
  if (e < 21) atomicAdd(H0.band + band_addr(H0, base + ci, base + cj), v);
  else if (e < 69) atomicAdd(H0.border + (size_t)(ci - 6) * H0.n_pad + base + cj, v);
  else if (e < 75) atomicAdd(grad + base + ci, v);
  else if (e < 83) cp_frame[(size_t)f * kCpFrame + 36 + (e - 75)] = v;
  else cp_frame[(size_t)f * kCpFrame + (e - 83)] = v;
}
// one warp per IMU factor whose first state has the given parity: 30 x 30 block of J^T J (lower part) and J^T r
__global__ void __launch_bounds__(128) k_dt_direct_imu(BandLayout H0, int nI, int parity, const int32_t* __restrict__ i_frame0, const double* __restrict__ r,
                                                       const double* __restrict__ J, double* __restrict__ grad) {
  const int k = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (k >= nI || (i_frame0[k] & 1) != parity) return;
  const double* Jk = J + 450 * (size_t)k; const double* rk = r + 15 * (size_t)k;
  const int base = 15 * i_frame0[k];
  for (int e = lane; e < 465; e += 32) {
    int i = (int)((sqrt(8.0 * e + 1.0) - 1.0) * 0.5);
    while ((i + 1) * (i + 2) / 2 <= e) ++i;
    while (i * (i + 1) / 2 > e) --i;
    const int j = e - i * (i + 1) / 2;
    double v = 0;
    for (int row = 0; row < 15; ++row) v += Jk[30 * row + i] * Jk[30 * row + j];
    atomicAdd(H0.band + band_addr(H0, base + i, base + j), v);
  }
  if (lane < 30) {
    double g = 0;
    for (int row = 0; row < 15; ++row) g += Jk[30 * row + lane] * rk[row];
    atomicAdd(grad + base + lane, g);
  }
}
__device__ __forceinline__ int gp_index(int c, int f0, int n_pad) {   // local column of the GP term -> unified index
  if (c < 6) return 15 * f0 + c;
  if (c < 12) return 15 * (f0 + 1) + (c - 6);
  if (c == 12) return n_pad + DGT_SCALE;
  if (c == 13) return n_pad + DGT_TGS;
  return n_pad + DGT_PBP + (c - 14);
}
// one block per run of global-position samples between the same two states (first state of the given parity), one thread per output
// entry: 17 x 17 lower (153) + gradient (17); the samples of the run are summed in their stored order.  Global x global entries and the
// globals' gradient go to the run's partial (15 + 5).
constexpr int kCpGp = 20;
__global__ void __launch_bounds__(192) k_dt_direct_gp(BandLayout H0, int nRuns, int parity, const int32_t* __restrict__ run_begin, const int32_t* __restrict__ gp_order,
                                                      const double* __restrict__ r, const double* __restrict__ J, GAct ga, double* __restrict__ grad,
                                                      double* __restrict__ cp_gp) {
  const int f0 = blockIdx.x, e = threadIdx.x;
  if (f0 >= nRuns || (f0 & 1) != parity || e >= 170) return;
  const int b0 = run_begin[f0], b1 = run_begin[f0 + 1];
  int i, j = -1;
  if (e < 153) { i = 0; while ((i + 1) * (i + 2) / 2 <= e) ++i; j = e - i * (i + 1) / 2; } else i = e - 153;
  const int ui = gp_index(i, f0, H0.n_pad), uj = j >= 0 ? gp_index(j, f0, H0.n_pad) : -1;
  const bool off = (ui >= H0.n_pad && !ga.a[ui - H0.n_pad]) || (uj >= H0.n_pad && !ga.a[uj - H0.n_pad]);
  double v = 0.0;
  if (!off)
    for (int q = b0; q < b1; ++q) {
      const size_t s = gp_order[q];
      const double* Js = J + 51 * s;
      if (j >= 0) v += Js[i] * Js[j] + Js[17 + i] * Js[17 + j] + Js[34 + i] * Js[34 + j];
      else v += Js[i] * r[3 * s] + Js[17 + i] * r[3 * s + 1] + Js[34 + i] * r[3 * s + 2];
    }
  if (j >= 0) {
    if (i >= 12 && j >= 12) { const int gi = i - 12, gj = j - 12; cp_gp[(size_t)f0 * kCpGp + gi * (gi + 1) / 2 + gj] = v; }
    else if (!off && b1 > b0) add_sym(H0, ui, uj, v);
  } else {
    if (i >= 12) cp_gp[(size_t)f0 * kCpGp + 15 + (i - 12)] = v;
    else if (b1 > b0) atomicAdd(grad + ui, v);
  }
}
// one thread per landmark: V0 = sum J_l^T J_l (xx xy xz yy yz zz), gl = sum J_l^T r   (list order: fixed)
__global__ void k_dt_lm_blocks(int L, const int32_t* __restrict__ lm_begin, const int32_t* __restrict__ lm_obs, const int32_t* __restrict__ o_frame,
                               const double* __restrict__ f_w, const double* __restrict__ rec, double* __restrict__ V0, double* __restrict__ gl) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  double V[6] = {0, 0, 0, 0, 0, 0}, g[3] = {0, 0, 0};
  for (int e = lm_begin[l]; e < lm_begin[l + 1]; ++e) {
    const int o = lm_obs[e];
    const double w = f_w[o_frame[o]], w2 = w * w;
    const double* R = rec + 36 * (size_t)o;
    const double a0 = R[16], a1 = R[17], a2 = R[18], b0 = R[33], b1 = R[34], b2 = R[35];
    V[0] += w2 * (a0 * a0 + b0 * b0); V[1] += w2 * (a0 * a1 + b0 * b1); V[2] += w2 * (a0 * a2 + b0 * b2);
    V[3] += w2 * (a1 * a1 + b1 * b1); V[4] += w2 * (a1 * a2 + b1 * b2); V[5] += w2 * (a2 * a2 + b2 * b2);
    g[0] += w2 * (a0 * R[0] + b0 * R[1]); g[1] += w2 * (a1 * R[0] + b1 * R[1]); g[2] += w2 * (a2 * R[0] + b2 * R[1]);
  }
  for (int k = 0; k < 6; ++k) V0[6 * (size_t)l + k] = V[k];
  for (int k = 0; k < 3; ++k) gl[3 * (size_t)l + k] = g[k];
}
// Tree reductions (fixed order) of the per-source partials into the 12 x 12 global corner and the globals' gradient / right-hand side.
// One block per output: corner entry (t1 >= t2) q = t1 (t1+1)/2 + t2 < 78, or gradient entry 78 + t.
//   mode 0 (accepted state): corner = frames + GP runs -> H.corner (=), grad[n_pad + t] (=)
//   mode 1 (LM iteration):   landmark partials          -> H.corner (+=), rhs_acc[n_pad + t] (=)
__global__ void __launch_bounds__(256) k_dt_corner_reduce(int mode, BandLayout H, int nA, const double* __restrict__ cpA /* frames | landmarks: 44 each */, int nRuns,
                                                          const double* __restrict__ cp_gp, double* __restrict__ vec /* grad | rhs_acc */) {
  __shared__ double sh[256];
  const int q = blockIdx.x;
  int t1, t2 = -1;
  if (q < 78) { t1 = 0; while ((t1 + 1) * (t1 + 2) / 2 <= q) ++t1; t2 = q - t1 * (t1 + 1) / 2; } else t1 = q - 78;
  double s = 0.0;
  if (cpA && t1 < 8) {
    const int idx = t2 >= 0 ? t1 * (t1 + 1) / 2 + t2 : 36 + t1;
    for (int i = threadIdx.x; i < nA; i += 256) s += cpA[(size_t)i * kCpFrame + idx];
  }
  sh[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
  double total = sh[0];
  __syncthreads();
  if (mode == 0 && cp_gp) {
    auto g_of = [](int t) { return t == DGT_SCALE ? 0 : t == DGT_TGS ? 1 : (t >= DGT_PBP && t < DGT_PBP + 3) ? 2 + (t - DGT_PBP) : -1; };
    const int g1 = g_of(t1), g2 = t2 >= 0 ? g_of(t2) : 0;
    if (g1 >= 0 && g2 >= 0) {
      const int a = g1 > g2 ? g1 : g2, b = g1 > g2 ? g2 : g1;
      const int idx = t2 >= 0 ? a * (a + 1) / 2 + b : 15 + g1;
      double s2 = 0.0;
      for (int i = threadIdx.x; i < nRuns; i += 256) s2 += cp_gp[(size_t)i * kCpGp + idx];
      sh[threadIdx.x] = s2;
      __syncthreads();
      for (int o = 128; o > 0; o >>= 1) { if (threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o]; __syncthreads(); }
      total += sh[0];
    }
  }
  if (threadIdx.x == 0) {
    if (t2 >= 0) { if (mode == 0) H.corner[t1 * kBorder + t2] = total; else H.corner[t1 * kBorder + t2] += total; }
    else vec[H.n_pad + t1] = total;
  }
}
__global__ void k_dt_diag(BandLayout H0, double* __restrict__ diag) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < H0.n_pad) diag[i] = H0.band[band_addr(H0, i, i)];
  else if (i < H0.n_pad + kBorder) { const int t = i - H0.n_pad; diag[i] = t < GT_SIZE ? H0.corner[t * kBorder + t] : 0.0; }
}
__global__ void k_dt_jscale(int n, const double* __restrict__ diag, int L, const double* __restrict__ V0, double* __restrict__ jscale, double* __restrict__ jscale_lm) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) jscale[i] = 1.0 / (1.0 + sqrt(fmax(diag[i], 0.0)));
  if (i < 3 * L) {
    const int di[3] = {0, 3, 5};
    jscale_lm[i] = 1.0 / (1.0 + sqrt(fmax(V0[6 * (size_t)(i / 3) + di[i % 3]], 0.0)));
  }
}

// ================= normal equations of one LM iteration: Hw = H0 + D - Schur(radius) =================
__device__ __forceinline__ double lm_damping(double h, double sc, double radius) { return fmin(fmax(h * sc * sc, 1e-6), 1e32) / (radius * sc * sc); }
__device__ __forceinline__ void sym3_mul(const double* Vi, double x, double y, double z, double* o) {   // o = V^-1 [x y z]
  o[0] = Vi[0] * x + Vi[1] * y + Vi[2] * z; o[1] = Vi[1] * x + Vi[3] * y + Vi[4] * z; o[2] = Vi[2] * x + Vi[4] * y + Vi[5] * z;
}
// one warp per landmark: V^-1 (damped), V^-1 g, per point W_p (pose coupling 6 x 3) and Y_p = W_p V^-1, summed global coupling W_g (8 x 3),
// the landmark's partial of the global corner / right-hand side
__global__ void __launch_bounds__(128) k_dt_schur_prepare(int L, double radius, const int32_t* __restrict__ lm_begin, const int32_t* __restrict__ lm_obs,
                                                          const int32_t* __restrict__ o_frame, const double* __restrict__ f_w, const double* __restrict__ rec,
                                                          const double* __restrict__ V0, const double* __restrict__ gl, const double* __restrict__ jscale_lm, GAct ga,
                                                          int any_global, double* __restrict__ Vi, double* __restrict__ Vg, double* __restrict__ WgS,
                                                          double* __restrict__ Wp, double* __restrict__ Yp, double* __restrict__ cp_lm, int32_t* __restrict__ flag) {
  const int l = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (l >= L) return;
  const int e0 = lm_begin[l], e1 = lm_begin[l + 1];
  if (e0 == e1) {
    if (any_global) for (int k = lane; k < kCpFrame; k += 32) cp_lm[(size_t)l * kCpFrame + k] = 0.0;
    return;
  }
  double Vinv[6];
  {
    const double* V = V0 + 6 * (size_t)l; const double* s = jscale_lm + 3 * (size_t)l;
    const double a = V[0] + lm_damping(V[0], s[0], radius), b = V[1], c = V[2], d = V[3] + lm_damping(V[3], s[1], radius), e = V[4], f = V[5] + lm_damping(V[5], s[2], radius);
    const double c00 = d * f - e * e, c01 = c * e - b * f, c02 = b * e - c * d;
    const double det = a * c00 + b * c01 + c * c02;
    if (lane == 0 && (!(det > 0.0) || !isfinite(det))) atomicExch(flag, 1);
    const double id = 1.0 / det;
    Vinv[0] = c00 * id; Vinv[1] = c01 * id; Vinv[2] = c02 * id; Vinv[3] = (a * f - c * c) * id; Vinv[4] = (b * c - a * e) * id; Vinv[5] = (a * d - b * b) * id;
  }
  double vg[3];
  sym3_mul(Vinv, gl[3 * (size_t)l], gl[3 * (size_t)l + 1], gl[3 * (size_t)l + 2], vg);
  if (lane == 0) {
    for (int k = 0; k < 6; ++k) Vi[6 * (size_t)l + k] = Vinv[k];
    for (int k = 0; k < 3; ++k) Vg[3 * (size_t)l + k] = vg[k];
  }
  double wg[24];
#pragma unroll
  for (int k = 0; k < 24; ++k) wg[k] = 0.0;
  for (int e = e0 + lane; e < e1; e += 32) {
    const int o = lm_obs[e];
    const double w = f_w[o_frame[o]], w2 = w * w;
    double R[36];
#pragma unroll
    for (int k = 0; k < 9; ++k) ld256c(rec + 36 * (size_t)o + 4 * k, R + 4 * k);
    const double a0 = R[16], a1 = R[17], a2 = R[18], b0 = R[33], b1 = R[34], b2 = R[35];
    double Wl[kWS], Yl[kWS];
    Wl[18] = Wl[19] = Yl[18] = Yl[19] = 0.0;
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      const double ja = w2 * R[2 + k], jb = w2 * R[19 + k];
      const double x = ja * a0 + jb * b0, y = ja * a1 + jb * b1, z = ja * a2 + jb * b2;
      Wl[3 * k] = x; Wl[3 * k + 1] = y; Wl[3 * k + 2] = z;
      sym3_mul(Vinv, x, y, z, Yl + 3 * k);
    }
#pragma unroll
    for (int k = 0; k < 5; ++k) { st256(Wp + kWS * (size_t)o + 4 * k, Wl + 4 * k); st256(Yp + kWS * (size_t)o + 4 * k, Yl + 4 * k); }
    if (any_global) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const double ja = w2 * R[8 + k], jb = w2 * R[25 + k];
        wg[3 * k] += ja * a0 + jb * b0; wg[3 * k + 1] += ja * a1 + jb * b1; wg[3 * k + 2] += ja * a2 + jb * b2;
      }
    }
  }
  if (any_global) {
#pragma unroll
    for (int k = 0; k < 24; ++k) { wg[k] = warp_sum(wg[k]); if (!ga.a[k / 3]) wg[k] = 0.0; }
    if (lane == 0) for (int k = 0; k < 24; ++k) WgS[24 * (size_t)l + k] = wg[k];
    // partial of the corner: -W_g V^-1 W_g^T (lower 36) and of the right-hand side: -W_g V^-1 g (8)
    double yg[24];
#pragma unroll
    for (int k = 0; k < 8; ++k) sym3_mul(Vinv, wg[3 * k], wg[3 * k + 1], wg[3 * k + 2], yg + 3 * k);
    if (lane == 0) {
      double* cp = cp_lm + (size_t)l * kCpFrame;
      int q = 0;
      for (int t1 = 0; t1 < 8; ++t1) for (int t2 = 0; t2 <= t1; ++t2) cp[q++] = -(yg[3 * t1] * wg[3 * t2] + yg[3 * t1 + 1] * wg[3 * t2 + 1] + yg[3 * t1 + 2] * wg[3 * t2 + 2]);
      for (int t = 0; t < 8; ++t) cp[36 + t] = -(wg[3 * t] * vg[0] + wg[3 * t + 1] * vg[1] + wg[3 * t + 2] * vg[2]);
    }
  }
}
// one warp per (state, distance) key: the 6 x 6 pose block -sum Y_i W_j^T over the landmarks seen from both states, summed in the stored
// order, one writer per entry.  distance 0: lower triangle only.
__global__ void __launch_bounds__(128) k_dt_schur_gather(BandLayout B, int nKeys, int T1, const int64_t* __restrict__ key_begin, const int2* __restrict__ pairs,
                                                         const double* __restrict__ Wp, const double* __restrict__ Yp) {
  const int key = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (key >= nKeys) return;
  const int64_t b0 = key_begin[key], b1 = key_begin[key + 1];
  if (b0 == b1) return;
  const int si = key / T1, d = key % T1;
  double acc[36];
#pragma unroll
  for (int k = 0; k < 36; ++k) acc[k] = 0.0;
  for (int64_t q = b0 + lane; q < b1; q += 32) {
    const int2 pr = pairs[q];
    double w[kWS], Y[kWS];
#pragma unroll
    for (int k = 0; k < 5; ++k) { ld256c(Yp + kWS * (size_t)pr.x + 4 * k, Y + 4 * k); ld256c(Wp + kWS * (size_t)pr.y + 4 * k, w + 4 * k); }
#pragma unroll
    for (int a = 0; a < 6; ++a) {
      const double y0 = Y[3 * a], y1 = Y[3 * a + 1], y2 = Y[3 * a + 2];
#pragma unroll
      for (int b = 0; b < 6; ++b) acc[6 * a + b] -= y0 * w[3 * b] + y1 * w[3 * b + 1] + y2 * w[3 * b + 2];
    }
  }
  double mine = 0.0, mine2 = 0.0;   // lane k holds entry k, lanes 0..3 also hold entries 32..35
#pragma unroll
  for (int k = 0; k < 36; ++k) { const double v = warp_sum(acc[k]); if (k < 32) { if (lane == k) mine = v; } else if (lane == k - 32) mine2 = v; }
  const int bi = 15 * si, bj = 15 * (si - d);
  {
    const int a = lane / 6, b = lane % 6;
    if (!(d == 0 && b > a)) atomicAdd(B.band + band_addr(B, bi + a, bj + b), mine);
  }
  if (lane < 4) {
    const int k = 32 + lane, a = k / 6, b = k % 6;
    if (!(d == 0 && b > a)) atomicAdd(B.band + band_addr(B, bi + a, bj + b), mine2);
  }
}
// one block per measurement frame: Schur part of pose x global (48) and of the right-hand side (6).  Threads take the frame's points
// round-robin and keep the 54 sums in registers; lanes are combined by a shuffle tree, warps in warp order (fixed order).
__global__ void __launch_bounds__(128) k_dt_schur_frame(BandLayout B, int nF, const int32_t* __restrict__ f_state, const int32_t* __restrict__ f_obs_begin,
                                                        const int32_t* __restrict__ o_lm, const double* __restrict__ Wp, const double* __restrict__ Yp,
                                                        const double* __restrict__ WgS, const double* __restrict__ Vg, GAct ga, int any_global, double* __restrict__ rhs_acc) {
  __shared__ double sh[4][54];
  const int f = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  double acc[54];
#pragma unroll
  for (int k = 0; k < 54; ++k) acc[k] = 0.0;
  for (int o = f_obs_begin[f] + tid; o < f_obs_begin[f + 1]; o += 128) {
    double W[kWS], Y[kWS];
#pragma unroll
    for (int k = 0; k < 5; ++k) { ld256c(Wp + kWS * (size_t)o + 4 * k, W + 4 * k); ld256c(Yp + kWS * (size_t)o + 4 * k, Y + 4 * k); }
    const int l = o_lm[o];
    const double g0 = Vg[3 * (size_t)l], g1 = Vg[3 * (size_t)l + 1], g2 = Vg[3 * (size_t)l + 2];
#pragma unroll
    for (int a = 0; a < 6; ++a) acc[48 + a] -= W[3 * a] * g0 + W[3 * a + 1] * g1 + W[3 * a + 2] * g2;
    if (any_global) {
      double G[24];
#pragma unroll
      for (int k = 0; k < 6; ++k) ld256c(WgS + 24 * (size_t)l + 4 * k, G + 4 * k);
#pragma unroll
      for (int b = 0; b < 8; ++b)
#pragma unroll
        for (int a = 0; a < 6; ++a) acc[6 * b + a] -= Y[3 * a] * G[3 * b] + Y[3 * a + 1] * G[3 * b + 1] + Y[3 * a + 2] * G[3 * b + 2];
    }
  }
#pragma unroll
  for (int k = 0; k < 54; ++k) { const double v = warp_sum(acc[k]); if (lane == 0) sh[warp][k] = v; }
  __syncthreads();
  if (tid < 54) {
    const double v = ((sh[0][tid] + sh[1][tid]) + sh[2][tid]) + sh[3][tid];
    const int base = 15 * f_state[f];
    if (tid < 48) { const int a = tid % 6, b = tid / 6; if (any_global && ga.a[b]) atomicAdd(B.border + (size_t)b * B.n_pad + base + a, v); }
    else atomicAdd(rhs_acc + base + (tid - 48), v);
  }
}
// damping, unit rows for the padding / inactive unknowns, right-hand side
__global__ void k_dt_finish(BandLayout B, int n_band, const uint8_t* __restrict__ sb_active, const double* __restrict__ jscale, const double* __restrict__ diag,
                            const double* __restrict__ grad, const double* __restrict__ rhs_acc, double radius, GAct ga, double* __restrict__ dvec) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B.n_pad) {
    bool active = i < n_band;
    if (active) { const int f = i / 15, k = i % 15; active = k < 6 || sb_active[f]; }
    const size_t ad = band_addr(B, i, i);
    if (active) {
      const double dd = lm_damping(diag[i], jscale[i], radius);
      B.band[ad] += dd; dvec[i] = dd;
      B.border[(size_t)kRhsRow * B.n_pad + i] = -(grad[i] + rhs_acc[i]);
    } else {
      B.band[ad] = 1.0; dvec[i] = 0.0;
      B.border[(size_t)kRhsRow * B.n_pad + i] = 0.0;
    }
  } else if (i < B.n_pad + GT_SIZE) {
    const int t = i - B.n_pad;
    if (ga.a[t]) {
      const double dd = lm_damping(diag[i], jscale[i], radius);
      B.corner[t * kBorder + t] += dd; dvec[i] = dd;
      B.corner[kRhsRow * kBorder + t] = -(grad[i] + rhs_acc[i]);
    } else {
      B.corner[t * kBorder + t] = 1.0; dvec[i] = 0.0;
      B.corner[kRhsRow * kBorder + t] = 0.0;
    }
  }
}

// ================= step =================
// one thread per landmark: delta_l = -V^-1 (g_l + sum_p W_p^T delta_pose + W_g^T delta_glob)
__global__ void k_dt_backsub(int L, const int32_t* __restrict__ lm_begin, const int32_t* __restrict__ lm_obs, const int32_t* __restrict__ o_frame,
                             const int32_t* __restrict__ f_state, const double* __restrict__ Vi, const double* __restrict__ WgS, const double* __restrict__ Wp,
                             const double* __restrict__ gl, const double* __restrict__ delta, int n_pad, int any_global, double* __restrict__ delta_lm) {
  const int l = blockIdx.x * blockDim.x + threadIdx.x;
  if (l >= L) return;
  double s[3] = {0, 0, 0};
  if (lm_begin[l] == lm_begin[l + 1]) { for (int k = 0; k < 3; ++k) delta_lm[3 * (size_t)l + k] = 0.0; return; }
  for (int k = 0; k < 3; ++k) s[k] = gl[3 * (size_t)l + k];
  for (int e = lm_begin[l]; e < lm_begin[l + 1]; ++e) {
    const int o = lm_obs[e];
    const double* d = delta + 15 * (size_t)f_state[o_frame[o]];
    const double* W = Wp + kWS * (size_t)o;
#pragma unroll
    for (int a = 0; a < 6; ++a) { s[0] += W[3 * a] * d[a]; s[1] += W[3 * a + 1] * d[a]; s[2] += W[3 * a + 2] * d[a]; }
  }
  if (any_global) {
    const double* G = WgS + 24 * (size_t)l;
    for (int b = 0; b < 8; ++b) { const double d = delta[n_pad + b]; s[0] += G[3 * b] * d; s[1] += G[3 * b + 1] * d; s[2] += G[3 * b + 2] * d; }
  }
  double o3[3];
  sym3_mul(Vi + 6 * (size_t)l, s[0], s[1], s[2], o3);
  for (int k = 0; k < 3; ++k) delta_lm[3 * (size_t)l + k] = -o3[k];
}
// Candidate state + the scalars of the step.  mode 0: candidate = Plus(x, delta) with model cost change and step norm; mode 1: x norm and
// gradient max norm (max | x - Plus(x, -g) | over the ambient coordinates of the free blocks) at the state S.
__global__ void __launch_bounds__(256) k_dt_step(int mode, int F, int L, int n_pad, const uint8_t* __restrict__ sb_active, const uint8_t* __restrict__ lm_active, GAct ga,
                                                 DtStateP S, DtStateP C, const double* __restrict__ delta, const double* __restrict__ delta_lm, const double* __restrict__ grad,
                                                 const double* __restrict__ gl, const double* __restrict__ dvec, const double* __restrict__ V0,
                                                 const double* __restrict__ jscale_lm, double radius, double* __restrict__ scal) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  double model = 0.0, step2 = 0.0, xn2 = 0.0, gmax = 0.0;
  if (i < F) {
    const double* T = S.pose + 7 * (size_t)i;
    double d[6], out[7];
    for (int k = 0; k < 6; ++k) d[k] = mode == 0 ? delta[15 * (size_t)i + k] : -grad[15 * (size_t)i + k];
    dt_se3_plus(T, d, out);
    for (int k = 0; k < 7; ++k) {
      const double df = out[k] - T[k];
      if (mode == 0) { C.pose[7 * (size_t)i + k] = out[k]; step2 += df * df; } else { xn2 += T[k] * T[k]; gmax = fmax(gmax, fabs(df)); }
    }
    if (mode == 0) for (int k = 0; k < 6; ++k) model += d[k] * (dvec[15 * (size_t)i + k] * d[k] - grad[15 * (size_t)i + k]);
    const double* x = S.sb + 9 * (size_t)i;
    if (sb_active[i]) {
      for (int k = 0; k < 9; ++k) {
        const size_t u = 15 * (size_t)i + 6 + k;
        if (mode == 0) { const double dk = delta[u]; C.sb[9 * (size_t)i + k] = x[k] + dk; const double df = (x[k] + dk) - x[k]; step2 += df * df; model += dk * (dvec[u] * dk - grad[u]); }
        else { xn2 += x[k] * x[k]; gmax = fmax(gmax, fabs((x[k] - grad[u]) - x[k])); }
      }
    } else if (mode == 0) {
      for (int k = 0; k < 9; ++k) C.sb[9 * (size_t)i + k] = x[k];
    }
  } else if (i < F + L) {
    const int l = i - F;
    const double* x = S.lm + 3 * (size_t)l;
    if (lm_active[l]) {
      const int di[3] = {0, 3, 5};
      for (int k = 0; k < 3; ++k) {
        const size_t u = 3 * (size_t)l + k;
        if (mode == 0) {
          const double dk = delta_lm[u]; C.lm[u] = x[k] + dk; const double df = (x[k] + dk) - x[k]; step2 += df * df;
          model += dk * (lm_damping(V0[6 * (size_t)l + di[k]], jscale_lm[u], radius) * dk - gl[u]);
        } else { xn2 += x[k] * x[k]; gmax = fmax(gmax, fabs((x[k] - gl[u]) - x[k])); }
      }
    } else if (mode == 0) {
      for (int k = 0; k < 3; ++k) C.lm[3 * (size_t)l + k] = x[k];
    }
  } else if (i == F + L) {
    const double* g = S.glb;
    const double* dl = mode == 0 ? delta + n_pad : grad + n_pad;
    const double sg = mode == 0 ? 1.0 : -1.0;
    double out[DG_SIZE];
    for (int k = 0; k < DG_SIZE; ++k) out[k] = g[k];
    if (ga.a[DGT_SCALE]) out[DG_SCALE] = g[DG_SCALE] + sg * dl[DGT_SCALE];
    if (ga.a[DGT_RBC]) {
      const quat q = quat_mul(quat{g[DG_RBC], g[DG_RBC + 1], g[DG_RBC + 2], g[DG_RBC + 3]}, so3_exp(v3{sg * dl[DGT_RBC], sg * dl[DGT_RBC + 1], sg * dl[DGT_RBC + 2]}));
      out[DG_RBC] = q.x; out[DG_RBC + 1] = q.y; out[DG_RBC + 2] = q.z; out[DG_RBC + 3] = q.w;
    }
    if (ga.a[DGT_PBC]) for (int k = 0; k < 3; ++k) out[DG_PBC + k] = g[DG_PBC + k] + sg * dl[DGT_PBC + k];
    if (ga.a[DGT_TCAM]) out[DG_TCAM] = g[DG_TCAM] + sg * dl[DGT_TCAM];
    if (ga.a[DGT_TGS]) out[DG_TGS] = g[DG_TGS] + sg * dl[DGT_TGS];
    if (ga.a[DGT_PBP]) for (int k = 0; k < 3; ++k) out[DG_PBP + k] = g[DG_PBP + k] + sg * dl[DGT_PBP + k];
    const int amb0[6] = {DG_SCALE, DG_RBC, DG_PBC, DG_TCAM, DG_TGS, DG_PBP}, ambn[6] = {1, 4, 3, 1, 1, 3}, tan0[6] = {DGT_SCALE, DGT_RBC, DGT_PBC, DGT_TCAM, DGT_TGS, DGT_PBP};
    for (int b = 0; b < 6; ++b) {
      if (!ga.a[tan0[b]]) continue;
      for (int k = 0; k < ambn[b]; ++k) {
        const double df = out[amb0[b] + k] - g[amb0[b] + k];
        if (mode == 0) step2 += df * df; else { xn2 += g[amb0[b] + k] * g[amb0[b] + k]; gmax = fmax(gmax, fabs(df)); }
      }
    }
    if (mode == 0) {
      for (int k = 0; k < DG_SIZE; ++k) C.glb[k] = out[k];
      for (int t = 0; t < DGT_COUNT; ++t) if (ga.a[t]) model += dl[t] * (dvec[n_pad + t] * dl[t] - grad[n_pad + t]);
    }
  }
  model = warp_sum(model); step2 = warp_sum(step2); xn2 = warp_sum(xn2);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) gmax = fmax(gmax, __shfl_xor_sync(0xffffffffu, gmax, o));
  if ((threadIdx.x & 31) == 0) {   // per-warp partials; k_dt_scal_reduce sums them in warp order
    double* o = scal + 4 * (size_t)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    o[0] = 0.5 * model; o[1] = step2; o[2] = xn2; o[3] = gmax;
  }
}
__global__ void __launch_bounds__(256) k_dt_scal_reduce(int nw, const double* __restrict__ part, double* __restrict__ scal) {
  __shared__ double sh[4][256];
  double a[4] = {0, 0, 0, 0};
  for (int i = threadIdx.x; i < nw; i += 256) { a[0] += part[4 * (size_t)i]; a[1] += part[4 * (size_t)i + 1]; a[2] += part[4 * (size_t)i + 2]; a[3] = fmax(a[3], part[4 * (size_t)i + 3]); }
  for (int k = 0; k < 4; ++k) sh[k][threadIdx.x] = a[k];
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) { for (int k = 0; k < 3; ++k) sh[k][threadIdx.x] += sh[k][threadIdx.x + o]; sh[3][threadIdx.x] = fmax(sh[3][threadIdx.x], sh[3][threadIdx.x + o]); }
    __syncthreads();
  }
  if (threadIdx.x == 0) { scal[DSC_MODEL] = sh[0][0]; scal[DSC_STEP2] = sh[1][0]; scal[DSC_XNORM2] = sh[2][0]; scal[DSC_GMAX] = sh[3][0]; }
}

// unit-noise error norms for evaluate*Errors
__global__ void k_dt_norms(size_t n, int width, const double* __restrict__ r, int stride, const double* __restrict__ inv_std_per, const int32_t* __restrict__ owner,
                           double* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  double s = 0;
  for (int k = 0; k < width; ++k) s += r[(size_t)stride * i + k] * r[(size_t)stride * i + k];
  const double inv = inv_std_per[owner ? owner[i] : i];
  out[i] = sqrt(s) / inv;
}

// ================= host side =================
int dt_sync(ctba_dt_problem* p) {
  cudaStream_t st = p->stream;
  if (!p->have_camera) { p->err = "set_camera first"; return CTBA_E_STATE; }
  if (p->F() < 2) { p->err = "initialize_states first (at least two states)"; return CTBA_E_STATE; }
  if (p->meas_dirty) {
    std::vector<int32_t> o_frame(p->M());
    for (size_t f = 0; f + 1 < p->f_obs_begin.size(); ++f) for (int o = p->f_obs_begin[f]; o < p->f_obs_begin[f + 1]; ++o) o_frame[o] = (int32_t)f;
    DT_CUDA(p, p->d_f_state.upload(p->f_state, st)); DT_CUDA(p, p->d_f_obs_begin.upload(p->f_obs_begin, st)); DT_CUDA(p, p->d_o_lm.upload(p->o_lm, st));
    DT_CUDA(p, p->d_o_frame.upload(o_frame, st)); DT_CUDA(p, p->d_f_inv_std.upload(p->f_inv_std, st)); DT_CUDA(p, p->d_o_uv.upload(p->o_uv, st));
    DT_CUDA(p, p->d_o_vel.upload(p->o_vel, st)); DT_CUDA(p, p->d_ts_imu.upload(p->ts_imu, st));
    DT_CUDA(p, p->d_i_frame0.upload(p->i_frame0, st)); DT_CUDA(p, p->d_i_begin.upload(p->i_begin, st)); DT_CUDA(p, p->d_i_t.upload(p->i_t, st));
    DT_CUDA(p, p->d_i_gyr.upload(p->i_gyr, st)); DT_CUDA(p, p->d_i_acc.upload(p->i_acc, st));
    DT_CUDA(p, p->d_g_frame0.upload(p->g_frame0, st)); DT_CUDA(p, p->d_g_t.upload(p->g_t, st)); DT_CUDA(p, p->d_g_pos.upload(p->g_pos, st));
    DT_CUDA(p, p->d_g_inv_std.upload(p->g_inv_std, st));
    DT_CUDA(p, cudaStreamSynchronize(st));   // o_frame is a local
    // new factors get a fresh cache (redo = 1), existing factors keep theirs
    const size_t nI = p->nI();
    if (nI > p->n_cache) {
      Dev<char> nc;
      DT_CUDA(p, nc.resize(nI * sizeof(DtImuCache)));
      if (p->n_cache) DT_CUDA(p, cudaMemcpyAsync(nc.p, p->d_imu_cache.p, p->n_cache * sizeof(DtImuCache), cudaMemcpyDeviceToDevice, st));
      std::vector<DtImuCache> fresh(nI - p->n_cache);
      for (auto& c : fresh) { std::memset(&c, 0, sizeof(c)); c.redo = 1; c.Delta_q[3] = 1.0; }
      DT_CUDA(p, cudaMemcpyAsync(nc.p + p->n_cache * sizeof(DtImuCache), fresh.data(), fresh.size() * sizeof(DtImuCache), cudaMemcpyHostToDevice, st));
      DT_CUDA(p, cudaStreamSynchronize(st));
      p->d_imu_cache.swap(nc);
      p->n_cache = nI;
      DT_CUDA(p, p->d_imu_scratch.resize(675 * nI));
    }
    DT_CUDA(p, p->d_imu_rc.resize(1));
    p->meas_dirty = false;
  }
  if (p->state_dirty) {
    DT_CUDA(p, p->st.pose.upload(p->pose, st)); DT_CUDA(p, p->st.sb.upload(p->sb, st)); DT_CUDA(p, p->st.lm.upload(p->lm, st));
    DT_CUDA(p, p->st.glb.resize(DG_SIZE));
    DT_CUDA(p, cudaMemcpyAsync(p->st.glb.p, p->glb, sizeof(p->glb), cudaMemcpyHostToDevice, st));
    DT_CUDA(p, cudaStreamSynchronize(st));
    p->state_dirty = false; p->host_stale = false;
  }
  return CTBA_OK;
}
int dt_download(ctba_dt_problem* p) {
  if (!p->host_stale) return CTBA_OK;
  cudaStream_t st = p->stream;
  DT_CUDA(p, cudaMemcpyAsync(p->pose.data(), p->st.pose.p, p->pose.size() * 8, cudaMemcpyDeviceToHost, st));
  DT_CUDA(p, cudaMemcpyAsync(p->sb.data(), p->st.sb.p, p->sb.size() * 8, cudaMemcpyDeviceToHost, st));
  if (!p->lm.empty()) DT_CUDA(p, cudaMemcpyAsync(p->lm.data(), p->st.lm.p, p->lm.size() * 8, cudaMemcpyDeviceToHost, st));
  DT_CUDA(p, cudaMemcpyAsync(p->glb, p->st.glb.p, sizeof(p->glb), cudaMemcpyDeviceToHost, st));
  DT_CUDA(p, cudaStreamSynchronize(st));
  p->host_stale = false;
  return CTBA_OK;
}

int dt_eval_buffers(ctba_dt_problem* p, DtEval& ev) {
  DT_CUDA(p, ev.rec.resize(36 * p->M())); DT_CUDA(p, ev.f_sq.resize(p->f_state.size()));
  DT_CUDA(p, ev.r_imu.resize(15 * p->nI())); DT_CUDA(p, ev.J_imu.resize(450 * p->nI()));
  DT_CUDA(p, ev.r_gp.resize(3 * p->nG())); DT_CUDA(p, ev.J_gp.resize(51 * p->nG()));
  DT_CUDA(p, ev.cost.resize(4));
  return CTBA_OK;
}
// residuals (+ Jacobians) of every term at S into ev, cost4 into ev.cost.  `cache` = the factors' caches (or a scratch copy)
int dt_evaluate(ctba_dt_problem* p, DtEval& ev, const DtStateP& S, bool jac, DtImuCache* cache) {
  cudaStream_t st = p->stream;
  int rc = dt_eval_buffers(p, ev); if (rc) return rc;
  const int nF = (int)p->f_state.size(), nI = (int)p->nI(), nG = (int)p->nG();
  if (nF) {
    if (jac) k_dt_reproj<true><<<cdiv((size_t)nF * 32, 128), 128, 0, st>>>(p->K, nF, p->d_f_state.p, p->d_f_obs_begin.p, p->d_f_inv_std.p, p->d_o_uv.p, p->d_o_vel.p, p->d_o_lm.p, S, ev.rec.p, ev.f_sq.p);
    else k_dt_reproj<false><<<cdiv((size_t)nF * 32, 128), 128, 0, st>>>(p->K, nF, p->d_f_state.p, p->d_f_obs_begin.p, p->d_f_inv_std.p, p->d_o_uv.p, p->d_o_vel.p, p->d_o_lm.p, S, ev.rec.p, ev.f_sq.p);
    ++p->launches;
  }
  if (nI) {
    DT_CUDA(p, cudaMemsetAsync(p->d_imu_rc.p, 0, sizeof(int32_t), st));
    k_dt_imu<<<cdiv(nI, 64), 64, 0, st>>>(p->K, nI, p->d_i_frame0.p, p->d_i_begin.p, p->d_i_t.p, p->d_i_gyr.p, p->d_i_acc.p, p->d_ts_imu.p, S, cache, p->d_imu_scratch.p,
                                          ev.r_imu.p, jac ? ev.J_imu.p : nullptr, p->d_imu_rc.p);
    ++p->launches;
  }
  if (nG) {
    if (jac) k_dt_gp<true><<<cdiv(nG, 128), 128, 0, st>>>(p->K, nG, p->d_g_frame0.p, p->d_g_t.p, p->d_g_pos.p, p->d_g_inv_std.p, p->d_ts_imu.p, S, ev.r_gp.p, ev.J_gp.p);
    else k_dt_gp<false><<<cdiv(nG, 128), 128, 0, st>>>(p->K, nG, p->d_g_frame0.p, p->d_g_t.p, p->d_g_pos.p, p->d_g_inv_std.p, p->d_ts_imu.p, S, ev.r_gp.p, ev.J_gp.p);
    ++p->launches;
  }
  k_dt_cost<<<1, 1024, 0, st>>>(p->K.cauchy_a, nF, ev.f_sq.p, 15 * (size_t)nI, ev.r_imu.p, 3 * (size_t)nG, ev.r_gp.p, ev.cost.p);
  ++p->launches;
  DT_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}
int dt_check_imu(ctba_dt_problem* p) {
  if (!p->nI()) return CTBA_OK;
  int32_t rc = 0;
  DT_CUDA(p, cudaMemcpyAsync(&rc, p->d_imu_rc.p, sizeof(rc), cudaMemcpyDeviceToHost, p->stream));
  DT_CUDA(p, cudaStreamSynchronize(p->stream));
  if (rc == -1) { p->err = "IMU factor: the samples do not reach the second state's time"; return CTBA_E_ARG; }
  if (rc == -2) { p->err = "IMU factor: singular pre-integration covariance"; return CTBA_E_NUMERIC; }
  return CTBA_OK;
}

GAct dt_gact(const ctba_dt_problem* p) { GAct g; for (int t = 0; t < GT_SIZE; ++t) g.a[t] = p->gactive[t]; return g; }

// host analysis: landmark lists sorted by state, band width, active masks, buffers, the solver context
int dt_structure(ctba_dt_problem* p) {
  cudaStream_t st = p->stream;
  const size_t F = p->F(), M = p->M(), L = p->lm.size() / 3;
  const int nFm = (int)p->f_state.size();
  p->Lm = (int)L;
  std::vector<int32_t> o_state(M);
  for (int f = 0; f < nFm; ++f) for (int o = p->f_obs_begin[f]; o < p->f_obs_begin[f + 1]; ++o) o_state[o] = p->f_state[f];
  std::vector<int32_t> lm_begin(L + 1, 0), lm_obs(M);
  for (size_t o = 0; o < M; ++o) lm_begin[p->o_lm[o] + 1]++;
  for (size_t l = 0; l < L; ++l) lm_begin[l + 1] += lm_begin[l];
  {
    std::vector<int32_t> cur(lm_begin.begin(), lm_begin.end() - 1);
    for (size_t o = 0; o < M; ++o) lm_obs[cur[p->o_lm[o]]++] = (int32_t)o;
  }
  int span = 29;   // IMU / GP terms couple two consecutive states
  std::vector<uint8_t> lm_active(L, 0), sb_active(F, 0);
  for (size_t l = 0; l < L; ++l) {
    auto b = lm_obs.begin() + lm_begin[l], e = lm_obs.begin() + lm_begin[l + 1];
    if (b == e) continue;
    lm_active[l] = 1;
    std::stable_sort(b, e, [&](int32_t x, int32_t y) { return o_state[x] < o_state[y]; });
    for (auto it = b; it + 1 < e; ++it) if (o_state[*it] == o_state[*(it + 1)]) { p->err = "a landmark is observed twice in one state"; return CTBA_E_ARG; }
    span = std::max(span, 15 * (o_state[*(e - 1)] - o_state[*b]) + 5);
  }
  for (size_t k = 0; k < p->nI(); ++k) { sb_active[p->i_frame0[k]] = 1; sb_active[p->i_frame0[k] + 1] = 1; }
  // (state, distance) keys -> the landmark pairs that couple the two poses: counting sort of all (i >= j) pairs of every landmark list
  int T = 0;
  for (size_t l = 0; l < L; ++l) if (lm_begin[l + 1] > lm_begin[l]) T = std::max(T, o_state[lm_obs[lm_begin[l + 1] - 1]] - o_state[lm_obs[lm_begin[l]]]);
  p->T1 = T + 1;
  const size_t nKeys = F * (size_t)p->T1;
  std::vector<int64_t> key_begin(nKeys + 1, 0);
  for (size_t l = 0; l < L; ++l)
    for (int i = lm_begin[l]; i < lm_begin[l + 1]; ++i) {
      const int si = o_state[lm_obs[i]];
      for (int j = lm_begin[l]; j <= i; ++j) key_begin[(size_t)si * p->T1 + (si - o_state[lm_obs[j]]) + 1]++;
    }
  for (size_t k = 0; k < nKeys; ++k) key_begin[k + 1] += key_begin[k];
  std::vector<int2> pairs((size_t)key_begin[nKeys]);
  {
    std::vector<int64_t> cur(key_begin.begin(), key_begin.end() - 1);
    for (size_t l = 0; l < L; ++l)
      for (int i = lm_begin[l]; i < lm_begin[l + 1]; ++i) {
        const int oi = lm_obs[i], si = o_state[oi];
        for (int j = lm_begin[l]; j <= i; ++j) { const int oj = lm_obs[j]; pairs[(size_t)cur[(size_t)si * p->T1 + (si - o_state[oj])]++] = make_int2(oi, oj); }
      }
  }
  // measurement frames must map to distinct states (one writer per pose block and pass)
  {
    std::vector<uint8_t> seen(F, 0);
    for (int f = 0; f < nFm; ++f) { if (seen[p->f_state[f]]) { p->err = "two measurement frames at one state"; return CTBA_E_ARG; } seen[p->f_state[f]] = 1; }
  }
  // global-position samples grouped by their first state, in the order they were added
  std::vector<int32_t> run_begin(F + 1, 0), gp_order(p->nG());
  for (size_t i = 0; i < p->nG(); ++i) run_begin[p->g_frame0[i] + 1]++;
  for (size_t f = 0; f < F; ++f) run_begin[f + 1] += run_begin[f];
  {
    std::vector<int32_t> cur(run_begin.begin(), run_begin.end() - 1);
    for (size_t i = 0; i < p->nG(); ++i) gp_order[cur[p->g_frame0[i]]++] = (int32_t)i;
  }
  // free global blocks, as DiscreteTimeEstimator wires them (estimator.h:149-165, 373-519)
  const bool has_fr = M > 0, has_gp = p->nG() > 0;
  for (int t = 0; t < GT_SIZE; ++t) p->gactive[t] = 0;
  if ((has_fr || has_gp) && p->optimize_scale) p->gactive[DGT_SCALE] = 1;
  if (!(has_fr && !p->optimize_camimu)) for (int k = 0; k < 3; ++k) p->gactive[DGT_RBC + k] = 1;
  if (has_fr && p->optimize_camimu) { for (int k = 0; k < 3; ++k) p->gactive[DGT_PBC + k] = 1; p->gactive[DGT_TCAM] = 1; }
  if (has_gp && p->optimize_t_gs) p->gactive[DGT_TGS] = 1;
  if (has_gp && p->optimize_p_bp) for (int k = 0; k < 3; ++k) p->gactive[DGT_PBP + k] = 1;
  p->n_band = 15 * (int)F;
  p->n_pad = (p->n_band + kNB - 1) / kNB * kNB;
  p->nbw = std::min((span + kNB - 1) / kNB + 0, p->n_pad / kNB - 1);
  if (p->nbw < 1) p->nbw = std::min(1, p->n_pad / kNB - 1);
  DT_CUDA(p, p->d_lm_begin.upload(lm_begin, st)); DT_CUDA(p, p->d_lm_obs.upload(lm_obs, st));
  DT_CUDA(p, p->d_sb_active.upload(sb_active, st)); DT_CUDA(p, p->d_lm_active.upload(lm_active, st));
  DT_CUDA(p, p->d_key_begin.upload(key_begin, st)); DT_CUDA(p, p->d_pairs.upload(pairs, st));
  DT_CUDA(p, p->d_run_begin.upload(run_begin, st)); DT_CUDA(p, p->d_gp_order.upload(gp_order, st));
  DT_CUDA(p, cudaStreamSynchronize(st));
  const size_t nu = (size_t)p->n_pad + kBorder;
  DT_CUDA(p, p->f_w.resize(nFm)); DT_CUDA(p, p->grad.resize(nu)); DT_CUDA(p, p->diag.resize(nu)); DT_CUDA(p, p->jscale.resize(nu)); DT_CUDA(p, p->dvec.resize(nu));
  DT_CUDA(p, p->rhs_acc.resize(nu));
  DT_CUDA(p, p->V0.resize(6 * L)); DT_CUDA(p, p->gl.resize(3 * L)); DT_CUDA(p, p->jscale_lm.resize(3 * L)); DT_CUDA(p, p->Vi.resize(6 * L)); DT_CUDA(p, p->WgS.resize(24 * L));
  DT_CUDA(p, p->Vg.resize(3 * L)); DT_CUDA(p, p->Yp.resize((size_t)kWS * M)); DT_CUDA(p, p->cp_frame.resize((size_t)kCpFrame * nFm)); DT_CUDA(p, p->cp_lm.resize((size_t)kCpFrame * L));
  DT_CUDA(p, p->cp_gp.resize((size_t)kCpGp * F));
  DT_CUDA(p, p->Wp.resize((size_t)kWS * M)); DT_CUDA(p, p->delta_lm.resize(3 * L)); DT_CUDA(p, p->scal.resize(DSC_COUNT));
  DT_CUDA(p, p->cand.pose.resize(7 * F)); DT_CUDA(p, p->cand.sb.resize(9 * F)); DT_CUDA(p, p->cand.lm.resize(3 * L)); DT_CUDA(p, p->cand.glb.resize(DG_SIZE));
  if (p->WgS.p) DT_CUDA(p, cudaMemsetAsync(p->WgS.p, 0, 24 * L * 8, st));
  // band solver context
  ctba_problem* c = p->ctx;
  if (c->solver) { solver_free(c->solver); c->solver = nullptr; }
  SolverData* sd = new SolverData();
  c->solver = sd;
  BandLayout& B = sd->Hw;
  B.n_band = p->n_band; B.n_pad = p->n_pad; B.nbc = p->n_pad / kNB; B.nbw = p->nbw;
  DT_CUDA(p, sd->bandw.resize(B.band_elems())); DT_CUDA(p, sd->borderw.resize((size_t)kBorder * B.n_pad)); DT_CUDA(p, sd->cornerw.resize(kBorder * kBorder));
  B.band = sd->bandw.p; B.border = sd->borderw.p; B.corner = sd->cornerw.p;
  DT_CUDA(p, sd->band0.resize(B.band_elems())); DT_CUDA(p, sd->border0.resize((size_t)kBorder * B.n_pad)); DT_CUDA(p, sd->corner0.resize(kBorder * kBorder));
  sd->H0 = B; sd->H0.band = sd->band0.p; sd->H0.border = sd->border0.p; sd->H0.corner = sd->corner0.p;
  DT_CUDA(p, sd->delta.resize(nu)); DT_CUDA(p, sd->flag.resize(4));
  DT_CUDA(p, cudaMemsetAsync(sd->delta.p, 0, nu * 8, st));
  sd->built = true;
  p->struct_dirty = false;
  return CTBA_OK;
}

// direct normal equations H0 (IMU, global position, landmark-free part of the reprojections), gradient, diag(J^T J), landmark blocks
// from the evaluation buffers of the current state
int dt_build(ctba_dt_problem* p, DtEval& ev) {
  cudaStream_t st = p->stream;
  SolverData* sd = p->ctx->solver;
  BandLayout& H0 = sd->H0;
  const int nF = (int)p->f_state.size(), nI = (int)p->nI(), nG = (int)p->nG(), L = p->Lm, F = (int)p->F();
  const size_t nu = (size_t)p->n_pad + kBorder;
  const GAct ga = dt_gact(p);
  int any_global = 0;
  for (int t = 0; t < 8; ++t) any_global |= ga.a[t];
  DT_CUDA(p, cudaMemsetAsync(H0.band, 0, H0.band_elems() * 8, st)); DT_CUDA(p, cudaMemsetAsync(H0.border, 0, (size_t)kBorder * H0.n_pad * 8, st));
  DT_CUDA(p, cudaMemsetAsync(H0.corner, 0, kBorder * kBorder * 8, st)); DT_CUDA(p, cudaMemsetAsync(p->grad.p, 0, nu * 8, st));
  if (nF) {
    k_dt_frame_w<<<cdiv(nF, 128), 128, 0, st>>>(p->K.cauchy_a, nF, ev.f_sq.p, p->f_w.p);
    k_dt_direct_frame<<<nF, 128, 0, st>>>(H0, nF, p->d_f_state.p, p->d_f_obs_begin.p, p->f_w.p, ev.rec.p, ga, any_global, p->grad.p, p->cp_frame.p);
    p->launches += 2;
  }
  if (nI) for (int par = 0; par < 2; ++par) { k_dt_direct_imu<<<cdiv((size_t)nI * 32, 128), 128, 0, st>>>(H0, nI, par, p->d_i_frame0.p, ev.r_imu.p, ev.J_imu.p, p->grad.p); ++p->launches; }
  if (nG) for (int par = 0; par < 2; ++par) {
    k_dt_direct_gp<<<F - 1, 192, 0, st>>>(H0, F - 1, par, p->d_run_begin.p, p->d_gp_order.p, ev.r_gp.p, ev.J_gp.p, ga, p->grad.p, p->cp_gp.p);
    ++p->launches;
  }
  if (L) { k_dt_lm_blocks<<<cdiv(L, 128), 128, 0, st>>>(L, p->d_lm_begin.p, p->d_lm_obs.p, p->d_o_frame.p, p->f_w.p, ev.rec.p, p->V0.p, p->gl.p); ++p->launches; }
  k_dt_corner_reduce<<<90, 256, 0, st>>>(0, H0, (nF && any_global) ? nF : 0, (nF && any_global) ? p->cp_frame.p : nullptr, nG ? F - 1 : 0, nG ? p->cp_gp.p : nullptr, p->grad.p);
  k_dt_diag<<<cdiv(nu, 128), 128, 0, st>>>(H0, p->diag.p);
  p->launches += 2;
  DT_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}
int dt_assemble(ctba_dt_problem* p, DtEval& ev, double radius) {
  (void)ev;
  cudaStream_t st = p->stream;
  SolverData* sd = p->ctx->solver;
  BandLayout& B = sd->Hw; BandLayout& H0 = sd->H0;
  const int nF = (int)p->f_state.size(), L = p->Lm, F = (int)p->F();
  const size_t nu = (size_t)p->n_pad + kBorder;
  const GAct ga = dt_gact(p);
  int any_global = 0;
  for (int t = 0; t < 8; ++t) any_global |= ga.a[t];
  DT_CUDA(p, cudaMemcpyAsync(B.band, H0.band, B.band_elems() * 8, cudaMemcpyDeviceToDevice, st));
  DT_CUDA(p, cudaMemcpyAsync(B.border, H0.border, (size_t)kBorder * B.n_pad * 8, cudaMemcpyDeviceToDevice, st));
  DT_CUDA(p, cudaMemcpyAsync(B.corner, H0.corner, kBorder * kBorder * 8, cudaMemcpyDeviceToDevice, st));
  DT_CUDA(p, cudaMemsetAsync(p->rhs_acc.p, 0, nu * 8, st));
  DT_CUDA(p, cudaMemsetAsync(sd->flag.p, 0, 4 * sizeof(int32_t), st));
  if (nF && L) {
    k_dt_schur_prepare<<<cdiv((size_t)L * 32, 128), 128, 0, st>>>(L, radius, p->d_lm_begin.p, p->d_lm_obs.p, p->d_o_frame.p, p->f_w.p, p->ev.rec.p, p->V0.p, p->gl.p, p->jscale_lm.p, ga,
                                                                  any_global, p->Vi.p, p->Vg.p, p->WgS.p, p->Wp.p, p->Yp.p, p->cp_lm.p, sd->flag.p + 2);
    const size_t nKeys = (size_t)F * p->T1;
    k_dt_schur_gather<<<cdiv(nKeys * 32, 128), 128, 0, st>>>(B, (int)nKeys, p->T1, p->d_key_begin.p, p->d_pairs.p, p->Wp.p, p->Yp.p);
    k_dt_schur_frame<<<nF, 128, 0, st>>>(B, nF, p->d_f_state.p, p->d_f_obs_begin.p, p->d_o_lm.p, p->Wp.p, p->Yp.p, p->WgS.p, p->Vg.p, ga, any_global, p->rhs_acc.p);
    p->launches += 3;
    if (any_global) { k_dt_corner_reduce<<<90, 256, 0, st>>>(1, B, L, p->cp_lm.p, 0, nullptr, p->rhs_acc.p); ++p->launches; }
  }
  k_dt_finish<<<cdiv(nu, 128), 128, 0, st>>>(B, p->n_band, p->d_sb_active.p, p->jscale.p, p->diag.p, p->grad.p, p->rhs_acc.p, radius, ga, p->dvec.p);
  ++p->launches;
  DT_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}
int dt_scalars(ctba_dt_problem* p, int mode, const DtStateP& S, const DtStateP& C, double radius, double* host_sc) {
  cudaStream_t st = p->stream;
  SolverData* sd = p->ctx->solver;
  const int F = (int)p->F(), L = p->Lm;
  const int nblk = cdiv((size_t)F + L + 1, 256);
  DT_CUDA(p, p->scal_part.resize((size_t)nblk * 8 * 4));
  k_dt_step<<<nblk, 256, 0, st>>>(mode, F, L, p->n_pad, p->d_sb_active.p, p->d_lm_active.p, dt_gact(p), S, C, sd->delta.p, p->delta_lm.p, p->grad.p, p->gl.p,
                                  p->dvec.p, p->V0.p, p->jscale_lm.p, radius, p->scal_part.p);
  k_dt_scal_reduce<<<1, 256, 0, st>>>(nblk * 8, p->scal_part.p, p->scal.p);
  p->launches += 2;
  DT_CUDA(p, cudaMemcpyAsync(host_sc, p->scal.p, DSC_COUNT * 8, cudaMemcpyDeviceToHost, st));
  DT_CUDA(p, cudaStreamSynchronize(st));
  return CTBA_OK;
}

struct EvT {
  cudaEvent_t a = nullptr, b = nullptr; cudaStream_t st; double* acc;
  EvT(cudaStream_t s, double* t) : st(s), acc(t) { cudaEventCreate(&a); cudaEventCreate(&b); cudaEventRecord(a, st); }
  void stop() { cudaEventRecord(b, st); cudaEventSynchronize(b); float ms = 0; cudaEventElapsedTime(&ms, a, b); *acc += ms * 1e-3; }
  ~EvT() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
};

int dt_solve(ctba_dt_problem* p, int max_num_iterations, ctba_summary* out) {
  // ceres::Solver::Options defaults [Ceres, recalled] — the same constants and decision order as ctba_solve.cu
  const double initial_radius = 1e4, max_radius = 1e16, min_radius = 1e-32, min_relative_decrease = 1e-3;
  const double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  const int max_consecutive_invalid = 5;
  auto t_begin = std::chrono::steady_clock::now();
  const int64_t launches0 = p->launches + p->ctx->launches;
  out->time_eval_s = out->time_build_s = out->time_solve_s = 0.0;
  int rc = dt_sync(p); if (rc) return rc;
  if (p->struct_dirty) { rc = dt_structure(p); if (rc) return rc; }
  cudaStream_t st = p->stream;
  SolverData* sd = p->ctx->solver;
  DtImuCache* cache = (DtImuCache*)p->d_imu_cache.p;
  double host_cost[4], host_sc[DSC_COUNT];
  int32_t host_flag[4];
  int nit = 0;
  auto log_it = [&](double cost, double cc, double gmax, double sn, double ratio, double radius, int ok) {
    if (out->iterations && nit < out->iterations_capacity) out->iterations[nit] = {cost, cc, gmax, sn, ratio, radius, ok};
    ++nit;
  };
  auto eval_at = [&](DtEval& ev, const DtStateP& S, double* cost) -> int {
    EvT t(st, &out->time_eval_s);
    int r = dt_evaluate(p, ev, S, true, cache); if (r) return r;
    DT_CUDA(p, cudaMemcpyAsync(host_cost, ev.cost.p, sizeof(host_cost), cudaMemcpyDeviceToHost, st));
    t.stop();
    r = dt_check_imu(p); if (r) return r;
    *cost = host_cost[3];
    return CTBA_OK;
  };
  auto build_at = [&]() -> int { EvT t(st, &out->time_build_s); int r = dt_build(p, p->ev); t.stop(); return r; };
  double x_cost = 0.0;
  rc = eval_at(p->ev, ptrs(p->st), &x_cost); if (rc) return rc;
  rc = build_at(); if (rc) return rc;
  k_dt_jscale<<<cdiv(std::max((size_t)p->n_pad + kBorder, 3 * (size_t)p->Lm), 128), 128, 0, st>>>(p->n_pad + kBorder, p->diag.p, p->Lm, p->V0.p, p->jscale.p, p->jscale_lm.p);
  ++p->launches;
  rc = dt_scalars(p, 1, ptrs(p->st), ptrs(p->st), -1.0, host_sc); if (rc) return rc;
  double x_norm = std::sqrt(host_sc[DSC_XNORM2]), gmax = host_sc[DSC_GMAX];
  out->initial_cost = x_cost;
  out->num_successful_steps = out->num_unsuccessful_steps = 0;
  out->termination_type = 1;
  log_it(x_cost, 0.0, gmax, 0.0, 0.0, initial_radius, 1);
  double radius = initial_radius, decrease_factor = 2.0;
  int invalid = 0, iter = 0;
  bool done = false;
  if (!std::isfinite(x_cost)) { out->termination_type = 2; done = true; }
  else if (!p->bench_no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; done = true; }
  while (!done) {
    if (iter >= max_num_iterations) { out->termination_type = 1; break; }
    if (radius < min_radius) { if (p->bench_no_termination) radius = initial_radius; else { out->termination_type = 0; break; } }
    ++iter;
    { EvT tb(st, &out->time_build_s); rc = dt_assemble(p, p->ev, radius); tb.stop(); if (rc) return rc; }
    int32_t lm_flag = 0;
    {
      EvT ts(st, &out->time_solve_s);
      DT_CUDA(p, cudaMemcpyAsync(&lm_flag, sd->flag.p + 2, sizeof(int32_t), cudaMemcpyDeviceToHost, st));   // band_factor_solve clears the flags
      rc = band_factor_solve(p->ctx);
      if (rc) { p->err = p->ctx->err; return rc; }
      DT_CUDA(p, cudaMemcpyAsync(host_flag, sd->flag.p, sizeof(host_flag), cudaMemcpyDeviceToHost, st));
      ts.stop();
    }
    bool ok = host_flag[0] == 0 && lm_flag == 0;
    double model_cost_change = 0.0, step_norm = 0.0;
    if (ok) {
      const GAct ga = dt_gact(p);
      int any_global = 0;
      for (int t = 0; t < 8; ++t) any_global |= ga.a[t];
      if (p->Lm) {
        k_dt_backsub<<<cdiv(p->Lm, 128), 128, 0, st>>>(p->Lm, p->d_lm_begin.p, p->d_lm_obs.p, p->d_o_frame.p, p->d_f_state.p, p->Vi.p, p->WgS.p, p->Wp.p, p->gl.p, sd->delta.p, p->n_pad,
                                                     any_global, p->delta_lm.p);
        ++p->launches;
      }
      rc = dt_scalars(p, 0, ptrs(p->st), ptrs(p->cand), radius, host_sc); if (rc) return rc;
      model_cost_change = host_sc[DSC_MODEL];
      step_norm = std::sqrt(host_sc[DSC_STEP2]);
      if (!std::isfinite(model_cost_change) || !std::isfinite(step_norm)) ok = false;
    }
    if (!ok || model_cost_change <= 0.0) {
      if (getenv("CTBA_DEBUG")) fprintf(stderr, "[ctba_dt %p] iter %d invalid step: flag %d lm %d model %.6g step %.6g radius %.3g\n", (void*)p, iter, host_flag[0], lm_flag, model_cost_change, step_norm, radius);
      if (++invalid >= max_consecutive_invalid && !p->bench_no_termination) { out->termination_type = 2; break; }
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, 0.0, gmax, 0.0, 0.0, radius, 0);
      continue;
    }
    invalid = 0;
    double cand_cost = 0.0;
    rc = eval_at(p->ev_alt, ptrs(p->cand), &cand_cost); if (rc) return rc;   // residuals and Jacobians at the candidate; swapped in when accepted
    const double cost_change = x_cost - cand_cost;
    const double ratio = cost_change / model_cost_change;
    if (!p->bench_no_termination && step_norm <= parameter_tolerance * (x_norm + parameter_tolerance)) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (!p->bench_no_termination && std::fabs(cost_change) <= function_tolerance * x_cost) {
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
      out->termination_type = 0; break;
    }
    if (std::isfinite(cand_cost) && (ratio > min_relative_decrease || p->bench_accept_all)) {
      p->st.swap(p->cand);
      p->ev.swap(p->ev_alt);
      p->host_stale = true;
      x_cost = cand_cost;
      rc = build_at(); if (rc) return rc;
      rc = dt_scalars(p, 1, ptrs(p->st), ptrs(p->st), -1.0, host_sc); if (rc) return rc;
      x_norm = std::sqrt(host_sc[DSC_XNORM2]); gmax = host_sc[DSC_GMAX];
      if (ratio > min_relative_decrease) { radius = std::min(max_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3))); decrease_factor = 2.0; }
      else { radius /= decrease_factor; decrease_factor *= 2.0; }
      out->num_successful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 1);
      if (!p->bench_no_termination && gmax <= gradient_tolerance) { out->termination_type = 0; break; }
    } else {
      radius /= decrease_factor; decrease_factor *= 2.0;
      out->num_unsuccessful_steps++;
      log_it(x_cost, cost_change, gmax, step_norm, ratio, radius, 0);
    }
  }
  out->final_cost = x_cost;
  out->num_iterations = nit;
  out->kernel_launches = p->launches + p->ctx->launches - launches0;
  DT_CUDA(p, cudaStreamSynchronize(st));
  out->time_total_s = std::chrono::duration<double>(std::chrono::steady_clock::now() - t_begin).count();
  return CTBA_OK;
}

}  // namespace

// ================= C ABI =================
extern "C" {

int ctba_dt_create(int device, ctba_dt_problem** out) {
  if (!out) return CTBA_E_ARG;
  *out = nullptr;
  ctba_problem* ctx = nullptr;
  const int rc = ctba_create(6, 1, 0, device, &ctx);   // validates the device, creates the stream
  if (rc) return rc;
  if (device < 0) { ctba_destroy(ctx); return CTBA_E_CUDA; }
  ctba_dt_problem* p = new ctba_dt_problem();
  p->device = device; p->ctx = ctx; p->stream = ctx->stream;
  p->K.cauchy_a = 3.0;   // CauchyLoss(3.0), estimator.h:432
  p->K.g_max = 7.8; p->K.a_max = 150.0;
  p->K.g_W[0] = 0; p->K.g_W[1] = 0; p->K.g_W[2] = 9.80665;
  *out = p;
  return CTBA_OK;
}
int ctba_dt_destroy(ctba_dt_problem* p) {
  DT_CHECK(p);
  cudaSetDevice(p->device);
  cudaStreamSynchronize(p->stream);
  ctba_problem* ctx = p->ctx;
  delete p;              // device buffers (Dev<> members)
  return ctba_destroy(ctx);
}
const char* ctba_dt_last_error(const ctba_dt_problem* p) { return p ? p->err.c_str() : "null handle"; }
int ctba_dt_set_stream(ctba_dt_problem* p, void* cuda_stream) {
  DT_CHECK(p);
  cudaStreamSynchronize(p->stream);
  const int rc = ctba_set_stream(p->ctx, cuda_stream);
  if (rc) { p->err = p->ctx->err; return rc; }
  p->stream = p->ctx->stream;
  return CTBA_OK;
}
// the host arrays are authoritative again: the next compute call uploads state and measurements (end-to-end timing through host buffers)
int ctba_dt_invalidate_device(ctba_dt_problem* p) {
  DT_CHECK(p);
  const int rc = dt_download(p); if (rc) return rc;
  p->meas_dirty = p->state_dirty = true;
  return CTBA_OK;
}

int ctba_dt_set_camera(ctba_dt_problem* p, int model, const double* intr, const double* dist, const double* q_bc, const double* p_bc, double t_off, double width,
                       double height) {
  DT_CHECK(p);
  if (!intr || !dist || !q_bc || !p_bc || (model != 4 && model != 5 && model != 7 && model != 1)) { p->err = "set_camera: bad argument"; return CTBA_E_ARG; }
  p->K.cam.model = model; p->K.cam.fx = intr[0]; p->K.cam.fy = intr[1]; p->K.cam.cx = intr[2]; p->K.cam.cy = intr[3];
  for (int i = 0; i < 4; ++i) { p->K.cam.d[i] = dist[i]; p->K.q_bc0[i] = q_bc[i]; }
  for (int i = 0; i < 3; ++i) p->K.p_bc0[i] = p_bc[i];
  p->K.t_offset_cam_imu = t_off; p->K.width = width; p->K.height = height;
  p->have_camera = true;
  return CTBA_OK;
}
int ctba_dt_set_imu(ctba_dt_problem* p, double gnd, double and_, double grw, double arw, double g_max, double a_max, const double* g_W) {
  DT_CHECK(p);
  if (!g_W) return CTBA_E_ARG;
  p->K.gyro_noise_density = gnd; p->K.acc_noise_density = and_; p->K.gyro_random_walk = grw; p->K.acc_random_walk = arw; p->K.g_max = g_max; p->K.a_max = a_max;
  for (int i = 0; i < 3; ++i) p->K.g_W[i] = g_W[i];
  p->have_imu = true;
  return CTBA_OK;
}
int ctba_dt_initialize_states(ctba_dt_problem* p, int F, const double* poses, const double* ts_cam) {
  DT_CHECK(p);
  if (F < 2 || !poses || !ts_cam) { p->err = "initialize_states: at least two states"; return CTBA_E_ARG; }
  if (!p->have_camera) { p->err = "initialize_states: set_camera first (the states' IMU times use the camera-imu offset)"; return CTBA_E_STATE; }
  p->pose.assign(poses, poses + 7 * (size_t)F); p->ts_cam.assign(ts_cam, ts_cam + F); p->ts_imu.resize(F); p->sb.assign(9 * (size_t)F, 0.0);
  for (int i = 0; i < F; ++i) p->ts_imu[i] = ts_cam[i] + p->K.t_offset_cam_imu;
  p->state_dirty = p->struct_dirty = p->meas_dirty = true; p->host_stale = false;
  return CTBA_OK;
}
int ctba_dt_initialize_linear_velocity(ctba_dt_problem* p) {
  DT_CHECK(p);
  int rc = dt_download(p); if (rc) return rc;
  const size_t n = p->F();
  if (n < 2) { p->err = "initialize_linear_velocity: initialize_states first"; return CTBA_E_STATE; }
  for (size_t i = 0; i + 1 < n; ++i) {
    const double dt = p->ts_imu[i + 1] - p->ts_imu[i];
    for (int c = 0; c < 3; ++c) p->sb[9 * i + c] = (p->pose[7 * (i + 1) + 4 + c] - p->pose[7 * i + 4 + c]) / dt;
  }
  for (int c = 0; c < 3; ++c) p->sb[9 * (n - 1) + c] = p->sb[9 * (n - 2) + c];
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_set_speed_and_bias(ctba_dt_problem* p, const double* sb) {
  DT_CHECK(p);
  if (!sb) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  std::memcpy(p->sb.data(), sb, p->sb.size() * 8);
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_set_states(ctba_dt_problem* p, const double* poses, const double* sb) {
  DT_CHECK(p);
  if (!poses || !sb) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  std::memcpy(p->pose.data(), poses, p->pose.size() * 8); std::memcpy(p->sb.data(), sb, p->sb.size() * 8);
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_get_states(ctba_dt_problem* p, double* poses, double* sb) {
  DT_CHECK(p);
  int rc = dt_download(p); if (rc) return rc;
  if (poses) std::memcpy(poses, p->pose.data(), p->pose.size() * 8);
  if (sb) std::memcpy(sb, p->sb.data(), p->sb.size() * 8);
  return CTBA_OK;
}
int ctba_dt_set_landmarks(ctba_dt_problem* p, int L, const double* xyz) {
  DT_CHECK(p);
  if (L < 0 || (L && !xyz)) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  if ((size_t)L != p->lm.size() / 3) p->struct_dirty = true;
  p->lm.assign(xyz, xyz + 3 * (size_t)L);
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_get_landmarks(ctba_dt_problem* p, double* xyz) {
  DT_CHECK(p);
  int rc = dt_download(p); if (rc) return rc;
  if (xyz && !p->lm.empty()) std::memcpy(xyz, p->lm.data(), p->lm.size() * 8);
  return CTBA_OK;
}
int ctba_dt_set_globals(ctba_dt_problem* p, double scale, const double* p_bp, double t_gs) {
  DT_CHECK(p);
  if (!p_bp) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  p->K.scale = scale; p->K.t_offset_gs_imu = t_gs;
  for (int i = 0; i < 3; ++i) p->glb[DG_PBP + i] = p_bp[i];
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_set_corrections(ctba_dt_problem* p, const double* c9, double t_gs_corr) {
  DT_CHECK(p);
  if (!c9) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  for (int i = 0; i < 9; ++i) p->glb[i] = c9[i];
  p->glb[DG_TGS] = t_gs_corr;
  p->state_dirty = true;
  return CTBA_OK;
}
int ctba_dt_get_globals(ctba_dt_problem* p, double* o) {
  DT_CHECK(p);
  if (!o) return CTBA_E_ARG;
  int rc = dt_download(p); if (rc) return rc;
  for (int i = 0; i < DG_SIZE; ++i) o[i] = p->glb[i];
  return CTBA_OK;
}
int ctba_dt_set_flags(ctba_dt_problem* p, int scale, int pbp, int camimu, int tgs) {
  DT_CHECK(p);
  p->optimize_scale = scale; p->optimize_p_bp = pbp; p->optimize_camimu = camimu; p->optimize_t_gs = tgs;
  p->struct_dirty = true;
  return CTBA_OK;
}
int ctba_dt_add_frames(ctba_dt_problem* p, int nF, const int32_t* state_id, const int32_t* obs_begin, const double* uv, const double* vel, const int32_t* lm, double sigma) {
  DT_CHECK(p);
  if (nF < 0 || !state_id || !obs_begin || !(sigma > 0.0)) { p->err = "add_frames: bad argument"; return CTBA_E_ARG; }
  const int L = (int)(p->lm.size() / 3), F = (int)p->F();
  for (int f = 0; f < nF; ++f) {
    if (state_id[f] < 0 || state_id[f] >= F || obs_begin[f + 1] < obs_begin[f]) { p->err = "add_frames: state id out of range"; return CTBA_E_ARG; }
    for (int q = obs_begin[f]; q < obs_begin[f + 1]; ++q) if (lm[q] < 0 || lm[q] >= L) { p->err = "add_frames: landmark index out of range (set_landmarks first)"; return CTBA_E_ARG; }
  }
  int added = 0;
  for (int f = 0; f < nF; ++f) {
    if (obs_begin[f + 1] == obs_begin[f]) continue;   // addImagePointsMeasurement returns early without valid points (estimator.h:377)
    p->f_state.push_back(state_id[f]); p->f_inv_std.push_back(1.0 / sigma);
    for (int q = obs_begin[f]; q < obs_begin[f + 1]; ++q) {
      p->o_uv.push_back(uv[2 * q]); p->o_uv.push_back(uv[2 * q + 1]); p->o_vel.push_back(vel ? vel[2 * q] : 0.0); p->o_vel.push_back(vel ? vel[2 * q + 1] : 0.0);
      p->o_lm.push_back(lm[q]);
    }
    p->f_obs_begin.push_back((int32_t)p->o_lm.size());
    ++added;
  }
  p->meas_dirty = p->struct_dirty = true;
  return added;
}
int ctba_dt_add_imu(ctba_dt_problem* p, int frame0, int n, const double* t, const double* gyr, const double* acc) {
  DT_CHECK(p);
  if (frame0 < 0 || frame0 + 1 >= (int)p->F() || n < 2 || !t || !gyr || !acc) { p->err = "add_imu: bad argument"; return CTBA_E_ARG; }
  if (!(t[n - 1] >= p->ts_imu[frame0 + 1])) { p->err = "add_imu: the samples do not reach the second state's time"; return CTBA_E_ARG; }
  p->i_frame0.push_back(frame0);
  p->i_t.insert(p->i_t.end(), t, t + n); p->i_gyr.insert(p->i_gyr.end(), gyr, gyr + 3 * (size_t)n); p->i_acc.insert(p->i_acc.end(), acc, acc + 3 * (size_t)n);
  p->i_begin.push_back((int32_t)p->i_t.size());
  p->meas_dirty = p->struct_dirty = true;
  return CTBA_OK;
}
int ctba_dt_add_gp(ctba_dt_problem* p, int n, const double* t, const double* pos, const int32_t* frame0, double sigma) {
  DT_CHECK(p);
  if (n < 0 || !t || !pos || !frame0 || !(sigma > 0.0)) { p->err = "add_gp: bad argument"; return CTBA_E_ARG; }
  for (int i = 0; i < n; ++i) if (frame0[i] < 0 || frame0[i] + 1 >= (int)p->F()) { p->err = "add_gp: frame id + 1 beyond the states"; return CTBA_E_ARG; }   // estimator.h:470-473
  for (int i = 0; i < n; ++i) {
    p->g_t.push_back(t[i]); p->g_frame0.push_back(frame0[i]); p->g_inv_std.push_back(1.0 / sigma);
    for (int c = 0; c < 3; ++c) p->g_pos.push_back(pos[3 * i + c]);
  }
  p->meas_dirty = p->struct_dirty = true;
  return n;
}
int ctba_dt_compute_valid_reprojections(ctba_dt_problem* p, int F, const int32_t* pt_begin, const double* uv, const int32_t* lm_index, int width, int height,
                                          int cell_size_px, uint8_t* valid, int32_t* n_valid) {
  DT_CHECK(p);
  if (F < 0 || !pt_begin || !valid || !n_valid || width <= 0 || height <= 0 || cell_size_px <= 1) { p->err = "compute_valid_reprojections: bad arguments"; return CTBA_E_ARG; }
  if ((size_t)F > p->F()) { p->err = "compute_valid_reprojections: more frames than states (frame k belongs to state k, estimator.h:192-194)"; return CTBA_E_ARG; }
  int rc = dt_download(p); if (rc) return rc;
  const m3 R_bc = quat_matrix(quat{p->K.q_bc0[0], p->K.q_bc0[1], p->K.q_bc0[2], p->K.q_bc0[3]});
  const v3 p_bc{p->K.p_bc0[0], p->K.p_bc0[1], p->K.p_bc0[2]};
  const double scale = p->K.scale + p->glb[DG_SCALE];   // getScale(), estimator.h:744
  auto pose = [&](int f, m3* R_wb, v3* p_wb) -> bool {   // frame k is judged at state k's pose (estimator.h:243-250)
    const double* T = &p->pose[7 * (size_t)f];
    *R_wb = quat_matrix(quat{T[0], T[1], T[2], T[3]});
    *p_wb = v3{T[4], T[5], T[6]};
    return true;
  };
  return select_valid_reprojections(F, pt_begin, uv, lm_index, (int64_t)p->lm.size() / 3, p->lm.data(), width, height, cell_size_px, R_bc, p_bc, scale, pose, valid, n_valid,
                                    &p->err);
}
int ctba_dt_counts(ctba_dt_problem* p, int64_t* c) {
  DT_CHECK(p);
  if (!c) return CTBA_E_ARG;
  c[0] = (int64_t)p->f_state.size(); c[1] = (int64_t)p->M(); c[2] = (int64_t)p->nI(); c[3] = (int64_t)p->nG();
  std::vector<uint8_t> sba(p->F(), 0), lma(p->lm.size() / 3, 0);
  for (size_t k = 0; k < p->nI(); ++k) { sba[p->i_frame0[k]] = 1; sba[p->i_frame0[k] + 1] = 1; }
  for (int32_t l : p->o_lm) lma[l] = 1;
  int64_t cols = 6 * (int64_t)p->F();
  for (uint8_t a : sba) cols += 9 * a;
  for (uint8_t a : lma) cols += 3 * a;
  const bool has_fr = p->M() > 0, has_gp = p->nG() > 0;
  if ((has_fr || has_gp) && p->optimize_scale) cols += 1;
  if (!(has_fr && !p->optimize_camimu)) cols += 3;
  if (has_fr && p->optimize_camimu) cols += 4;
  if (has_gp && p->optimize_t_gs) cols += 1;
  if (has_gp && p->optimize_p_bp) cols += 3;
  c[4] = cols;
  return CTBA_OK;
}

int ctba_dt_evaluate(ctba_dt_problem* p, int kind, double* r, double* j) {
  DT_CHECK(p);
  if (kind < 0 || kind > 2) return CTBA_E_ARG;
  cudaSetDevice(p->device);
  int rc = dt_sync(p); if (rc) return rc;
  rc = dt_evaluate(p, p->ev, ptrs(p->st), true, (DtImuCache*)p->d_imu_cache.p); if (rc) return rc;
  rc = dt_check_imu(p); if (rc) return rc;
  cudaStream_t st = p->stream;
  if (kind == CTBA_DT_REPROJ) {
    const size_t M = p->M();
    std::vector<double> rec(36 * M);
    if (M) DT_CUDA(p, cudaMemcpyAsync(rec.data(), p->ev.rec.p, rec.size() * 8, cudaMemcpyDeviceToHost, st));
    DT_CUDA(p, cudaStreamSynchronize(st));
    for (size_t o = 0; o < M; ++o) {
      if (r) { r[2 * o] = rec[36 * o]; r[2 * o + 1] = rec[36 * o + 1]; }
      if (j) std::memcpy(j + 34 * o, &rec[36 * o + 2], 34 * 8);
    }
  } else if (kind == CTBA_DT_IMU) {
    if (r && p->nI()) DT_CUDA(p, cudaMemcpyAsync(r, p->ev.r_imu.p, 15 * p->nI() * 8, cudaMemcpyDeviceToHost, st));
    if (j && p->nI()) DT_CUDA(p, cudaMemcpyAsync(j, p->ev.J_imu.p, 450 * p->nI() * 8, cudaMemcpyDeviceToHost, st));
    DT_CUDA(p, cudaStreamSynchronize(st));
  } else {
    if (r && p->nG()) DT_CUDA(p, cudaMemcpyAsync(r, p->ev.r_gp.p, 3 * p->nG() * 8, cudaMemcpyDeviceToHost, st));
    if (j && p->nG()) DT_CUDA(p, cudaMemcpyAsync(j, p->ev.J_gp.p, 51 * p->nG() * 8, cudaMemcpyDeviceToHost, st));
    DT_CUDA(p, cudaStreamSynchronize(st));
  }
  return CTBA_OK;
}
int ctba_dt_cost(ctba_dt_problem* p, double* cost4) {
  DT_CHECK(p);
  if (!cost4) return CTBA_E_ARG;
  cudaSetDevice(p->device);
  int rc = dt_sync(p); if (rc) return rc;
  rc = dt_evaluate(p, p->ev, ptrs(p->st), false, (DtImuCache*)p->d_imu_cache.p); if (rc) return rc;
  rc = dt_check_imu(p); if (rc) return rc;
  DT_CUDA(p, cudaMemcpyAsync(cost4, p->ev.cost.p, 4 * 8, cudaMemcpyDeviceToHost, p->stream));
  DT_CUDA(p, cudaStreamSynchronize(p->stream));
  return CTBA_OK;
}
int ctba_dt_evaluate_errors(ctba_dt_problem* p, int kind, double* out) {
  DT_CHECK(p);
  if (kind < 0 || kind > 2 || !out) return CTBA_E_ARG;
  cudaSetDevice(p->device);
  int rc = dt_sync(p); if (rc) return rc;
  cudaStream_t st = p->stream;
  // fresh pre-integration caches: evaluateImuErrors builds new functors (estimator.h:593-625), the problem's factors are not touched
  Dev<char> tmp;
  const size_t nI = p->nI();
  if (nI) {
    DT_CUDA(p, tmp.resize(nI * sizeof(DtImuCache)));
    std::vector<DtImuCache> fresh(nI);
    for (auto& c : fresh) { std::memset(&c, 0, sizeof(c)); c.redo = 1; c.Delta_q[3] = 1.0; }
    DT_CUDA(p, cudaMemcpyAsync(tmp.p, fresh.data(), nI * sizeof(DtImuCache), cudaMemcpyHostToDevice, st));
    DT_CUDA(p, cudaStreamSynchronize(st));
  }
  rc = dt_evaluate(p, p->ev_alt, ptrs(p->st), false, (DtImuCache*)tmp.p); if (rc) return rc;
  rc = dt_check_imu(p); if (rc) return rc;
  if (kind == CTBA_DT_IMU) {
    if (nI) DT_CUDA(p, cudaMemcpyAsync(out, p->ev_alt.r_imu.p, 15 * nI * 8, cudaMemcpyDeviceToHost, st));
  } else {
    const size_t n = kind == CTBA_DT_REPROJ ? p->M() : p->nG();
    Dev<double> d;
    DT_CUDA(p, d.resize(n));
    if (n) {
      if (kind == CTBA_DT_REPROJ) k_dt_norms<<<cdiv(n, 128), 128, 0, st>>>(n, 2, p->ev_alt.rec.p, 36, p->d_f_inv_std.p, p->d_o_frame.p, d.p);
      else k_dt_norms<<<cdiv(n, 128), 128, 0, st>>>(n, 3, p->ev_alt.r_gp.p, 3, p->d_g_inv_std.p, nullptr, d.p);
      ++p->launches;
      DT_CUDA(p, cudaMemcpyAsync(out, d.p, n * 8, cudaMemcpyDeviceToHost, st));
    }
  }
  DT_CUDA(p, cudaStreamSynchronize(st));
  return CTBA_OK;
}
int ctba_dt_imu_redo_count(ctba_dt_problem* p, int k, int* count) {
  DT_CHECK(p);
  if (k < 0 || (size_t)k >= p->n_cache || !count) return CTBA_E_ARG;
  DtImuCache c;
  DT_CUDA(p, cudaMemcpy(&c, p->d_imu_cache.p + (size_t)k * sizeof(DtImuCache), sizeof(c), cudaMemcpyDeviceToHost));
  *count = c.redo_counter;
  return CTBA_OK;
}
int ctba_dt_solve(ctba_dt_problem* p, int max_num_iterations, ctba_summary* summary) {
  DT_CHECK(p);
  if (!summary || max_num_iterations < 0) return CTBA_E_ARG;
  cudaSetDevice(p->device);
  return dt_solve(p, max_num_iterations, summary);
}
int ctba_dt_set_option(ctba_dt_problem* p, const char* name, double value) {
  DT_CHECK(p);
  if (!name) return CTBA_E_ARG;
  const std::string n(name);
  if (n == "bench_no_termination") p->bench_no_termination = value != 0.0;
  else if (n == "bench_accept_all") p->bench_accept_all = value != 0.0;
  else if (n == "cauchy_a") p->K.cauchy_a = value;
  else if (n == "band_partitions" || n == "band_graph" || n == "band_nested") return ctba_set_option(p->ctx, name, value);
  else { p->err = "unknown option " + n; return CTBA_E_ARG; }
  return CTBA_OK;
}
int ctba_dt_get_stat(ctba_dt_problem* p, const char* name, double* value) {
  DT_CHECK(p);
  if (!name || !value) return CTBA_E_ARG;
  const std::string n(name);
  if (n == "unknowns_band") { *value = p->n_band; return CTBA_OK; }
  if (!p->ctx->solver) { p->err = "get_stat: no solve yet"; return CTBA_E_STATE; }
  return ctba_get_stat(p->ctx, name, value);
}

}  // extern "C"
