// ctba_band.cu — bordered block-banded Cholesky factorisation + substitutions on the device (sm_100a, FP64).
//
// Solves the reduced (landmark-eliminated) LM system  [H_band  Bd^T; Bd  G] [x; z] = [y; c]  of ctba_solver.h exactly.
// This is the linear solve that Ceres' SPARSE_NORMAL_CHOLESKY performs inside ceres::Solve for the reference
// (include/spline/bspline_so3r3.h:1771, 1778) — here on the block band that the spline's locality produces.
//
//   k_band_panel / k_band_update   one launch pair per 32-column block, right-looking with a one-column look-ahead; every 32x32x32
//                     product runs on the FP64 tensor pipe (mma.sync m8n8k4 f64 -> DMMA)
//   k_band_border     the border rows of a time partition as a trailing triangular solve with many right-hand sides (left-looking K loop)
//   k_corner_partial / k_corner_solve   Schur complement of the band onto the <= 15 global parameters, 15x15 solve
//   k_back_rhs + k_band_backsolve       backward substitution  L^T x = y - Lb^T z
#include <cstdlib>
#include <cooperative_groups.h>
#include <algorithm>
#include <cmath>
#include <initializer_list>
#include <vector>
#include "ctba_solver.h"

namespace ctba {
namespace cg = cooperative_groups;

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

// ---- FP64 tensor-core tile product -------------------------------------------------------------------------------
// mma.sync m8n8k4 f64 (SASS: DMMA).  Fragment ownership (lane = 4*g + t): A[g][t], B[t][g], C[g][2t], C[g][2t+1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// Operand fragments for K = 32: lane (g, t) holds row g, columns 8t .. 8t+7 — 64 contiguous bytes, two 256-bit loads — and k-step ks
// multiplies column 8t + ks.  A and B use the same column permutation, so the sum over k is complete; compared with the natural
// layout (column 4 ks + t: eight 8-byte loads per lane, each touching 8 cache lines per warp) the L1 sees a quarter of the requests,
// and L1 request throughput is what bounds these kernels (ncu: l1tex 87-99 % busy before the change).
__device__ __forceinline__ void load_frag8(const double* __restrict__ p, double v[8]) {
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[0]), "=d"(v[1]), "=d"(v[2]), "=d"(v[3]) : "l"(p));
  asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(v[4]), "=d"(v[5]), "=d"(v[6]), "=d"(v[7]) : "l"(p + 4));
}
// One 8x8 tile of  P = A * Bm^T  over K = 32:  A rows [8*tr, 8*tr+8) (row stride lda), Bm rows [8*tc, 8*tc+8) (row stride ldb).
// Operands are read straight from global memory / L2.
__device__ __forceinline__ void tile_abt(const double* __restrict__ A, int lda, const double* __restrict__ Bm, int ldb, int tr, int tc, int lane,
                                         double& c0, double& c1) {
  const int g = lane >> 2, t = lane & 3;
  double a[8], b[8];
  load_frag8(A + (size_t)(8 * tr + g) * lda + 8 * t, a);
  load_frag8(Bm + (size_t)(8 * tc + g) * ldb + 8 * t, b);
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) dmma884(c0, c1, a[ks], b[ks]);
}

constexpr int kLd = kNB + 1;       // shared-memory leading dimension: odd, so that lane = row accesses are conflict-free
constexpr int kStepThreads = 256;  // 8 warps: warp w owns tiles (tr = w>>1, tc = 2*(w&1) + {0,1}) of a 32x32 block

// Panel staging for one CTA (all global loads are issued before the first DMMA, so they share one L2 round trip):
//   sD = Cd - Ld Ld^T                       (diagonal block; Ld = L(j,j-1))
//   sX = Cx - Ax Ld^T   (rows < nrows)      (this CTA's block of column j; Ax = its block of column j-1)
// sub_d / sub_x switch the products off (first column / block outside the band).  want_x = false: only sD.
// mma.sync is warp-collective: nrows is 16 or 32 and tiles are 8 rows high, so every predicate below is warp-uniform.
__device__ __forceinline__ void panel_stage(double* sD, double* sX, const double* __restrict__ Cd, const double* __restrict__ Ld, bool sub_d,
                                            bool want_x, const double* __restrict__ Cx, int ldcx, const double* __restrict__ Ax, int ldax,
                                            int nrows, bool sub_x, int warp, int lane) {
  const int tr = warp >> 1, tc0 = 2 * (warp & 1), g = lane >> 2, t = lane & 3;
  const int r = 8 * tr + g;
  const bool live_x = want_x && (8 * tr) < nrows;
  sub_x = sub_x && live_x;
  double ad[8], ax[8], b0[8], b1[8];
  if (sub_d || sub_x) {
    const double* pb0 = Ld + (size_t)(8 * tc0 + g) * kNB + 8 * t;
    load_frag8(pb0, b0); load_frag8(pb0 + 8 * kNB, b1);
  }
  if (sub_d) load_frag8(Ld + (size_t)r * kNB + 8 * t, ad);
  if (sub_x) load_frag8(Ax + (size_t)r * ldax + 8 * t, ax);
  const int c = 8 * tc0 + 2 * t;
  double2 d0 = *reinterpret_cast<const double2*>(Cd + (size_t)r * kNB + c);
  double2 d1 = *reinterpret_cast<const double2*>(Cd + (size_t)r * kNB + c + 8);
  double2 x0 = make_double2(0.0, 0.0), x1 = make_double2(0.0, 0.0);
  if (live_x) {
    x0 = *reinterpret_cast<const double2*>(Cx + (size_t)r * ldcx + c);
    x1 = *reinterpret_cast<const double2*>(Cx + (size_t)r * ldcx + c + 8);
  }
  if (sub_d) {
    double p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) { dmma884(p0, p1, ad[ks], b0[ks]); dmma884(q0, q1, ad[ks], b1[ks]); }
    d0.x -= p0; d0.y -= p1; d1.x -= q0; d1.y -= q1;
  }
  if (sub_x) {
    double p0 = 0.0, p1 = 0.0, q0 = 0.0, q1 = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) { dmma884(p0, p1, ax[ks], b0[ks]); dmma884(q0, q1, ax[ks], b1[ks]); }
    x0.x -= p0; x0.y -= p1; x1.x -= q0; x1.y -= q1;
  }
  sD[r * kLd + c] = d0.x; sD[r * kLd + c + 1] = d0.y; sD[r * kLd + c + 8] = d1.x; sD[r * kLd + c + 9] = d1.y;
  if (want_x) { sX[r * kLd + c] = x0.x; sX[r * kLd + c + 1] = x0.y; sX[r * kLd + c + 8] = x1.x; sX[r * kLd + c + 9] = x1.y; }
}

// Branch-free 1/sqrt(x) for x > 0 in the normal range: MUFU.RSQ64H seed (~2^-22) + one third-order (Halley) step
//   y <- y (1 + e/2 + 3e^2/8),  e = 1 - x y^2        (relative error ~ e^3: full double precision up to an ulp or two, which the
// factorisation tolerates: L_cc is formed as pivot * rinv).  Four dependent FP64 operations.  Bad input gives NaN / inf, caught by the caller.
__device__ __forceinline__ double fast_rsqrt(double x) {
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  const double t = x * y;
  const double e = fma(-t, y, 1.0);
  const double p = fma(0.375, e, 0.5);
  const double ye = y * e;
  return fma(ye, p, y);
}

// Fused 32x32 Cholesky (sD) + triangular solve X L^T = ACC (sX, in place) by the whole CTA, blocked by 8 columns.  The rows of X
// behave like extra rows of the matrix below the diagonal block, so the solve is pipelined one column block behind the factorisation
// and leaves the critical path (small code: the loop over column blocks is not unrolled).
//   P(cb): warp 0 (lane = row of D, 8 columns in registers): right-looking panel; per column the dependent chain is
//            broadcast 1/sqrt(pivot) -> scale -> update of the next pivot on its own lane -> rsqrt of it   (one shuffle per column)
//          warp 1 (lane = row of X): columns of block cb-1:  x_c = (x_c - sum_{m<c} x_m L[c][m]) / L[c][c]   (registers, broadcast LDS)
//   T(cb): all warps (column kk -> warp kk%8, lane = row): trailing columns kk of D for panel cb and of X for panel cb-1:
//            A[r][kk] -= sum_m L[r][c0+m] L[kk][c0+m]
// On exit sD holds L (lower, zeros above the diagonal), sRinv[c] = 1 / L[c][c], sX holds X.
__device__ __forceinline__ void potrf_trsm32_cta(double* sD, double* sX, double* sRinv, int warp, int lane, int32_t* flag) {
  auto solve_x_block = [&](int x0) {   // warp-level: columns x0..x0+7 of X with the final 8x8 diagonal block of L
    double x[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) x[k] = sX[lane * kLd + x0 + k];
#pragma unroll
    for (int c = 0; c < 8; ++c) {
      x[c] *= sRinv[x0 + c];
#pragma unroll
      for (int k = c + 1; k < 8; ++k) x[k] = fma(-x[c], sD[(x0 + k) * kLd + x0 + c], x[k]);
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) sX[lane * kLd + x0 + k] = x[k];
  };
  auto trailing = [&](double* M, int c0, bool lower_only) {   // columns kk >= c0+8 of M (rows = lanes) with the panel L[:, c0..c0+7]
    double lr[8];
#pragma unroll
    for (int m = 0; m < 8; ++m) lr[m] = M[lane * kLd + c0 + m];
    for (int kk = c0 + 8 + warp; kk < kNB; kk += 8) {
      const double* Lk = sD + kk * kLd + c0;
      double s0 = 0.0, s1 = 0.0;
#pragma unroll
      for (int m = 0; m < 8; m += 2) { s0 = fma(lr[m], Lk[m], s0); s1 = fma(lr[m + 1], Lk[m + 1], s1); }
      if (!lower_only || lane >= kk) M[lane * kLd + kk] -= s0 + s1;
    }
  };
#pragma unroll 1
  for (int c0 = 0; c0 < kNB; c0 += 8) {
    if (warp == 0) {
      double a[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) a[k] = sD[lane * kLd + c0 + k];
      bool bad = false;
      // every lane runs the (branch-free) arithmetic of the pivot lane on its own data; only the pivot lane's value is broadcast
      double rinv_own = fast_rsqrt(a[0]);
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        const double rinv = __shfl_sync(0xffffffffu, rinv_own, c0 + c);
        bad |= !(rinv > 0.0) || !(rinv < 1e150);   // pivot <= 0, NaN or 0: flagged, the (garbage) factor is discarded by the LM loop
        const double l = a[c] * rinv;
        a[c] = l;
        if (lane == c0 + c) sRinv[c0 + c] = rinv;
        if (c < 7) rinv_own = fast_rsqrt(fma(-l, l, a[c + 1]));   // next pivot: final on its own lane before the shuffles below complete
#pragma unroll
        for (int k = c + 1; k < 8; ++k) { const double lk = __shfl_sync(0xffffffffu, l, c0 + k); a[k] = fma(-l, lk, a[k]); }
      }
#pragma unroll
      for (int k = 0; k < 8; ++k) sD[lane * kLd + c0 + k] = (c0 + k <= lane) ? a[k] : 0.0;
      if (bad && lane == 0) atomicExch(flag, 1);
    } else if (warp == 1 && c0 > 0) {
      solve_x_block(c0 - 8);
    }
    __syncthreads();
    if (c0 + 8 < kNB) trailing(sD, c0, true);
    if (c0 > 0) trailing(sX, c0 - 8, false);
    __syncthreads();
  }
  if (warp == 1) solve_x_block(kNB - 8);
}

// ---- time partitions ---------------------------------------------------------------------------------------------------------
// The factorisation is sequential along the band, so the block columns are cut into P "interiors" I_0 .. I_{P-1} separated by P-1
// "separators" S_q of nbw block columns each (S_q lies between I_q and I_{q+1}; an interior is at least nbw columns wide, so separators
// do not couple directly).  All interiors are eliminated CONCURRENTLY (one block column of every interior per launch).  Interior p >= 1
// sees its left separator S_{p-1} as extra ("fat") border rows: the coupling fills in along the whole interior exactly like the rows of
// the global parameters do.  What remains is a small block-tridiagonal interface system in the separators (+ globals), solved by the same
// code with P = 1.  Sequential block steps: max|I_p| + (P-1) nbw instead of nbc.  This is also the multi-GPU decomposition (partition
// per rank, interface system all-gathered).
//
// Border storage of partition p ("FB"): row-major, R = 32*(nrb-1) + 16 rows (nrb-1 = nbw row blocks of S_{p-1} for p >= 1, none for
// p = 0; the last 16 rows are the 15 globals + right-hand side), ld = 32 * ncols doubles, ncols = |I_p| + (nbw columns of S_p if p < P-1).
constexpr int kMaxParts = 64;
struct BandPart {
  int c0;        // first block column of the interior
  int nI;        // interior block columns
  int nrb;       // border row blocks (the last one has 16 rows)
  int ncols;     // block columns covered by fb: nI (+ nbw for the right separator)
  int ld;        // 32 * ncols
  double* fb;    // border rows of this partition
};
// sepw: block columns per separator (nbw at level 0).  defer = 1 (every level that HAS partitions): the border rows are not part of the block
// steps; they trail the band factorisation as a triangular solve with many right-hand sides (k_band_border).
struct BandParts { int P; int sepw; int defer; BandPart part[kMaxParts]; };

// task counts of partition q at step t (shared by host and device)
__host__ __device__ inline void part_tasks(const BandLayout& B, const BandPart& q, int t, int defer, int* n_panel, int* m_upd, int* k_lo) {
  const int j = q.c0 + t;
  const bool has_panel = t < q.nI;
  // a partition never reaches past its right separator: at level 0 that is the band reach anyway, at the interface levels (band 2 sepw - 1,
  // interiors one separator wide) the blocks beyond are structurally zero AND belong to a partition that is being factored concurrently
  const int last = q.c0 + q.ncols - 1 < B.nbc - 1 ? q.c0 + q.ncols - 1 : B.nbc - 1;
  const int rows = has_panel ? (B.nbw < last - j ? B.nbw : last - j) : 0;
  *n_panel = has_panel ? 1 + rows + (defer ? 0 : q.nrb) : 0;
  *k_lo = has_panel ? j + 1 : j;
  int k_hi = j - 1 + B.nbw; if (k_hi > last) k_hi = last;
  int m = (t >= 1 && t <= q.nI) ? k_hi - *k_lo + 1 : 0;
  *m_upd = m > 0 ? m : 0;
}
// update tasks of one partition at one step: 64 x 64 super-tiles (2 x 2 blocks) over the lower triangle of the m trailing block columns
// and over (border row blocks) x (trailing block columns)
__host__ __device__ inline int part_update_tasks(int m, int nrb) {
  const int ms = (m + 1) >> 1, nrs = (nrb + 1) >> 1;
  return m > 0 ? ms * (ms + 1) / 2 + ms * nrs : 0;
}



// Debug build only (-DCTBA_BAND_TRACE, profiles/tools/band_trace.py): first CTA start / last CTA end of every step launch in nanoseconds
// (globaltimer), per kernel kind (0 panel, 1 gather, 2 border chunk) — the timeline of one factorisation as it runs inside the graph.
#ifdef CTBA_BAND_TRACE
constexpr int kTraceSteps = 8192;
__device__ unsigned long long g_band_trace[3][kTraceSteps][2];
__device__ __forceinline__ unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#define BAND_TRACE_BEGIN(kind, t) do { if (threadIdx.x == 0 && (t) < kTraceSteps) atomicMin(&g_band_trace[kind][t][0], trace_now()); } while (0)
#define BAND_TRACE_END(kind, t) do { if (threadIdx.x == 0 && (t) < kTraceSteps) atomicMax(&g_band_trace[kind][t][1], trace_now()); } while (0)
#else
#define BAND_TRACE_BEGIN(kind, t) do {} while (0)
#define BAND_TRACE_END(kind, t) do {} while (0)
#endif
#ifdef CTBA_BAND_TIMING
__device__ long long g_band_clk[8];
#define BAND_CLK(k) do { if (j == 100 && b == 1 && tid == 0) g_band_clk[k] = clock64(); } while (0)
__device__ long long g_border_clk[8][8];   // [column of the chunk][phase] of one border CTA (blockIdx 30, chunk starting at step 304)
#define BORDER_CLK(k) do { if (blockIdx.x == 30 && t0 == 304 && threadIdx.x == 0) g_border_clk[t - t0][k] = clock64(); } while (0)
#else
#define BAND_CLK(k) do {} while (0)
#define BORDER_CLK(k) do {} while (0)
#endif

// Two launches per step t (they run concurrently, on two streams / two graph branches): every partition p advances to its block column
// j = c0_p + t (right-looking with a one-column look-ahead):
//   k_band_panel  : apply column j-1's update to the CTA's OWN block of column j (and to the diagonal block, redundantly), factor the diagonal
//                   block, triangular-solve the block, store it.  Task 0 = diagonal (also stores L_jj^-1 for the backward substitution),
//                   tasks 1..rows = sub-diagonal blocks, then one task per border row block.  This is the critical path.
//   k_band_update : remaining trailing update of column j-1:  T(i,k) -= L(i,j-1) L(k,j-1)^T  for k_lo <= k <= i <= j-1+nbw, and the border
//                   row blocks of those block columns.  k_lo = j+1 while the partition still has a panel (the panel CTAs cover column j),
//                   k_lo = j in the partition's final flush step t = |I_p|.  Small register footprint, so that many CTAs are resident.
// Both depend on BOTH kernels of step t-1.  Pre-condition of step t: columns < j of the interior final, every trailing block carries the
// updates of columns <= j-2.
__global__ void __launch_bounds__(kStepThreads) k_band_panel(BandLayout B, const __grid_constant__ BandParts PS, int t, double* __restrict__ dinv,
                                                             int32_t* __restrict__ flag) {
  __shared__ double sD[kNB * kLd];
  __shared__ double sX[kNB * kLd];
  __shared__ double sRinv[kNB];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nb1 = B.nbw + 1;
  int b = blockIdx.x, pi = 0, n_panel = 0, m = 0, k_lo = 0;
  for (; pi < PS.P; ++pi) {
    part_tasks(B, PS.part[pi], t, PS.defer, &n_panel, &m, &k_lo);
    if (b < n_panel) break;
    b -= n_panel;
  }
  if (pi >= PS.P) return;
  const BandPart& Q = PS.part[pi];
  const int j = Q.c0 + t;
  BAND_TRACE_BEGIN(0, t);
  const bool prev = t > 0 && B.nbw >= 1;
  const double* Lprev = B.band + ((size_t)max(j - 1, 0) * nb1) * 1024;  // column j-1: block (j-1+o, j-1) at offset o
  const int lcol = j - Q.c0;                                            // column of fb
  const int rows = n_panel - 1 - (PS.defer ? 0 : Q.nrb);
  double* Ajj = B.band + ((size_t)j * nb1) * 1024;
  BAND_CLK(0);
  int rb = -1, nrows = kNB;
  double* Fb = nullptr;
  if (b == 0) {
    panel_stage(sD, sX, Ajj, Lprev + 1024, prev, false, nullptr, 0, nullptr, 0, 0, false, warp, lane);
    // identity right-hand side: X L^T = I gives X = L_jj^-T, whose transpose L_jj^-1 is kept for the backward substitution
    for (int e = tid; e < kNB * kNB; e += kStepThreads) sX[(e >> 5) * kLd + (e & 31)] = ((e >> 5) == (e & 31)) ? 1.0 : 0.0;
  } else if (b <= rows) {
    const bool have_prev = prev && (b + 1 <= B.nbw);   // block (j+b, j-1) lies inside the band
    panel_stage(sD, sX, Ajj, Lprev + 1024, prev, true, Ajj + (size_t)b * 1024, kNB, Lprev + (size_t)min(b + 1, B.nbw) * 1024, kNB, kNB, have_prev,
                warp, lane);
  } else {
    rb = b - rows - 1;
    nrows = (rb == Q.nrb - 1) ? kBorder : kNB;
    Fb = Q.fb + (size_t)rb * kNB * Q.ld + (size_t)lcol * kNB;
    panel_stage(sD, sX, Ajj, Lprev + 1024, prev, true, Fb, Q.ld, Fb - kNB, Q.ld, nrows, prev, warp, lane);
  }
  __syncthreads();
  BAND_CLK(1);
  potrf_trsm32_cta(sD, sX, sRinv, warp, lane, flag);
  BAND_CLK(2);
  __syncthreads();
  BAND_CLK(3);
  if (b == 0) {
    // The diagonal block is NOT written back: every other CTA of this launch stages the un-factored A_jj from the same address, and a CTA
    // that is scheduled late (busy GPU, many partitions) would otherwise read L_jj instead.  Nothing downstream needs L_jj itself — the
    // substitutions use L_jj^-1.
    for (int e = tid; e < kNB * kNB; e += kStepThreads) {
      const int r = e >> 5, c = e & 31;
      dinv[(size_t)j * 1024 + e] = sX[c * kLd + r];   // (L^-1)[r][c] = X[c][r]
    }
  } else if (b <= rows) {
    double* Lij = Ajj + (size_t)b * 1024;
    for (int e = tid; e < kNB * kNB; e += kStepThreads) Lij[e] = sX[(e >> 5) * kLd + (e & 31)];
    BAND_CLK(4);
  } else {
    for (int e = tid; e < nrows * kNB; e += kStepThreads) { const int r = e >> 5, c = e & 31; Fb[(size_t)r * Q.ld + c] = sX[r * kLd + c]; }
  }
  BAND_TRACE_END(0, t);
}

// CTA = one super-tile; warp (wr, wc) owns rows 16 (wr & 1) .. +16 of row block (wr >> 1) and the 32 columns of column block wc of it:
// 2 x 4 DMMA tiles from two A fragments and four B fragments (the one-block-per-CTA version loaded three fragments per two tiles), all
// operand and target loads issued before the first DMMA.
__global__ void __launch_bounds__(kStepThreads, 2) k_band_update(BandLayout B, const __grid_constant__ BandParts PS, int t) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int nb1 = B.nbw + 1;
  int u = blockIdx.x, pi = 0, n_panel = 0, m = 0, k_lo = 0;
  for (; pi < PS.P; ++pi) {
    part_tasks(B, PS.part[pi], t, PS.defer, &n_panel, &m, &k_lo);
    const int nu = part_update_tasks(m, PS.defer ? 0 : PS.part[pi].nrb);
    if (u < nu) break;
    u -= nu;
  }
  if (pi >= PS.P) return;
  const BandPart& Q = PS.part[pi];
  const int j = Q.c0 + t;
  BAND_TRACE_BEGIN(1, t);
  const double* Lprev = B.band + ((size_t)(j - 1) * nb1) * 1024;  // column j-1: block (j-1+o, j-1) at offset o
  const int ms = (m + 1) >> 1, ntri = ms * (ms + 1) / 2;
  const int g = lane >> 2, tq = lane & 3, wr = warp >> 1, wc = warp & 1;
  const int half = wr & 1;                        // which 16 rows of the row block
  const double* A; const double* Lk; double* T; int lda, ldt;
  if (u < ntri) {
    int sa = (int)((sqrtf(8.0f * u + 1.0f) - 1.0f) * 0.5f);
    while ((sa + 1) * (sa + 2) / 2 <= u) ++sa;
    while (sa * (sa + 1) / 2 > u) --sa;
    const int sb = u - sa * (sa + 1) / 2;
    const int a = 2 * sa + (wr >> 1), c = 2 * sb + wc;          // block row / column inside the trailing window, 0 <= c, a < m
    if (a >= m || c >= m || c > a) return;                      // warp-uniform: outside the window or above the diagonal
    const int i = k_lo + a, k = k_lo + c;
    A = Lprev + (size_t)(i - (j - 1)) * 1024; lda = kNB;
    Lk = Lprev + (size_t)(k - (j - 1)) * 1024;
    T = B.band + ((size_t)k * nb1 + (i - k)) * 1024; ldt = kNB;
  } else {                                        // border row blocks (2 per task) x block columns (2 per task)
    u -= ntri;
    const int rs = u / ms, cs = u - rs * ms;
    const int rb = 2 * rs + (wr >> 1), c = 2 * cs + wc;
    if (rb >= Q.nrb || c >= m) return;
    if (rb == Q.nrb - 1 && half * 16 >= kBorder) return;        // the last row block has 16 rows
    const int k = k_lo + c;
    Lk = Lprev + (size_t)(k - (j - 1)) * 1024;
    A = Q.fb + (size_t)rb * kNB * Q.ld + (size_t)(j - 1 - Q.c0) * kNB; lda = Q.ld;
    T = Q.fb + (size_t)rb * kNB * Q.ld + (size_t)(k - Q.c0) * kNB; ldt = Q.ld;
  }
  const int r0 = 16 * half + g;
  double a0[8], a1[8], b0[8], b1[8], b2[8], b3[8];
  load_frag8(A + (size_t)r0 * lda + 8 * tq, a0);
  load_frag8(A + (size_t)(r0 + 8) * lda + 8 * tq, a1);
  const double* pb = Lk + (size_t)g * kNB + 8 * tq;
  load_frag8(pb, b0); load_frag8(pb + 8 * kNB, b1); load_frag8(pb + 16 * kNB, b2); load_frag8(pb + 24 * kNB, b3);
  double2* d0 = reinterpret_cast<double2*>(T + (size_t)r0 * ldt + 2 * tq);
  double2* d1 = reinterpret_cast<double2*>(T + (size_t)(r0 + 8) * ldt + 2 * tq);
  double2 v[2][4];
#pragma unroll
  for (int h = 0; h < 4; ++h) { v[0][h] = d0[4 * h]; v[1][h] = d1[4 * h]; }
  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) {
    dmma884(acc[0][0][0], acc[0][0][1], a0[ks], b0[ks]); dmma884(acc[1][0][0], acc[1][0][1], a1[ks], b0[ks]);
    dmma884(acc[0][1][0], acc[0][1][1], a0[ks], b1[ks]); dmma884(acc[1][1][0], acc[1][1][1], a1[ks], b1[ks]);
    dmma884(acc[0][2][0], acc[0][2][1], a0[ks], b2[ks]); dmma884(acc[1][2][0], acc[1][2][1], a1[ks], b2[ks]);
    dmma884(acc[0][3][0], acc[0][3][1], a0[ks], b3[ks]); dmma884(acc[1][3][0], acc[1][3][1], a1[ks], b3[ks]);
  }
#pragma unroll
  for (int h = 0; h < 4; ++h) {
    v[0][h].x -= acc[0][h][0]; v[0][h].y -= acc[0][h][1]; v[1][h].x -= acc[1][h][0]; v[1][h].y -= acc[1][h][1];
    d0[4 * h] = v[0][h]; d1[4 * h] = v[1][h];
  }
  BAND_TRACE_END(1, t);
}

// ---- K loop shared by the border kernels ---------------------------------------------------------------------------------
// One warp accumulates  acc += sum_k A_k Bm_k^T  for a whole 32 x 32 tile (16 accumulator tiles of 8 x 8 = 16 independent DMMA chains) over
// the blocks k = 0 .. nk-1 it is given (A_k = pa + k a_step with `a_rs` doubles between rows, Bm_k = pb + k b_step, row-major 32 x 32 with
// leading dimension 32).  Operand fragments come straight from L2 — profiles/fp64_peaks.json: 31-33 TFLOP/s of the 37.1 TFLOP/s DMMA peak for
// this loop shape against 25 TFLOP/s when the blocks are staged through shared memory by cp.async.bulk — and a block is consumed in two
// halves (fragment columns 8 t + 4 half .. + 4: one 256-bit load per 8-row group and operand); the next half is in flight while the current
// one feeds 64 DMMAs (1024 issue cycles per warp: more than the L2 latency, so one stage of prefetch hides it).
// (measured: DMMA.8x8x4 issues every ~16 cycles per SM sub-partition with a dependent latency of ~27 cycles, so 2 chains per warp suffice.)
// COHERENT: the A operand was written by this CTA earlier in the same launch (plain loads instead of the read-only path).
// rows_a: A rows >= rows_a are not read (taken as zero) — the 16-row border of an unpartitioned system.
struct KFrag { double a[4][4], b[4][4]; };
template <bool COHERENT>
__device__ __forceinline__ void kfrag_load(KFrag& f, const double* pa, size_t a_rs, const double* __restrict__ pb, int half, int rows_a) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (8 * i < rows_a) {
      if (COHERENT) asm volatile("ld.global.L1::evict_last.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(f.a[i][0]), "=d"(f.a[i][1]), "=d"(f.a[i][2]), "=d"(f.a[i][3]) : "l"(pa + (size_t)(8 * i) * a_rs + 4 * half) : "memory");
      else asm volatile("ld.global.nc.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(f.a[i][0]), "=d"(f.a[i][1]), "=d"(f.a[i][2]), "=d"(f.a[i][3]) : "l"(pa + (size_t)(8 * i) * a_rs + 4 * half));
    } else {
      f.a[i][0] = f.a[i][1] = f.a[i][2] = f.a[i][3] = 0.0;
    }
    asm volatile("ld.global.nc.L1::no_allocate.v4.f64 {%0,%1,%2,%3}, [%4];" : "=d"(f.b[i][0]), "=d"(f.b[i][1]), "=d"(f.b[i][2]), "=d"(f.b[i][3]) : "l"(pb + (size_t)(8 * i) * kNB + 4 * half));
  }
}
__device__ __forceinline__ void kfrag_mma(double (&acc)[4][4][2], const KFrag& f) {
#pragma unroll
  for (int ks = 0; ks < 4; ++ks)
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int h = 0; h < 4; ++h) dmma884(acc[i][h][0], acc[i][h][1], f.a[i][ks], f.b[h][ks]);
}
// pa / pb point at this lane's fragment origin (row g, column 8 tq) of the first block
template <bool COHERENT>
__device__ __forceinline__ void kloop(double (&acc)[4][4][2], const double* pa, size_t a_rs, size_t a_step, const double* __restrict__ pb, size_t b_step,
                                      int nk, int rows_a) {
  if (nk <= 0) return;
  KFrag f0, f1;
  kfrag_load<COHERENT>(f0, pa, a_rs, pb, 0, rows_a);
  for (int k = 0; k < nk; ++k) {
    kfrag_load<COHERENT>(f1, pa, a_rs, pb, 1, rows_a);
    kfrag_mma(acc, f0);
    pa += a_step; pb += b_step;
    if (k + 1 < nk) kfrag_load<COHERENT>(f0, pa, a_rs, pb, 0, rows_a);
    kfrag_mma(acc, f1);
  }
}
// The same sum over the SAME half (fragment columns 8 t + 4 half .. + 3) of every second block: n half-blocks starting at pa / pb, the next one
// a_step / b_step further.  Used by the border pipeline to deal the k range to its 4 warps in half-blocks (warps 0, 1: the two halves of the
// even blocks, warps 2, 3: of the odd blocks), so that 21 blocks split 11 / 11 / 10 / 10 half-blocks instead of 12 / 10 / 10 / 10.
template <bool COHERENT>
__device__ __forceinline__ void kloop_half(double (&acc)[4][4][2], const double* pa, size_t a_rs, ptrdiff_t a_step, const double* __restrict__ pb,
                                           ptrdiff_t b_step, int n, int half, int rows_a) {
  if (n <= 0) return;
  KFrag f0, f1;
  kfrag_load<COHERENT>(f0, pa, a_rs, pb, half, rows_a);
  for (int i = 0; i < n; i += 2) {
    if (i + 1 < n) kfrag_load<COHERENT>(f1, pa + a_step, a_rs, pb + b_step, half, rows_a);
    kfrag_mma(acc, f0);
    pa += 2 * a_step; pb += 2 * b_step;
    if (i + 1 < n) {
      if (i + 2 < n) kfrag_load<COHERENT>(f0, pa, a_rs, pb, half, rows_a);
      kfrag_mma(acc, f1);
    }
  }
}
__device__ __forceinline__ void cp_async16(double* dst_smem, const double* src) {   // 16 bytes global -> shared, through L2, no registers
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(dst_smem)), "l"(src) : "memory");
}
constexpr int kLdS = 36;   // shared-memory leading dimension: (4 g + t) mod 16 distinct -> conflict-free 8-byte fragment reads (col = 4 ks + t)
__device__ __forceinline__ void acc_to_smem(double* sP, const double (&acc)[4][4][2], int g, int tq) {
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) {
      double* d = &sP[(8 * i + g) * kLdS + 8 * h + 2 * tq];
      d[0] = acc[i][h][0]; d[1] = acc[i][h][1];
    }
}

// ---- deferred border rows:  W = FB L^-T  as a pipeline that trails the band factorisation ------------------------------------------------
// The border rows of a partition (the coupling to its left separator + the global rows) are not needed by any band block, only by later
// columns of the SAME border row block, by the separator gather and by the Gram matrices.  They are therefore taken off the block steps
// (where each border row block cost one more CTA with its own redundant 32 x 32 factorisation on the critical path) and computed by this
// kernel, launched once per chunk of block steps on its own stream:  CTA = (partition, border row block) walks the chunk's block columns
//     W(r, j) = ( FB(r, j) - sum_{k = max(c0, j - nbw)}^{j-1} W(r, k) L(j, k)^T ) L_jj^-T
// in order; every dependency is either final (L, L_jj^-1 from the panel kernel) or the CTA's own earlier output.  4 warps split the k range,
// the partial sums meet in shared memory, then the product with L_jj^-1 (staged in shared memory while the k loop runs) gives the block.
constexpr int kBorderThreads = 128;
constexpr size_t kBorderSmem = (size_t)6 * kNB * kLdS * sizeof(double);   // 4 partial sums + L_jj^-1 + FB(r, j)
__global__ void __launch_bounds__(kBorderThreads) k_band_border(BandLayout B, const __grid_constant__ BandParts PS, int t0, int t1,
                                                                const double* __restrict__ dinv) {
  extern __shared__ __align__(16) double border_smem[];   // kBorderSmem bytes (above the 48 KB a kernel may declare statically)
  double (*sP)[kNB * kLdS] = reinterpret_cast<double (*)[kNB * kLdS]>(border_smem);   // [4] per-warp partial sums; sP[0] becomes T = FB - sum
  double* sL = border_smem + 4 * kNB * kLdS;              // L_jj^-1
  double* sF = border_smem + 5 * kNB * kLdS;              // FB(r, j), the column's own block before the update
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tq = lane & 3;
  const int nb1 = B.nbw + 1;
  int u = blockIdx.x, pi = 0;
  for (; pi < PS.P; ++pi) { if (u < PS.part[pi].nrb) break; u -= PS.part[pi].nrb; }
  if (pi >= PS.P) return;
  const BandPart& Q = PS.part[pi];
  double* F = Q.fb + (size_t)u * kNB * Q.ld;                       // row block u of the partition's border, row stride Q.ld
  const int tend = t1 < Q.nI ? t1 : Q.nI;
  BAND_TRACE_BEGIN(2, t0);
  for (int t = t0; t < tend; ++t) {
    const int j = Q.c0 + t;
    const int k_lo = Q.c0 > j - B.nbw ? Q.c0 : j - B.nbw;
    BORDER_CLK(0);
    // L_jj^-1 and the column's own block -> shared memory by cp.async: in flight during the k loop, no registers, consumed after the barriers
    // below (the in-kernel clocks showed ~1000 cycles of exposed latency for a register-staged L_jj^-1 before the k loop could start and
    // ~1700 for the global read of FB(r, j) between the two barriers)
    double* Fj = F + (size_t)t * kNB;
    for (int e = threadIdx.x; e < kNB * kNB / 2; e += kBorderThreads) {
      const int r = e >> 4, c = (e & 15) << 1;
      cp_async16(&sL[r * kLdS + c], dinv + (size_t)j * 1024 + r * kNB + c);
      cp_async16(&sF[r * kLdS + c], Fj + (size_t)r * Q.ld + c);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    BORDER_CLK(1);
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
    // warp w: half w & 1 of the blocks k_lo + (w >> 1), + 2, + 4, ...
    const int nk_all = j - k_lo, b0 = warp >> 1, nh = b0 < nk_all ? (nk_all - b0 + 1) / 2 : 0;
    // The CTA re-reads its own last nbw blocks W(r, k) every column (they are the A operand): those loads allocate in L1 with evict-last, the
    // L blocks and every other fragment stream of the factorisation do not allocate, and the k order alternates from column to column
    // (ascending on even t, descending on odd t) so that a working set slightly larger than L1 still hits on the blocks used last.
    {
      const double* pa0 = F + (size_t)g * Q.ld + (size_t)(k_lo + b0 - Q.c0) * kNB + 8 * tq;
      const double* pb0 = B.band + ((size_t)(k_lo + b0) * nb1 + (j - k_lo - b0)) * 1024 + (size_t)g * kNB + 8 * tq;
      const ptrdiff_t as = 2 * kNB, bs = (ptrdiff_t)2 * (nb1 - 1) * 1024;
      const bool down = (t & 1) && nh > 0;
      kloop_half<true>(acc, down ? pa0 + (nh - 1) * as : pa0, Q.ld, down ? -as : as, down ? pb0 + (nh - 1) * bs : pb0, down ? -bs : bs, nh, warp & 1, kNB);
    }
    BORDER_CLK(2);
    acc_to_smem(sP[warp], acc, g, tq);
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    BORDER_CLK(3);
    for (int e = threadIdx.x; e < kNB * kNB; e += kBorderThreads) {
      const int r = e >> 5, c = e & 31, o = r * kLdS + c;
      sP[0][o] = sF[o] - ((sP[0][o] + sP[1][o]) + (sP[2][o] + sP[3][o]));
    }
    __syncthreads();
    BORDER_CLK(4);
    // W(r, j) = T L_jj^-T:  X[r][c] = sum_k T[r][k] Linv[c][k];  warp w: row group w, all four column groups; natural k layout (col = 4 ks + t)
    double x[4][2][2];
#pragma unroll
    for (int h = 0; h < 4; ++h) x[h][0][0] = x[h][0][1] = x[h][1][0] = x[h][1][1] = 0.0;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) {
      const double av = sP[0][(8 * warp + g) * kLdS + 4 * ks + tq];
#pragma unroll
      for (int h = 0; h < 4; ++h) dmma884(x[h][ks & 1][0], x[h][ks & 1][1], av, sL[(8 * h + g) * kLdS + 4 * ks + tq]);
    }
#pragma unroll
    for (int h = 0; h < 4; ++h)
      *reinterpret_cast<double2*>(Fj + (size_t)(8 * warp + g) * Q.ld + 8 * h + 2 * tq) = make_double2(x[h][0][0] + x[h][1][0], x[h][0][1] + x[h][1][1]);
    BORDER_CLK(5);
    __syncthreads();   // the block is visible to the whole CTA before the next column reads it back; sP / sL may be overwritten
    BORDER_CLK(6);
  }
  BAND_TRACE_END(2, t0);
}

// Border rows over the right separator of a partition: after the border pipeline has finished the interior,
//     FB(r, c) -= sum_{k = max(c0, c - nbw)}^{c0 + nI - 1} W(r, k) L(c, k)^T      for every separator column c
// (the band blocks of the separator columns are updated by k_band_update step by step; the border rows are not part of the steps).
// CTA = (partition, separator column, border row block); 4 warps split the k range.
__global__ void __launch_bounds__(kBorderThreads) k_band_border_sep(BandLayout B, const __grid_constant__ BandParts PS) {
  __shared__ double sP[4][kNB * kLdS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tq = lane & 3;
  const int nb1 = B.nbw + 1;
  int u = blockIdx.x, pi = 0;
  for (; pi < PS.P; ++pi) { const int nu = (PS.part[pi].ncols - PS.part[pi].nI) * PS.part[pi].nrb; if (u < nu) break; u -= nu; }
  if (pi >= PS.P) return;
  const BandPart& Q = PS.part[pi];
  const int tc = u / Q.nrb, rb = u - tc * Q.nrb;
  const int c = Q.c0 + Q.nI + tc, k_hi = Q.c0 + Q.nI - 1;
  const int k_lo = Q.c0 > c - B.nbw ? Q.c0 : c - B.nbw;
  if (k_lo > k_hi) return;
  double* F = Q.fb + (size_t)rb * kNB * Q.ld;
  double acc[4][4][2];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
  const int nk_all = k_hi - k_lo + 1, nk = warp < nk_all ? (nk_all - warp + 3) / 4 : 0;
  kloop<false>(acc, F + (size_t)g * Q.ld + (size_t)(k_lo + warp - Q.c0) * kNB + 8 * tq, Q.ld, 4 * kNB,
               B.band + ((size_t)(k_lo + warp) * nb1 + (c - k_lo - warp)) * 1024 + (size_t)g * kNB + 8 * tq, (size_t)4 * (nb1 - 1) * 1024, nk, kNB);
  acc_to_smem(sP[warp], acc, g, tq);
  __syncthreads();
  double* T = F + (size_t)(c - Q.c0) * kNB;
  for (int e = threadIdx.x; e < kNB * kNB; e += kBorderThreads) {
    const int r = e >> 5, cc = e & 31, o = r * kLdS + cc;
    T[(size_t)r * Q.ld + cc] -= (sP[0][o] + sP[1][o]) + (sP[2][o] + sP[3][o]);
  }
}

// ---- partition border initialisation -------------------------------------------------------------------------------------------
// fb_p = [ coupling of the left separator S_{p-1} to the partition's columns (transposed band blocks, zero beyond the band) ; global rows ]
__global__ void __launch_bounds__(256) k_fb_init(BandLayout B, const __grid_constant__ BandParts PS) {
  int blk = blockIdx.x, pi = 0;
  for (; pi < PS.P; ++pi) {
    const int nb = PS.part[pi].nrb * PS.part[pi].ncols;
    if (blk < nb) break;
    blk -= nb;
  }
  if (pi >= PS.P) return;
  const BandPart& Q = PS.part[pi];
  const int rb = blk / Q.ncols, col = blk % Q.ncols;
  double* dst = Q.fb + (size_t)rb * kNB * Q.ld + (size_t)col * kNB;
  const int i = Q.c0 + col;
  if (rb == Q.nrb - 1) {   // 15 globals + right-hand side (the 16 pad rows below are never read as data)
    for (int e = threadIdx.x; e < kNB * kNB; e += 256) {
      const int r = e >> 5, c = e & 31;
      dst[(size_t)r * Q.ld + c] = r < kBorder ? B.border[(size_t)r * B.n_pad + (size_t)i * kNB + c] : 0.0;
    }
    return;
  }
  const int k = Q.c0 - (Q.nrb - 1) + rb;    // block column of the left separator that owns this row block
  const bool inband = (i - k) <= B.nbw;
  const double* T = B.band + ((size_t)k * (B.nbw + 1) + (inband ? i - k : 0)) * 1024;   // block (i, k)
  for (int e = threadIdx.x; e < kNB * kNB; e += 256) {
    const int r = e >> 5, c = e & 31;       // fb[r][c] = A[(k,r),(i,c)] = T(i,k)[c][r]
    dst[(size_t)r * Q.ld + c] = inband ? T[c * kNB + r] : 0.0;
  }
}

// ---- Gram matrices of the factored borders:  G_p = W_p W_p^T over the interior columns (lower 32x32 tiles, split over column chunks) ----
constexpr int kGramChunk = 16;
struct GramPlan { int P; int tile0[kMaxParts + 1]; int nsplit[kMaxParts]; int task0[kMaxParts + 1]; int stask0[kMaxParts + 1]; };   // tiles / tasks / super-tile tasks prefix sums
// CTA = a 64 x 64 super-tile (2 x 2 tiles of 32 x 32, lower triangle of super-tiles) over one column chunk; warp (wr, wc) owns rows
// 16 wr .. +16 and columns 32 wc .. +32 of it: 2 x 4 tiles of 8 x 8, so one pair of A fragments and two pairs of B fragments feed 32 DMMAs
// (the 32 x 32 version loaded three fragments per 16).  The partials land in the same [task][1024] scratch layout (task = chunk * ntile
// + 32-tile), so the reduction kernel is unchanged.
__global__ void __launch_bounds__(kStepThreads, 2) k_gram_partial(const __grid_constant__ BandParts PS, const __grid_constant__ GramPlan GP,
                                                                  double* __restrict__ scratch /* [task][1024] */) {
  int task = blockIdx.x, pi = 0;
  while (pi + 1 < GP.P && task >= GP.stask0[pi + 1]) ++pi;
  task -= GP.stask0[pi];
  const BandPart& Q = PS.part[pi];
  const int ntile = GP.tile0[pi + 1] - GP.tile0[pi];
  const int nst = (Q.nrb + 1) >> 1, nsuper = nst * (nst + 1) / 2;
  const int sup = task % nsuper, ks = task / nsuper;
  int sa = (int)((sqrtf(8.0f * sup + 1.0f) - 1.0f) * 0.5f);
  while ((sa + 1) * (sa + 2) / 2 <= sup) ++sa;
  while (sa * (sa + 1) / 2 > sup) --sa;
  const int sb = sup - sa * (sa + 1) / 2;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wr = warp >> 1, wc = warp & 1, g = lane >> 2, tq = lane & 3;
  const int R = kNB * (Q.nrb - 1) + kBorder, Rpad = kNB * Q.nrb;
  const int row0 = 64 * sa + 16 * wr, col0 = 64 * sb + 32 * wc;
  if (row0 >= Rpad || col0 >= Rpad) return;                       // no 32-tile here (odd number of row blocks)
  if ((row0 >> 5) < (col0 >> 5)) return;                          // the upper 32-tile of a diagonal super-tile
  double acc[2][4][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h) acc[i][h][0] = acc[i][h][1] = 0.0;
  const int col_first = ks * kGramChunk, col_last = min(col_first + kGramChunk, Q.nI);
  const bool va0 = row0 < R, va1 = row0 + 8 < R;
  const bool vb0 = col0 < R, vb1 = col0 + 8 < R, vb2 = col0 + 16 < R, vb3 = col0 + 24 < R;
  const double* Fa = Q.fb + (size_t)(row0 + g) * Q.ld + 8 * tq;
  const double* Fb = Q.fb + (size_t)(col0 + g) * Q.ld + 8 * tq;
  const size_t r8 = (size_t)8 * Q.ld;
  for (int col = col_first; col < col_last; ++col) {
    const size_t co = (size_t)col * kNB;
    double a0[8], a1[8], b0[8], b1[8];
    if (va0) load_frag8(Fa + co, a0);
    if (va1) load_frag8(Fa + r8 + co, a1);
    if (vb0) load_frag8(Fb + co, b0);
    if (vb1) load_frag8(Fb + r8 + co, b1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (va0 && vb0) dmma884(acc[0][0][0], acc[0][0][1], a0[k], b0[k]);
      if (va0 && vb1) dmma884(acc[0][1][0], acc[0][1][1], a0[k], b1[k]);
      if (va1 && vb0) dmma884(acc[1][0][0], acc[1][0][1], a1[k], b0[k]);
      if (va1 && vb1) dmma884(acc[1][1][0], acc[1][1][1], a1[k], b1[k]);
    }
    if (vb2) load_frag8(Fb + 2 * r8 + co, b0);
    if (vb3) load_frag8(Fb + 3 * r8 + co, b1);
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      if (va0 && vb2) dmma884(acc[0][2][0], acc[0][2][1], a0[k], b0[k]);
      if (va0 && vb3) dmma884(acc[0][3][0], acc[0][3][1], a0[k], b1[k]);
      if (va1 && vb2) dmma884(acc[1][2][0], acc[1][2][1], a1[k], b0[k]);
      if (va1 && vb3) dmma884(acc[1][3][0], acc[1][3][1], a1[k], b1[k]);
    }
  }
  // rows / columns past R inside an existing 32-tile are written as zeros (the reduction reads whole tiles)
  const int ta = row0 >> 5, tb = col0 >> 5;
  double* out = scratch + ((size_t)(GP.task0[pi] + ks * ntile + ta * (ta + 1) / 2 + tb) << 10);
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int h = 0; h < 4; ++h)
      *reinterpret_cast<double2*>(out + ((row0 & 31) + 8 * i + g) * kNB + 8 * h + 2 * tq) = make_double2(acc[i][h][0], acc[i][h][1]);
}
// G_p tile = sum of its column-chunk partials in fixed order (deterministic); G_p is stored row-major with ld = 32 * nrb
__global__ void __launch_bounds__(256) k_gram_reduce(const __grid_constant__ BandParts PS, const __grid_constant__ GramPlan GP,
                                                     const double* __restrict__ scratch, double* __restrict__ gram /* concatenated G_p */,
                                                     const size_t* __restrict__ gram_off) {
  int t = blockIdx.x, pi = 0;
  while (pi + 1 < GP.P && t >= GP.tile0[pi + 1]) ++pi;
  const int tile = t - GP.tile0[pi];
  const int ntile = GP.tile0[pi + 1] - GP.tile0[pi];
  int ta = (int)((sqrtf(8.0f * tile + 1.0f) - 1.0f) * 0.5f);
  while ((ta + 1) * (ta + 2) / 2 <= tile) ++ta;
  while (ta * (ta + 1) / 2 > tile) --ta;
  const int tb = tile - ta * (ta + 1) / 2;
  const int ldg = kNB * PS.part[pi].nrb;
  double* G = gram + gram_off[pi];
  for (int e = threadIdx.x; e < kNB * kNB; e += 256) {
    double s = 0.0;
    for (int ks = 0; ks < GP.nsplit[pi]; ++ks) s += scratch[((size_t)(GP.task0[pi] + ks * ntile + tile) << 10) + e];
    G[(size_t)(ta * kNB + (e >> 5)) * ldg + tb * kNB + (e & 31)] = s;
  }
}

// ---- interface system in the separators -------------------------------------------------------------------------------------------
// B2: block columns = the separator columns in time order (nbw per separator), half bandwidth 2 nbw - 1 blocks (dense block tridiagonal).
// Sharded solves (ctba_shard.h): every rank assembles only the terms it holds — the raw + right-looking-updated separator blocks of its
// own band copy (zero outside its column range), the Gram matrices and couplings of its own partitions [q_lo, q_hi) — and the interface
// system is the SUM over ranks (one all-reduce).  q_lo = 0, q_hi = P is the single-GPU case.
struct IfaceArgs { const double* gram; size_t gram_off[kMaxParts]; int q_lo, q_hi; double* extra; };
__global__ void __launch_bounds__(256) k_iface_assemble(BandLayout B, BandLayout B2, const __grid_constant__ BandParts PS, const __grid_constant__ IfaceArgs IA,
                                                        const int32_t* __restrict__ flag) {
  const int nb1_2 = B2.nbw + 1;
  const int nblk = B2.nbc * nb1_2;
  const int blk = blockIdx.x;
  const int nbw = PS.sepw;      // separator width (== B.nbw at level 0; the band stride of B is B.nbw + 1)
  auto local = [&](int q) { return q >= IA.q_lo && q < IA.q_hi; };
  if (blk < nblk) {
    const int K2 = blk / nb1_2, off = blk % nb1_2, I2 = K2 + off;
    double* dst = B2.band + ((size_t)blk << 10);
    if (I2 >= B2.nbc) { for (int e = threadIdx.x; e < 1024; e += 256) dst[e] = 0.0; return; }
    const int qi = I2 / nbw, a = I2 % nbw, qk = K2 / nbw, c = K2 % nbw;
    if (qi == qk) {
      const bool adjacent = qk >= IA.q_lo - 1 && qk <= IA.q_hi - 1;      // separator qk borders one of this rank's partitions
      const bool have_g = local(qk + 1);
      const int s = PS.part[qk + 1].c0 - nbw;                         // first block column of separator qk
      const double* T = B.band + ((size_t)(s + c) * (B.nbw + 1) + (a - c)) * 1024;
      const double* G = IA.gram + IA.gram_off[qk + 1];
      const int ldg = kNB * PS.part[qk + 1].nrb;
      for (int e = threadIdx.x; e < 1024; e += 256) {
        double v = adjacent ? T[e] : 0.0;
        if (have_g) v -= G[(size_t)(a * kNB + (e >> 5)) * ldg + c * kNB + (e & 31)];
        dst[e] = v;
      }
    } else if (qi == qk + 1 && local(qi)) {
      const BandPart& Q = PS.part[qi];                                 // the interior between the two separators
      const double* F = Q.fb + (size_t)c * kNB * Q.ld + (size_t)(Q.nI + a) * kNB;   // rows of S_qk (block c), columns of S_qi (block a)
      for (int e = threadIdx.x; e < 1024; e += 256) dst[e] = F[(size_t)(e & 31) * Q.ld + (e >> 5)];
    } else {
      for (int e = threadIdx.x; e < 1024; e += 256) dst[e] = 0.0;
    }
    return;
  }
  const int bb = blk - nblk;
  if (bb < B2.nbc) {   // global rows over separator column K2
    const int K2 = bb, qk = K2 / nbw, c = K2 % nbw;
    const BandPart& Q = PS.part[qk];                                   // left neighbour: carries the global rows over S_qk's columns
    const bool have_f = local(qk), raw = !have_f && qk == IA.q_lo - 1, have_g = local(qk + 1);
    const double* F = have_f ? Q.fb + (size_t)(Q.nrb - 1) * kNB * Q.ld + (size_t)(Q.nI + c) * kNB : nullptr;
    const int s = PS.part[qk + 1].c0 - nbw;
    const double* Rw = B.border + (size_t)(s + c) * kNB;               // this rank's own contributions to those columns (left partition elsewhere)
    const double* G = IA.gram + IA.gram_off[qk + 1];
    const int ldg = kNB * PS.part[qk + 1].nrb;
    for (int e = threadIdx.x; e < kBorder * kNB; e += 256) {
      const int g = e >> 5, cc = e & 31;
      double v = have_f ? F[(size_t)g * Q.ld + cc] : (raw ? Rw[(size_t)g * B.n_pad + cc] : 0.0);
      if (have_g) v -= G[(size_t)(nbw * kNB + g) * ldg + c * kNB + cc];
      B2.border[(size_t)g * B2.n_pad + (size_t)K2 * kNB + cc] = v;
    }
    return;
  }
  // corner
  for (int e = threadIdx.x; e < kBorder * kBorder; e += 256) {
    const int a = e / kBorder, b = e % kBorder;
    double v = B.corner[e];
    for (int p = IA.q_lo; p < IA.q_hi; ++p) {
      const int ldg = kNB * PS.part[p].nrb, o = (PS.part[p].nrb - 1) * kNB;
      v -= IA.gram[IA.gram_off[p] + (size_t)(o + (a > b ? a : b)) * ldg + o + (a > b ? b : a)];
    }
    B2.corner[e] = v;
  }
  if (threadIdx.x < 32) IA.extra[threadIdx.x] = threadIdx.x == 0 ? (double)(flag[0] != 0) : 0.0;   // a failed interior anywhere fails the solve everywhere
}
__global__ void k_flag_from_sum(const double* __restrict__ extra, int32_t* __restrict__ flag) {
  if (threadIdx.x == 0 && blockIdx.x == 0 && extra[0] != 0.0) atomicExch(flag, 1);
}
// interface solution -> separator columns of the full solution vector (+ globals)
__global__ void k_iface_scatter(BandLayout B, BandLayout B2, const __grid_constant__ BandParts PS, const double* __restrict__ delta2, double* __restrict__ delta) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < B2.n_pad) {
    const int K2 = i >> 5, q = K2 / PS.sepw, c = K2 % PS.sepw;
    const int s = PS.part[q + 1].c0 - PS.sepw;
    delta[(size_t)(s + c) * kNB + (i & 31)] = delta2[i];
  } else if (i < B2.n_pad + kBorder) {
    delta[B.n_pad + (i - B2.n_pad)] = delta2[i];
  }
}

// ---- dense border system of a single-partition solve ---------------------------------------------------------------------------
// partial[blk][a*16+b] = sum over the block's 256 columns of Lb[a][c] Lb[b][c]
__global__ void __launch_bounds__(256) k_corner_partial(BandLayout B, double* __restrict__ partial) {
  __shared__ double s[kBorder][257];
  const int c0 = blockIdx.x * 256;
  for (int e = threadIdx.x; e < kBorder * 256; e += 256) {
    const int r = e >> 8, c = e & 255;
    s[r][c] = (c0 + c < B.n_pad) ? B.border[(size_t)r * B.n_pad + c0 + c] : 0.0;
  }
  __syncthreads();
  const int a = threadIdx.x >> 4, b = threadIdx.x & 15;
  double v = 0.0;
  for (int c = 0; c < 256; ++c) v += s[a][c] * s[b][c];
  partial[(size_t)blockIdx.x * 256 + threadIdx.x] = v;
}
// G' = G - Lb Lb^T ; c' = c - Lb y ; z = G'^-1 c'   (15x15 in shared memory, 256 threads: right-looking Cholesky, the right-hand side
// rides along as a 16th row, then a serial 15-step back substitution) ; delta[n_pad + t] = z[t]
__global__ void __launch_bounds__(256) k_corner_solve(BandLayout B, int nblk, const double* __restrict__ partial, double* __restrict__ delta,
                                                      int32_t* __restrict__ flag) {
  __shared__ double M[16][17];
  const int a = threadIdx.x >> 4, b = threadIdx.x & 15;
  {
    double v = 0.0;
    for (int k = 0; k < nblk; ++k) v += partial[(size_t)k * 256 + threadIdx.x];
    // rows 0..14: G' (lower part used), row 15 = kRhsRow: c'
    M[a][b] = (b <= a || a == kRhsRow) ? B.corner[a * kBorder + b] - v : 0.0;
  }
  __syncthreads();
  const int n = GT_SIZE;
  for (int j = 0; j < n; ++j) {
    double d = M[j][j];
    const bool bad = !(d > 0.0) || !isfinite(d);
    if (bad) { if (threadIdx.x == 0) atomicExch(flag, 2); d = 1.0; }
    const double rinv = 1.0 / sqrt(d);
    const double la = (a > j) ? M[a][j] * rinv : 0.0, lb = (b > j && b < n) ? M[b][j] * rinv : 0.0;
    __syncthreads();
    if (a > j && b > j && b <= a && b < n) M[a][b] -= la * lb;      // trailing update (the right-hand-side row a = 15 included)
    if (b == j) M[a][j] = (a > j) ? la : (a == j ? sqrt(d) : M[a][j]);
    __syncthreads();
  }
  // row 15 now holds y^T = (L^-1 c')^T ; back substitution L^T z = y
  if (threadIdx.x == 0) {
    double z[GT_SIZE];
#pragma unroll
    for (int i = n - 1; i >= 0; --i) {
      double sacc = M[kRhsRow][i];
#pragma unroll
      for (int k = i + 1; k < n; ++k) sacc -= M[k][i] * z[k];
      z[i] = sacc / M[i][i];
    }
#pragma unroll
    for (int i = 0; i < n; ++i) delta[B.n_pad + i] = z[i];
    delta[B.n_pad + kRhsRow] = 0.0;
  }
}

// t = y - W^T [x_S ; z] for the interior columns of every partition (y = forward-substituted right-hand side = last border row),
// x_S = solution on the partition's left separator, z = globals.  Writes delta[interior columns].
__global__ void __launch_bounds__(256) k_back_rhs(BandLayout B, const __grid_constant__ BandParts PS, double* __restrict__ delta) {
  __shared__ double red[8][kNB];
  const int jb = blockIdx.x, lane = threadIdx.x & 31, grp = threadIdx.x >> 5;   // CTA per block column, warp = row group
  int pi = -1;
  for (int q = 0; q < PS.P; ++q) if (jb >= PS.part[q].c0 && jb < PS.part[q].c0 + PS.part[q].nI) pi = q;
  if (pi < 0) return;   // separator column: solved by the interface system
  const BandPart& Q = PS.part[pi];
  const double* F = Q.fb + (size_t)(jb - Q.c0) * kNB + lane;
  const int rs = (Q.nrb - 1) * kNB;                 // separator rows
  const double* z = delta + B.n_pad;
  const double* xs = delta + (size_t)(Q.c0 - (Q.nrb - 1)) * kNB;   // left separator; only dereferenced when rs > 0 (partition >= 1)
  double t0 = 0.0, t1 = 0.0;
  for (int r = grp; r < rs; r += 16) {
    t0 = fma(F[(size_t)r * Q.ld], xs[r], t0);
    if (r + 8 < rs) t1 = fma(F[(size_t)(r + 8) * Q.ld], xs[r + 8], t1);
  }
  for (int g = grp; g < GT_SIZE; g += 8) t0 = fma(F[(size_t)(rs + g) * Q.ld], z[g], t0);
  red[grp][lane] = t0 + t1;
  __syncthreads();
  if (grp == 0) {
    double t = F[(size_t)(rs + kRhsRow) * Q.ld];
#pragma unroll
    for (int w = 0; w < 8; ++w) t -= red[w][lane];
    delta[(size_t)jb * kNB + lane] = t;
  }
}

// x = L^-T t on the interior of every partition, block columns from last to first; partition p is handled by ONE thread-block cluster of
// kBackCtas CTAs (a single SM cannot stream the factor fast enough: every step reads nbw+1 blocks of 8 KB).  Sub-diagonal offset o (1..nbw)
// belongs to CTA (o-1) % kBackCtas, warp 1 + (o-1) / kBackCtas:
//   warp w>=1 : p_o = L(i+o, i)^T x_{i+o}     (block rows streamed with coalesced loads, the next column prefetched into registers)
//   warp 0    : sums its CTA's p_o, writes the partial into every CTA's exchange buffer through distributed shared memory;
//               after ONE cluster barrier every CTA redundantly forms  x_i = L_ii^-T (t_i - sum of partials)  and keeps its own ring of x blocks.
// The ring starts with the (already known) solution on the partition's right separator.  delta holds t on entry and x on exit.
constexpr int kBackCtas = 8;
template <int MAXT>   // 288 threads cover bands of up to 64 block diagonals; the 448-thread build (wider bands) spills a little
__global__ void __launch_bounds__(MAXT) k_band_backsolve(BandLayout B, const __grid_constant__ BandParts PS, const double* __restrict__ dinv,
                                                        double* __restrict__ delta) {
  extern __shared__ double sh[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const BandPart& Q = PS.part[blockIdx.x / kBackCtas];
  const int nb1 = B.nbw + 1;
  const int nwarps = blockDim.x >> 5;
  double* xs = sh;                               // nb1 * 32      ring of x blocks
  double* part = xs + nb1 * kNB;                 // nwarps * 32   per-warp products
  double* exch = part + nwarps * kNB;            // 2 * kBackCtas * 32
  double* ts = exch + 2 * kBackCtas * kNB;       // 32
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int off = (warp - 1) * kBackCtas + rank + 1;   // this warp's sub-diagonal offset (warp >= 1)
  const int first = Q.c0, last = Q.c0 + Q.nI - 1;
  auto src = [&](int i) -> const double* {
    if (i < first) return nullptr;
    if (warp == 0) return dinv + (size_t)i * 1024 + lane;                                   // (L_ii^-1)[c][lane]
    if (off <= B.nbw && i + off < B.nbc && i + off < Q.c0 + Q.ncols) return B.band + ((size_t)i * nb1 + off) * 1024 + lane;   // L(i+off, i)[r][lane]
    return nullptr;
  };
  for (int e = threadIdx.x; e < B.nbw * kNB; e += blockDim.x) {   // solution on the right separator (columns last+1 .. last+nbw)
    const int col = last + 1 + (e >> 5);
    if (col < B.nbc) xs[(col % nb1) * kNB + (e & 31)] = delta[(size_t)col * kNB + (e & 31)];
  }
  double cur[kNB], nxt[kNB];
  {
    const double* s0 = src(last);
#pragma unroll
    for (int r = 0; r < kNB; ++r) cur[r] = s0 ? s0[r * kNB] : 0.0;
  }
  auto dot32 = [&](const double* v) -> double {
    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
    for (int r = 0; r < kNB; r += 4) { p0 = fma(cur[r], v[r], p0); p1 = fma(cur[r + 1], v[r + 1], p1); p2 = fma(cur[r + 2], v[r + 2], p2); p3 = fma(cur[r + 3], v[r + 3], p3); }
    return (p0 + p1) + (p2 + p3);
  };
  double t_cur = (warp == 0 && last >= first) ? delta[(size_t)last * kNB + lane] : 0.0;
  cluster.sync();   // every CTA of the cluster is running before the first remote store; also orders the ring initialisation
  for (int i = last; i >= first; --i) {
    const double* g = src(i);
    const double* gn = src(i - 1);
    const double t_nxt = (warp == 0 && i > first) ? delta[(size_t)(i - 1) * kNB + lane] : 0.0;
#pragma unroll
    for (int r = 0; r < kNB; ++r) nxt[r] = gn ? gn[r * kNB] : 0.0;
    if (warp > 0) part[warp * kNB + lane] = g ? dot32(xs + ((i + off) % nb1) * kNB) : 0.0;
    __syncthreads();
    double* ex = exch + (i & 1) * kBackCtas * kNB;
    if (warp == 0) {
      double s = 0.0;
      for (int w = 1; w < nwarps; ++w) s += part[w * kNB + lane];
#pragma unroll
      for (int r = 0; r < kBackCtas; ++r) cluster.map_shared_rank(ex, r)[rank * kNB + lane] = s;
    }
    cluster.sync();
    if (warp == 0) {
      double t = t_cur;
#pragma unroll
      for (int r = 0; r < kBackCtas; ++r) t -= ex[r * kNB + lane];
      ts[lane] = t;
      __syncwarp();
      const double x = dot32(ts);
      xs[(i % nb1) * kNB + lane] = x;
      if (rank == 0) delta[(size_t)i * kNB + lane] = x;
    }
    __syncthreads();
    t_cur = t_nxt;
#pragma unroll
    for (int r = 0; r < kNB; ++r) cur[r] = nxt[r];
  }
  cluster.sync();   // no CTA exits while a peer may still address its shared memory
}

// ================================================== host side =====================================================
// One level of the nested solve.  Level 0 is the reduced system itself (Hw); level l + 1 is the interface system of level l (its separators:
// block tridiagonal with sepw-wide dense blocks, stored as a band of half width 2 sepw - 1).  Every level with P > 1 eliminates its
// interiors concurrently, one block column per launch; at the interface levels an interior is ONE separator block of the level below
// (odd-even reduction), so the sequential depth is  max|I_q| + sepw * ceil(log2 P0)  block steps instead of  max|I_q| + (P0 - 1) sepw.
struct BandLevel {
  BandLayout B{};           // the system of this level
  int P = 1, max_steps = 0;
  BandParts PS{};           // every partition of this level (P == 1: one partition whose border is B.border itself)
  BandParts PSloc{};        // the partitions this handle factors (all of them, except at level 0 of a problem sharded over ranks)
  GramPlan GP{};
  IfaceArgs IA{};
  int gram_tasks = 0, gram_tiles = 0, gram_stasks = 0;
  size_t sys_elems = 0;     // next level's band | border | corner | extra, ONE buffer (a sharded solve all-reduces level 0's)
  DBuf<double> fb, gram, gram_scratch, sys, delta_own, dinv, cpart;
  double* delta = nullptr;  // solution of this level (level 0: the solver's delta; otherwise delta_own)
  size_t* d_gram_off = nullptr;
  ~BandLevel() {
    DBuf<double>* dd[] = {&fb, &gram, &gram_scratch, &sys, &delta_own, &dinv, &cpart};
    for (auto* d : dd) d->release();
    if (d_gram_off) cudaFree(d_gram_off);
  }
};
struct BandPlan {
  // key
  int nbc = -1, nbw = -1, req = -2; const double* band_ptr = nullptr; const double* border_ptr = nullptr;
  std::vector<BandLevel*> lv;
  void clear() { for (auto* l : lv) delete l; lv.clear(); }
  ~BandPlan() { clear(); }
};

// sequential block steps of the nested solve with P partitions at level 0
static int nested_depth(int nbc, int nbw, int P) {
  if (P <= 1) return nbc;
  int d = (nbc - (P - 1) * nbw + P - 1) / P + 1, m = P - 1;   // largest interior + its flush step
  while (m > 2) { d += nbw + 1; m = (m + 1) / 2 - 1; }          // one separator block per interface level
  return d + m * nbw;                                           // last level: one partition of m blocks
}

int band_partition_layout(int nbc, int nbw, int requested, int world, int nested, PartLayout* out) {
  int P = 1;
  const auto fits = [&](int q) { return q >= 1 && q <= kMaxParts && (nbc - (q - 1) * nbw) / q >= std::max(nbw, 2); };
  const bool can_split = !(nbw < 1 || 2 * nbw > 63);
  if (can_split) {
    if (requested > 0) {
      P = std::max(1, std::min(requested, kMaxParts));
      if (world > 1) P = std::max(world, (P / world) * world);
      while (P > world && !fits(P)) P -= world;
    } else if (!nested) {   // flat interface system: balance interior length against the sequential interface solve
      P = (int)lround(std::sqrt((double)nbc / (1.2 * nbw)));
      // wide bands are throughput-bound, not latency-bound: one partition-step is ~8 ns * nbw^2 of DMMA work against a ~8.5 us
      // launch-chain floor, and beyond ~2.5x that many concurrent partitions more partitions only add interface steps
      // (measured on config 3, nbw = 21, solve phase: P = 6 -> 24.5, 8 -> 24.1, 9 -> 23.9, 10 -> 24.0, 12 -> 24.2, 14 -> 24.6, 17 -> 25.3, 20 -> 25.8 ms)
      if (world == 1) P = std::min(P, std::max(2, (int)lround(4000.0 / ((double)nbw * nbw))));
      P = std::max(1, std::min(P, kMaxParts));
      if (world > 1) P = std::max(world, (P / world) * world);
      while (P > (world > 1 ? world : 1) && !fits(P)) P -= (world > 1 ? world : 1);
    } else {   // nested interface levels: the P that minimises the sequential depth (ties: fewer partitions)
      int best = nested_depth(nbc, nbw, 1);
      for (int q = world > 1 ? world : 2; q <= kMaxParts; q += (world > 1 ? world : 1)) {
        if (!fits(q)) break;
        const int d = nested_depth(nbc, nbw, q);
        if (d < best || (world > 1 && P < world)) { best = d; P = q; }
      }
    }
  }
  if (world > 1 && (P < world || !can_split || !fits(P))) return -1;
  if (world == 1 && P > 1 && !fits(P)) P = 1;
  out->P = P; out->c0.assign(P, 0); out->nI.assign(P, nbc);
  if (P == 1) return 0;
  const int interior = nbc - (P - 1) * nbw;
  int c = 0;
  for (int q = 0; q < P; ++q) {
    out->c0[q] = c; out->nI[q] = interior / P + (q < interior % P ? 1 : 0);
    c += out->nI[q] + nbw;
  }
  return 0;
}

// partitions of an interface level: m separator blocks of sepw columns; blocks 0, 2, 4, ... are interiors, 1, 3, 5, ... separators;
// the last interior absorbs a trailing block so that the level ends with an interior
static void interface_layout(int m, int sepw, PartLayout* out) {
  const int P = m <= 2 ? 1 : (m + 1) / 2;
  out->P = P; out->c0.assign(P, 0); out->nI.assign(P, m * sepw);
  if (P == 1) return;
  for (int q = 0; q < P; ++q) {
    out->c0[q] = 2 * q * sepw;
    out->nI[q] = (q == P - 1) ? (m - 2 * q) * sepw : sepw;
  }
}

// builds level `li` (its system BL.B must be set) and, when it has more than one partition, allocates the next level's system
static int level_build(ctba_problem* p, BandPlan& pl, int li, const PartLayout& lay, int sepw, int q_lo, int q_hi) {
  BandLevel& L = *pl.lv[li];
  const BandLayout& B = L.B;
  L.P = lay.P;
  CTBA_CUDA(p, L.dinv.resize((size_t)B.nbc * 1024));
  CTBA_CUDA(p, L.cpart.resize((size_t)cdiv(B.n_pad, 256) * 256));
  L.PS = BandParts{};
  if (L.P == 1) {
    L.PS.P = 1; L.PS.sepw = 0; L.PS.defer = 0;
    L.PS.part[0] = BandPart{0, B.nbc, 1, B.nbc, B.n_pad, B.border};
    L.PSloc = L.PS;
    L.max_steps = B.nbc;
    return CTBA_OK;
  }
  const int P = L.P;
  size_t fb_total = 0, gram_total = 0;
  std::vector<size_t> fb_off(P, 0), g_off(P, 0), g_off_loc;
  int maxI = 0;
  L.PS.P = P; L.PS.sepw = sepw; L.PS.defer = 1;
  for (int q = 0; q < P; ++q) {
    const int nI = lay.nI[q];
    BandPart& Q = L.PS.part[q];
    Q.c0 = lay.c0[q]; Q.nI = nI; Q.nrb = (q == 0 ? 1 : sepw + 1); Q.ncols = nI + (q < P - 1 ? sepw : 0); Q.ld = kNB * Q.ncols; Q.fb = nullptr;
    if (q < q_lo || q >= q_hi) continue;
    fb_off[q] = fb_total; fb_total += (size_t)Q.nrb * kNB * Q.ld;
    g_off[q] = gram_total; gram_total += (size_t)(Q.nrb * kNB) * (Q.nrb * kNB);
    g_off_loc.push_back(g_off[q]);
    maxI = std::max(maxI, nI);
  }
  L.max_steps = maxI + 1;   // + the flush step that applies the last interior column's update to the right separator
  CTBA_CUDA(p, L.fb.resize(fb_total)); CTBA_CUDA(p, L.gram.resize(gram_total));
  L.GP = GramPlan{}; L.GP.P = q_hi - q_lo;
  L.PSloc = BandParts{}; L.PSloc.P = q_hi - q_lo; L.PSloc.sepw = sepw; L.PSloc.defer = 1;
  L.IA = IfaceArgs{};
  int tiles = 0, tasks = 0, stasks = 0;
  for (int q = q_lo; q < q_hi; ++q) {
    BandPart& Q = L.PS.part[q];
    Q.fb = L.fb.p + fb_off[q];
    L.PSloc.part[q - q_lo] = Q;
    L.IA.gram_off[q] = g_off[q];
    const int nt = Q.nrb * (Q.nrb + 1) / 2, ns = (Q.nI + kGramChunk - 1) / kGramChunk;
    L.GP.tile0[q - q_lo] = tiles; L.GP.task0[q - q_lo] = tasks; L.GP.nsplit[q - q_lo] = ns; L.GP.stask0[q - q_lo] = stasks;
    const int nst = (Q.nrb + 1) / 2;
    tiles += nt; tasks += nt * ns; stasks += nst * (nst + 1) / 2 * ns;
  }
  L.GP.tile0[q_hi - q_lo] = tiles; L.GP.task0[q_hi - q_lo] = tasks; L.GP.stask0[q_hi - q_lo] = stasks;
  L.gram_tiles = tiles; L.gram_tasks = tasks; L.gram_stasks = stasks;
  L.IA.gram = L.gram.p; L.IA.q_lo = q_lo; L.IA.q_hi = q_hi;
  CTBA_CUDA(p, L.gram_scratch.resize((size_t)tasks * 1024));
  if (!L.d_gram_off) CTBA_CUDA(p, cudaMalloc((void**)&L.d_gram_off, kMaxParts * sizeof(size_t)));
  CTBA_CUDA(p, cudaMemcpy(L.d_gram_off, g_off_loc.data(), g_off_loc.size() * sizeof(size_t), cudaMemcpyHostToDevice));
  // the next level: band | border | corner | extra in ONE buffer
  BandLayout B2{};
  B2.nbc = (P - 1) * sepw; B2.n_pad = B2.nbc * kNB; B2.n_band = B2.n_pad; B2.nbw = std::min(2 * sepw - 1, B2.nbc - 1);
  const size_t e_band = B2.band_elems(), e_border = (size_t)kBorder * B2.n_pad;
  L.sys_elems = e_band + e_border + kBorder * kBorder + 32;
  CTBA_CUDA(p, L.sys.resize(L.sys_elems));
  B2.band = L.sys.p; B2.border = L.sys.p + e_band; B2.corner = B2.border + e_border;
  L.IA.extra = B2.corner + kBorder * kBorder;
  BandLevel* N = new BandLevel();
  N->B = B2;
  CTBA_CUDA(p, N->delta_own.resize(B2.n_pad + kBorder));
  CTBA_CUDA(p, cudaMemset(N->delta_own.p, 0, (B2.n_pad + kBorder) * sizeof(double)));
  N->delta = N->delta_own.p;
  pl.lv.push_back(N);
  return CTBA_OK;
}

static int plan_build(ctba_problem* p, BandPlan& pl, const BandLayout& B) {
  pl.clear();
  pl.nbc = B.nbc; pl.nbw = B.nbw; pl.req = p->band_partitions + 1000 * p->band_nested; pl.band_ptr = B.band; pl.border_ptr = B.border;
  PartLayout lay;
  int q_lo = 0, q_hi = 1;
  if (p->sharded) {
    lay = p->shard->layout; q_lo = p->shard->q_lo; q_hi = p->shard->q_hi;
    if (p->shard->nbc != B.nbc || p->shard->nbw != B.nbw) { p->err = "band plan: structure differs from the sharding plan"; return CTBA_E_STATE; }
  } else {
    band_partition_layout(B.nbc, B.nbw, p->band_partitions, 1, p->band_nested, &lay);
    q_hi = lay.P;
  }
  BandLevel* L0 = new BandLevel();
  L0->B = B; L0->delta = p->solver->delta.p;
  pl.lv.push_back(L0);
  int rc = level_build(p, pl, 0, lay, B.nbw, q_lo, q_hi); if (rc) return rc;
  const int sepw = B.nbw;
  for (size_t li = 1; li < pl.lv.size(); ++li) {   // level_build appends the next level while the current one still has partitions
    PartLayout l2;
    interface_layout(pl.lv[li]->B.nbc / sepw, sepw, &l2);
    if (p->band_nested == 0) { l2.P = 1; l2.c0.assign(1, 0); l2.nI.assign(1, pl.lv[li]->B.nbc); }   // flat: the interface system as one partition
    rc = level_build(p, pl, (int)li, l2, sepw, 0, l2.P); if (rc) return rc;
  }
  return CTBA_OK;
}

// Step launches on three streams (inside a capture: three graph branches): the panel kernel on `st`, the update kernel on a side stream
// (each kernel of step t waits for both kernels of step t-1) and, on levels with partitions, the border pipeline on a third stream: one
// launch per chunk of steps, which waits for the chunk's last panel step and for the previous chunk.  After the last step and the last
// chunk one launch carries the border rows over the right separators.
constexpr int kBorderChunk = 8;
static int enqueue_steps(ctba_problem* p, cudaStream_t st, const BandLayout& B, const BandParts& PS, int nsteps, double* dinv, int64_t* nl) {
  SolverData* sd = p->solver;
  cudaStream_t s2 = (cudaStream_t)sd->band_stream2, s3 = (cudaStream_t)sd->band_stream3;
  cudaEvent_t evA = (cudaEvent_t)sd->band_evA, evB = (cudaEvent_t)sd->band_evB, evC = (cudaEvent_t)sd->band_evC;
  CTBA_CUDA(p, cudaEventRecord(evA, st));
  CTBA_CUDA(p, cudaStreamWaitEvent(s2, evA, 0));       // fork
  CTBA_CUDA(p, cudaStreamWaitEvent(s3, evA, 0));
  CTBA_CUDA(p, cudaEventRecord(evB, s2));              // the joins below are valid even when a side stream gets no work
  CTBA_CUDA(p, cudaEventRecord(evC, s3));
  int nborder = 0, nsep = 0, max_panel = 0;
  if (PS.defer) for (int q = 0; q < PS.P; ++q) { nborder += PS.part[q].nrb; nsep += (PS.part[q].ncols - PS.part[q].nI) * PS.part[q].nrb; max_panel = std::max(max_panel, PS.part[q].nI); }
  int chunk_begin = 0;
  for (int t = 0; t < nsteps; ++t) {
    int gp = 0, gu = 0;
    for (int q = 0; q < PS.P; ++q) {
      int np, m, kl;
      part_tasks(B, PS.part[q], t, PS.defer, &np, &m, &kl);
      gp += np; gu += part_update_tasks(m, PS.defer ? 0 : PS.part[q].nrb);
    }
    if (gp || gu) {
      if (t > 0) { CTBA_CUDA(p, cudaStreamWaitEvent(st, evB, 0)); CTBA_CUDA(p, cudaStreamWaitEvent(s2, evA, 0)); }
      if (gp) { k_band_panel<<<gp, kStepThreads, 0, st>>>(B, PS, t, dinv, sd->flag.p); ++*nl; }
      if (gu) { k_band_update<<<gu, kStepThreads, 0, s2>>>(B, PS, t); ++*nl; }
      CTBA_CUDA(p, cudaEventRecord(evA, st));
      CTBA_CUDA(p, cudaEventRecord(evB, s2));
    }
    if (nborder && t < max_panel && (t + 1 - chunk_begin == kBorderChunk || t + 1 == max_panel)) {
      CTBA_CUDA(p, cudaStreamWaitEvent(s3, evA, 0));   // evA: the most recent panel step = the chunk's last
      CTBA_CUDA(p, cudaFuncSetAttribute(k_band_border, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kBorderSmem));
      k_band_border<<<nborder, kBorderThreads, kBorderSmem, s3>>>(B, PS, chunk_begin, t + 1, dinv); ++*nl;
      CTBA_CUDA(p, cudaEventRecord(evC, s3));
      chunk_begin = t + 1;
    }
  }
  CTBA_CUDA(p, cudaStreamWaitEvent(st, evB, 0));       // join
  CTBA_CUDA(p, cudaStreamWaitEvent(st, evC, 0));
  if (nsep) { k_band_border_sep<<<nsep, kBorderThreads, 0, st>>>(B, PS); ++*nl; }
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

static int enqueue_backsolve(ctba_problem* p, cudaStream_t st, const BandLayout& B, const BandParts& PS, const double* dinv, double* delta, int64_t* nl) {
  k_back_rhs<<<B.nbc, 256, 0, st>>>(B, PS, delta);
  const int nb1 = B.nbw + 1;
  const int nwarps = 1 + std::max(1, (B.nbw + kBackCtas - 1) / kBackCtas);   // <= 9 for nbw <= 63
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(kBackCtas * PS.P); cfg.blockDim = dim3(32 * nwarps); cfg.stream = st;
  cfg.dynamicSmemBytes = (size_t)(nb1 + nwarps + 2 * kBackCtas + 1) * kNB * sizeof(double);
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = kBackCtas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr; cfg.numAttrs = 1;
  if (32 * nwarps <= 288) CTBA_CUDA(p, cudaLaunchKernelEx(&cfg, k_band_backsolve<288>, B, PS, dinv, delta));
  else CTBA_CUDA(p, cudaLaunchKernelEx(&cfg, k_band_backsolve<448>, B, PS, dinv, delta));
  *nl += 2;
  return CTBA_OK;
}

// factor + solve of a system that is ONE partition (the whole system when P == 1, the interface system otherwise)
static int enqueue_single(ctba_problem* p, cudaStream_t st, const BandLayout& B, const BandParts& PS1, double* dinv, double* delta, double* cpart, int64_t* nl) {
  SolverData* sd = p->solver;
  int rc = enqueue_steps(p, st, B, PS1, B.nbc, dinv, nl); if (rc) return rc;
  const int nblk = cdiv(B.n_pad, 256);
  k_corner_partial<<<nblk, 256, 0, st>>>(B, cpart);
  k_corner_solve<<<1, 256, 0, st>>>(B, nblk, cpart, delta, sd->flag.p);
  *nl += 2;
  return enqueue_backsolve(p, st, B, PS1, dinv, delta, nl);
}

// phase 0: the whole solve (one handle holds every partition); phases 1 / 2: the two halves of a sharded solve — level-0 interiors + this
// rank's terms of the level-1 system | (all-reduce of that system in between) | remaining levels (replicated) + all back-substitutions
static int band_enqueue(ctba_problem* p, cudaStream_t st, BandPlan& pl, int phase, int64_t* nl) {
  SolverData* sd = p->solver;
  int rc;
  if (phase <= 1) CTBA_CUDA(p, cudaMemsetAsync(sd->flag.p, 0, 4 * sizeof(int32_t), st));
  const int nlev = (int)pl.lv.size();
  // ---- downward: eliminate the interiors of every level that has partitions, assemble the next level
  for (int li = (phase == 2 ? 1 : 0); li < nlev - 1; ++li) {
    BandLevel& L = *pl.lv[li];
    const BandLayout& B2 = pl.lv[li + 1]->B;
    if (phase == 2 && li == 1) { k_flag_from_sum<<<1, 32, 0, st>>>(pl.lv[0]->IA.extra, sd->flag.p); ++*nl; }
    int nblk = 0;
    for (int q = 0; q < L.PSloc.P; ++q) nblk += L.PSloc.part[q].nrb * L.PSloc.part[q].ncols;
    k_fb_init<<<nblk, 256, 0, st>>>(L.B, L.PSloc);
    ++*nl;
    rc = enqueue_steps(p, st, L.B, L.PSloc, L.max_steps, L.dinv.p, nl); if (rc) return rc;
    k_gram_partial<<<L.gram_stasks, kStepThreads, 0, st>>>(L.PSloc, L.GP, L.gram_scratch.p);
    k_gram_reduce<<<L.gram_tiles, 256, 0, st>>>(L.PSloc, L.GP, L.gram_scratch.p, L.gram.p, L.d_gram_off);
    k_iface_assemble<<<B2.nbc * (B2.nbw + 1) + B2.nbc + 1, 256, 0, st>>>(L.B, B2, L.PS, L.IA, sd->flag.p);
    *nl += 3;
    if (phase == 1) break;   // the level-1 system is summed over the ranks before anything else happens
  }
  if (phase == 1) { CTBA_CUDA(p, cudaGetLastError()); return CTBA_OK; }
  if (phase == 2 && nlev == 2) { k_flag_from_sum<<<1, 32, 0, st>>>(pl.lv[0]->IA.extra, sd->flag.p); ++*nl; }
  // ---- last level: one partition, factor + solve
  {
    BandLevel& L = *pl.lv[nlev - 1];
    rc = enqueue_single(p, st, L.B, L.PS, L.dinv.p, L.delta, L.cpart.p, nl); if (rc) return rc;
  }
  // ---- upward: separator solutions -> level below, interior back-substitution
  for (int li = nlev - 2; li >= 0; --li) {
    BandLevel& L = *pl.lv[li];
    const BandLayout& B2 = pl.lv[li + 1]->B;
    k_iface_scatter<<<cdiv(B2.n_pad + kBorder, 256), 256, 0, st>>>(L.B, B2, L.PS, pl.lv[li + 1]->delta, L.delta);
    ++*nl;
    rc = enqueue_backsolve(p, st, L.B, L.PSloc, L.dinv.p, L.delta, nl); if (rc) return rc;
  }
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

void band_graph_release(SolverData* sd) {
  for (void** g : {&sd->band_graph_exec, &sd->band_graph_exec1, &sd->band_graph_exec2})
    if (*g) { cudaGraphExecDestroy((cudaGraphExec_t)*g); *g = nullptr; }
}
void band_plan_release(SolverData* sd) {
  band_graph_release(sd);
  if (sd->band_plan) { delete (BandPlan*)sd->band_plan; sd->band_plan = nullptr; }
  if (sd->band_stream2) {
    cudaStreamDestroy((cudaStream_t)sd->band_stream2); cudaEventDestroy((cudaEvent_t)sd->band_evA); cudaEventDestroy((cudaEvent_t)sd->band_evB);
    cudaStreamDestroy((cudaStream_t)sd->band_stream3); cudaEventDestroy((cudaEvent_t)sd->band_evC);
    sd->band_stream2 = nullptr; sd->band_stream3 = nullptr;
  }
}

// Factor + solve.  The launch sequence depends only on the band structure and the buffer addresses, so it is captured once into a
// CUDA graph (cheaper kernel-to-kernel hand-over than hundreds of stream launches) and replayed every LM iteration.
int band_factor_solve(ctba_problem* p) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  BandLayout& B = sd->Hw;
  if (B.nbw > 103) { p->err = "band_factor_solve: band wider than 103 blocks of 32 scalars (use linear_solver = 1)"; return CTBA_E_ARG; }
  if (!sd->band_plan) sd->band_plan = new BandPlan();
  if (!sd->band_stream2) {
    cudaStream_t s2, s3; cudaEvent_t ea, eb, ec;
    CTBA_CUDA(p, cudaStreamCreateWithFlags(&s2, cudaStreamNonBlocking));
    CTBA_CUDA(p, cudaStreamCreateWithFlags(&s3, cudaStreamNonBlocking));
    CTBA_CUDA(p, cudaEventCreateWithFlags(&ea, cudaEventDisableTiming));
    CTBA_CUDA(p, cudaEventCreateWithFlags(&eb, cudaEventDisableTiming));
    CTBA_CUDA(p, cudaEventCreateWithFlags(&ec, cudaEventDisableTiming));
    sd->band_stream2 = s2; sd->band_stream3 = s3; sd->band_evA = ea; sd->band_evB = eb; sd->band_evC = ec;
  }
  BandPlan& pl = *(BandPlan*)sd->band_plan;
  if (pl.nbc != B.nbc || pl.nbw != B.nbw || pl.req != p->band_partitions + 1000 * p->band_nested || pl.band_ptr != B.band || pl.border_ptr != B.border) {
    band_graph_release(sd);
    const int rc = plan_build(p, pl, B); if (rc) return rc;
  }
  // run one phase: eagerly, or as a CUDA graph captured on first use
  auto run_phase = [&](int phase, void** exec_slot, int64_t* launches_slot) -> int {
    int64_t nl = 0;
    if (!p->band_graph) { const int rc = band_enqueue(p, st, pl, phase, &nl); p->launches += nl; *launches_slot = nl; return rc; }
    if (!*exec_slot) {
      cudaGraph_t graph = nullptr;
      CTBA_CUDA(p, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
      const int rc = band_enqueue(p, st, pl, phase, &nl);
      const cudaError_t e = cudaStreamEndCapture(st, &graph);
      if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
      CTBA_CUDA(p, e);
      cudaGraphExec_t exec = nullptr;
      const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
      cudaGraphDestroy(graph);
      CTBA_CUDA(p, ei);
      *exec_slot = exec; *launches_slot = nl;
    }
    p->launches += *launches_slot;
    CTBA_CUDA(p, cudaGraphLaunch((cudaGraphExec_t)*exec_slot, st));
    return CTBA_OK;
  };
  const int key = 16 * p->band_nested;
  if (sd->band_graph_key != key) band_graph_release(sd);
  sd->band_graph_key = key;
  if (!p->sharded) return run_phase(0, &sd->band_graph_exec, &sd->band_launches);
  int rc = run_phase(1, &sd->band_graph_exec1, &sd->band_launches1); if (rc) return rc;
  if (p->comm->allreduce_sum(pl.lv[0]->sys.p, pl.lv[0]->sys_elems, st)) { p->err = p->comm->error(); return CTBA_E_NCCL; }
  rc = run_phase(2, &sd->band_graph_exec2, &sd->band_launches2);
  sd->band_launches = sd->band_launches1 + sd->band_launches2;
  return rc;
}

}  // namespace ctba

#ifdef CTBA_BAND_TIMING
extern "C" int ctba_debug_border_clocks(long long* out64) {
  return cudaMemcpyFromSymbol(out64, ctba::g_border_clk, sizeof(long long) * 64) == cudaSuccess ? 0 : -2;
}
extern "C" int ctba_debug_band_clocks(long long* out8) {
  return cudaMemcpyFromSymbol(out8, ctba::g_band_clk, sizeof(long long) * 8) == cudaSuccess ? 0 : -2;
}
#endif
#ifdef CTBA_BAND_TRACE
// reset != 0: arm the trace (starts = max, ends = 0); otherwise copy it out: 3 x 8192 x 2 unsigned 64-bit nanosecond stamps
extern "C" int ctba_debug_band_trace(unsigned long long* out, int reset) {
  const size_t n = (size_t)3 * ctba::kTraceSteps * 2;
  if (reset) {
    std::vector<unsigned long long> h(n);
    for (size_t i = 0; i < n; ++i) h[i] = (i & 1) ? 0ull : ~0ull;
    return cudaMemcpyToSymbol(ctba::g_band_trace, h.data(), n * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
  }
  cudaDeviceSynchronize();
  return cudaMemcpyFromSymbol(out, ctba::g_band_trace, n * sizeof(unsigned long long)) == cudaSuccess ? 0 : -2;
}
#endif
