// ctba_band.cu — bordered block-banded Cholesky factorisation + substitutions on the device (sm_100a, FP64).
//
// Solves the reduced (landmark-eliminated) LM system  [H_band  Bd^T; Bd  G] [x; z] = [y; c]  of ctba_solver.h exactly.
// This is the linear solve that Ceres' SPARSE_NORMAL_CHOLESKY performs inside ceres::Solve for the reference
// (include/spline/bspline_so3r3.h:1771, 1778) — here on the block band that the spline's locality produces.
//
//   k_band_step       one launch per 32-column block, right-looking with a one-column look-ahead; every 32x32x32 product
//                     runs on the FP64 tensor pipe (mma.sync m8n8k4 f64 -> DMMA)
//   k_corner_partial / k_corner_solve   Schur complement of the band onto the <= 15 global parameters, 15x15 solve
//   k_back_rhs + k_band_backsolve       backward substitution  L^T x = y - Lb^T z
#include <cooperative_groups.h>
#include "ctba_solver.h"

namespace ctba {
namespace cg = cooperative_groups;

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

// ---- FP64 tensor-core tile product -------------------------------------------------------------------------------
// mma.sync m8n8k4 f64 (SASS: DMMA).  Fragment ownership (lane = 4*g + t): A[g][t], B[t][g], C[g][2t], C[g][2t+1].
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// One 8x8 tile of  P = A * Bm^T  over K = 32:  A rows [8*tr, 8*tr+8) (row stride lda), Bm rows [8*tc, 8*tc+8) (row stride ldb).
// Operands are read straight from global memory / L2: per k-step a warp touches 8 rows x one 32 B sector; every sector is used once.
__device__ __forceinline__ void tile_abt(const double* __restrict__ A, int lda, const double* __restrict__ Bm, int ldb, int tr, int tc, int lane,
                                         double& c0, double& c1) {
  const int g = lane >> 2, t = lane & 3;
  const double* pa = A + (size_t)(8 * tr + g) * lda + t;
  const double* pb = Bm + (size_t)(8 * tc + g) * ldb + t;
  double a[8], b[8];
#pragma unroll
  for (int ks = 0; ks < 8; ++ks) { a[ks] = pa[4 * ks]; b[ks] = pb[4 * ks]; }
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
    const double* pb0 = Ld + (size_t)(8 * tc0 + g) * kNB + t;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) { b0[ks] = pb0[4 * ks]; b1[ks] = pb0[8 * kNB + 4 * ks]; }
  }
  if (sub_d) {
    const double* pa = Ld + (size_t)r * kNB + t;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) ad[ks] = pa[4 * ks];
  }
  if (sub_x) {
    const double* pa = Ax + (size_t)r * ldax + t;
#pragma unroll
    for (int ks = 0; ks < 8; ++ks) ax[ks] = pa[4 * ks];
  }
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

#ifdef CTBA_BAND_TIMING
__device__ long long g_band_clk[8];
#define BAND_CLK(k) do { if (j == 100 && b == 1 && tid == 0) g_band_clk[k] = clock64(); } while (0)
#else
#define BAND_CLK(k) do {} while (0)
#endif

// One launch per block column j (right-looking with a one-column look-ahead):
//   panel CTAs  [0, rows+2)      : apply column j-1's update to THEIR block of column j (and to the diagonal block, redundantly),
//                                  factor the diagonal block, triangular-solve their block, store it.  CTA 0 = diagonal (also stores
//                                  L_jj^-1 for the backward substitution), CTA rows+1 = the 16 border rows (15 globals + right-hand side).
//   update CTAs [rows+2, ...)    : remaining trailing update of column j-1:  T(i,k) -= L(i,j-1) L(k,j-1)^T  for j+1 <= k <= i <= j-1+nbw,
//                                  and the border rows of those block columns.
// Pre-condition of launch j: columns < j final, every trailing block carries the updates of columns <= j-2.
__global__ void __launch_bounds__(kStepThreads) k_band_step(BandLayout B, int j, double* __restrict__ dinv, int32_t* __restrict__ flag, int pdl) {
  __shared__ double sD[kNB * kLd];
  __shared__ double sX[kNB * kLd];
  __shared__ double sRinv[kNB];
  // programmatic dependent launch (pdl != 0): column j+1's grid may become resident early (pdl 1: now; pdl 2: once this CTA has staged its
  // operands); it waits here until column j-1's grid has completed and flushed.  Without the launch attribute both are no-ops.
  if (pdl == 1) asm volatile("griddepcontrol.launch_dependents;");
  if (pdl) asm volatile("griddepcontrol.wait;" ::: "memory");
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nb1 = B.nbw + 1;
  const int rows = min(B.nbw, B.nbc - 1 - j);
  const int b = blockIdx.x;
  const bool prev = j > 0 && B.nbw >= 1;
  const double* Lprev = B.band + ((size_t)max(j - 1, 0) * nb1) * 1024;  // column j-1: block (j-1+o, j-1) at offset o
  const double* Bprev = B.border + (size_t)max(j - 1, 0) * kNB;
  if (b < rows + 2) {
    // ---------------- panel ----------------
    double* Ajj = B.band + ((size_t)j * nb1) * 1024;
    const bool is_border = (b == rows + 1);
    BAND_CLK(0);
    if (b == 0) {
      panel_stage(sD, sX, Ajj, Lprev + 1024, prev, false, nullptr, 0, nullptr, 0, 0, false, warp, lane);
      // identity right-hand side: X L^T = I gives X = L_jj^-T, whose transpose L_jj^-1 is kept for the backward substitution
      for (int e = tid; e < kNB * kNB; e += kStepThreads) sX[(e >> 5) * kLd + (e & 31)] = ((e >> 5) == (e & 31)) ? 1.0 : 0.0;
    } else if (!is_border) {
      const bool have_prev = prev && (b + 1 <= B.nbw);   // block (j+b, j-1) lies inside the band
      panel_stage(sD, sX, Ajj, Lprev + 1024, prev, true, Ajj + (size_t)b * 1024, kNB, Lprev + (size_t)min(b + 1, B.nbw) * 1024, kNB, kNB, have_prev,
                  warp, lane);
    } else {
      panel_stage(sD, sX, Ajj, Lprev + 1024, prev, true, B.border + (size_t)j * kNB, B.n_pad, Bprev, B.n_pad, kBorder, prev, warp, lane);
    }
    __syncthreads();
    if (pdl == 2) asm volatile("griddepcontrol.launch_dependents;");
    BAND_CLK(1);
    potrf_trsm32_cta(sD, sX, sRinv, warp, lane, flag);
    BAND_CLK(2);
    __syncthreads();
    BAND_CLK(3);
    if (b == 0) {
      for (int e = tid; e < kNB * kNB; e += kStepThreads) {
        const int r = e >> 5, c = e & 31;
        Ajj[e] = sD[r * kLd + c];
        dinv[(size_t)j * 1024 + e] = sX[c * kLd + r];   // (L^-1)[r][c] = X[c][r]
      }
    } else if (!is_border) {
      double* Lij = Ajj + (size_t)b * 1024;
      for (int e = tid; e < kNB * kNB; e += kStepThreads) Lij[e] = sX[(e >> 5) * kLd + (e & 31)];
      BAND_CLK(4);
    } else {
      for (int e = tid; e < kBorder * kNB; e += kStepThreads) { const int r = e >> 5, c = e & 31; B.border[(size_t)r * B.n_pad + j * kNB + c] = sX[r * kLd + c]; }
    }
    return;
  }
  // ---------------- trailing update of column j-1 (only launched for j >= 1) ----------------
  if (pdl == 2) asm volatile("griddepcontrol.launch_dependents;");
  const int m = min(B.nbw - 1, B.nbc - 1 - j);   // rows j+1 .. j+m carry a block in column j-1
  int u = b - (rows + 2);
  const int ntri = m * (m + 1) / 2;
  const int g = lane >> 2, t = lane & 3;
  if (u < ntri) {
    int a = (int)((sqrtf(8.0f * u + 1.0f) - 1.0f) * 0.5f);
    while ((a + 1) * (a + 2) / 2 <= u) ++a;
    while (a * (a + 1) / 2 > u) --a;
    const int c = u - a * (a + 1) / 2;          // 0 <= c <= a < m
    const int i = j + 1 + a, k = j + 1 + c;
    const double* Li = Lprev + (size_t)(i - (j - 1)) * 1024;
    const double* Lk = Lprev + (size_t)(k - (j - 1)) * 1024;
    double* T = B.band + ((size_t)k * nb1 + (i - k)) * 1024;
    const int tr = warp >> 1;
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int tc = 2 * (warp & 1) + h;
      double c0 = 0.0, c1 = 0.0;
      tile_abt(Li, kNB, Lk, kNB, tr, tc, lane, c0, c1);
      double2* dst = reinterpret_cast<double2*>(T + (size_t)(8 * tr + g) * kNB + 8 * tc + 2 * t);
      double2 v = *dst; v.x -= c0; v.y -= c1; *dst = v;
    }
    return;
  }
  u -= ntri;
  if (u < m) {   // border rows of block column k
    const int k = j + 1 + u;
    const double* Lk = Lprev + (size_t)(k - (j - 1)) * 1024;
    const int tr = warp >> 2, tc = warp & 3;     // 2 x 4 tiles of the 16 x 32 border block
    double c0 = 0.0, c1 = 0.0;
    tile_abt(Bprev, B.n_pad, Lk, kNB, tr, tc, lane, c0, c1);
    double2* dst = reinterpret_cast<double2*>(B.border + (size_t)(8 * tr + g) * B.n_pad + (size_t)k * kNB + 8 * tc + 2 * t);
    double2 v = *dst; v.x -= c0; v.y -= c1; *dst = v;
  }
}

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
// G' = G - Lb Lb^T ; c' = c - Lb y ; z = G'^-1 c'   (15x15, one thread) ; delta[n_pad + t] = z[t]
__global__ void k_corner_solve(BandLayout B, int nblk, const double* __restrict__ partial, double* __restrict__ delta, int32_t* __restrict__ flag) {
  __shared__ double Msum[256];
  {
    double v = 0.0;
    for (int k = 0; k < nblk; ++k) v += partial[(size_t)k * 256 + threadIdx.x];
    Msum[threadIdx.x] = v;
  }
  __syncthreads();
  if (threadIdx.x != 0) return;
  const int n = GT_SIZE;
  double Gm[GT_SIZE][GT_SIZE], c[GT_SIZE];
  for (int a = 0; a < n; ++a) {
    for (int b = 0; b <= a; ++b) { Gm[a][b] = B.corner[a * kBorder + b] - Msum[a * 16 + b]; Gm[b][a] = Gm[a][b]; }
    c[a] = B.corner[kRhsRow * kBorder + a] - Msum[kRhsRow * 16 + a];
  }
  for (int j = 0; j < n; ++j) {
    double d = Gm[j][j];
    for (int k = 0; k < j; ++k) d -= Gm[j][k] * Gm[j][k];
    if (!(d > 0.0) || !isfinite(d)) { atomicExch(flag, 2); d = 1.0; }
    d = sqrt(d);
    Gm[j][j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = Gm[i][j];
      for (int k = 0; k < j; ++k) s -= Gm[i][k] * Gm[j][k];
      Gm[i][j] = s / d;
    }
  }
  for (int i = 0; i < n; ++i) { double s = c[i]; for (int k = 0; k < i; ++k) s -= Gm[i][k] * c[k]; c[i] = s / Gm[i][i]; }
  for (int i = n - 1; i >= 0; --i) { double s = c[i]; for (int k = i + 1; k < n; ++k) s -= Gm[k][i] * c[k]; c[i] = s / Gm[i][i]; }
  for (int i = 0; i < n; ++i) delta[B.n_pad + i] = c[i];
  delta[B.n_pad + kRhsRow] = 0.0;
}

// t = y - Lb^T z  (y = forward-substituted right-hand side = border row 15) into delta[0, n_pad)
__global__ void k_back_rhs(BandLayout B, double* __restrict__ delta) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= B.n_pad) return;
  double t = B.border[(size_t)kRhsRow * B.n_pad + c];
#pragma unroll
  for (int g = 0; g < GT_SIZE; ++g) t -= B.border[(size_t)g * B.n_pad + c] * delta[B.n_pad + g];
  delta[c] = t;
}

// x = L^-T t, block columns from last to first, by ONE thread-block cluster of kBackCtas CTAs (a single SM cannot stream the factor
// fast enough: every step reads nbw+1 blocks of 8 KB).  Sub-diagonal offset o (1..nbw) belongs to CTA (o-1) % kBackCtas, warp 1 + (o-1) / kBackCtas:
//   warp w>=1 : p_o = L(i+o, i)^T x_{i+o}     (block rows streamed with coalesced loads, the next column prefetched into registers)
//   warp 0    : sums its CTA's p_o, writes the partial into every CTA's exchange buffer through distributed shared memory;
//               after ONE cluster barrier every CTA redundantly forms  x_i = L_ii^-T (t_i - sum of partials)  and keeps its own ring of x blocks.
// The exchange buffer is double-buffered by step parity.  delta[0, n_pad) holds t on entry and x on exit.
constexpr int kBackCtas = 8;
__global__ void __launch_bounds__(160) k_band_backsolve(BandLayout B, const double* __restrict__ dinv, double* __restrict__ delta) {
  extern __shared__ double sh[];
  cg::cluster_group cluster = cg::this_cluster();
  const int rank = (int)cluster.block_rank();
  const int nb1 = B.nbw + 1;
  const int nwarps = blockDim.x >> 5;
  double* xs = sh;                               // nb1 * 32      ring of x blocks
  double* part = xs + nb1 * kNB;                 // nwarps * 32   per-warp products
  double* exch = part + nwarps * kNB;            // 2 * kBackCtas * 32
  double* ts = exch + 2 * kBackCtas * kNB;       // 32
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int off = (warp - 1) * kBackCtas + rank + 1;   // this warp's sub-diagonal offset (warp >= 1)
  auto src = [&](int i) -> const double* {
    if (i < 0) return nullptr;
    if (warp == 0) return dinv + (size_t)i * 1024 + lane;                                   // (L_ii^-1)[c][lane]
    if (off <= B.nbw && i + off < B.nbc) return B.band + ((size_t)i * nb1 + off) * 1024 + lane;   // L(i+off, i)[r][lane]
    return nullptr;
  };
  double cur[kNB], nxt[kNB];
  {
    const double* s0 = src(B.nbc - 1);
#pragma unroll
    for (int r = 0; r < kNB; ++r) cur[r] = s0 ? s0[r * kNB] : 0.0;
  }
  auto dot32 = [&](const double* v) -> double {
    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
#pragma unroll
    for (int r = 0; r < kNB; r += 4) { p0 = fma(cur[r], v[r], p0); p1 = fma(cur[r + 1], v[r + 1], p1); p2 = fma(cur[r + 2], v[r + 2], p2); p3 = fma(cur[r + 3], v[r + 3], p3); }
    return (p0 + p1) + (p2 + p3);
  };
  double t_cur = (warp == 0) ? delta[(size_t)(B.nbc - 1) * kNB + lane] : 0.0;
  cluster.sync();   // every CTA of the cluster is running before the first remote store
  for (int i = B.nbc - 1; i >= 0; --i) {
    const double* g = src(i);
    const double* gn = src(i - 1);
    const double t_nxt = (warp == 0 && i > 0) ? delta[(size_t)(i - 1) * kNB + lane] : 0.0;
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

static int band_enqueue(ctba_problem* p, cudaStream_t st) {
  SolverData* sd = p->solver;
  BandLayout& B = sd->Hw;
  CTBA_CUDA(p, cudaMemsetAsync(sd->flag.p, 0, 4 * sizeof(int32_t), st));
  for (int j = 0; j < B.nbc; ++j) {
    const int rows = std::min(B.nbw, B.nbc - 1 - j);
    int grid = rows + 2;
    if (j > 0) { const int m = std::min(B.nbw - 1, B.nbc - 1 - j); if (m > 0) grid += m * (m + 1) / 2 + m; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid); cfg.blockDim = dim3(kStepThreads); cfg.dynamicSmemBytes = 0; cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    const int pdl = (j > 0) ? p->band_pdl : 0;   // column 0 follows non-kernel work: plain stream order
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    CTBA_CUDA(p, cudaLaunchKernelEx(&cfg, k_band_step, B, j, sd->dinv.p, sd->flag.p, pdl));
  }
  const int nblk = cdiv(B.n_pad, 256);
  k_corner_partial<<<nblk, 256, 0, st>>>(B, sd->corner_partial.p);
  k_corner_solve<<<1, 256, 0, st>>>(B, nblk, sd->corner_partial.p, sd->delta.p, sd->flag.p);
  k_back_rhs<<<cdiv(B.n_pad, 256), 256, 0, st>>>(B, sd->delta.p);
  {
    const int nb1 = B.nbw + 1;
    const int nwarps = 1 + std::max(1, (B.nbw + kBackCtas - 1) / kBackCtas);   // <= 5 for nbw <= 31
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kBackCtas); cfg.blockDim = dim3(32 * nwarps); cfg.stream = st;
    cfg.dynamicSmemBytes = (size_t)(nb1 + nwarps + 2 * kBackCtas + 1) * kNB * sizeof(double);
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kBackCtas; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    CTBA_CUDA(p, cudaLaunchKernelEx(&cfg, k_band_backsolve, B, (const double*)sd->dinv.p, sd->delta.p));
  }
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

void band_graph_release(SolverData* sd) {
  if (sd->band_graph_exec) { cudaGraphExecDestroy((cudaGraphExec_t)sd->band_graph_exec); sd->band_graph_exec = nullptr; }
}

// Factor + solve.  The launch sequence depends only on the band structure and the buffer addresses, so it is captured once into a
// CUDA graph (cheaper kernel-to-kernel hand-over than 400+ stream launches) and replayed every LM iteration.
int band_factor_solve(ctba_problem* p) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  BandLayout& B = sd->Hw;
  if (B.nbw + 1 > 32) { p->err = "band_factor_solve: band wider than 31 blocks of 32 scalars"; return CTBA_E_ARG; }
  const bool fresh = sd->dinv.n < (size_t)B.nbc * 1024 || sd->corner_partial.n < (size_t)cdiv(B.n_pad, 256) * 256;
  CTBA_CUDA(p, sd->dinv.resize((size_t)B.nbc * 1024));
  CTBA_CUDA(p, sd->corner_partial.resize((size_t)cdiv(B.n_pad, 256) * 256));
  p->launches += B.nbc + 4;
  if (!p->band_graph) return band_enqueue(p, st);
  const int key = p->band_pdl;
  if (fresh || sd->band_graph_key != key) band_graph_release(sd);
  if (!sd->band_graph_exec) {
    cudaGraph_t graph = nullptr;
    CTBA_CUDA(p, cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal));
    const int rc = band_enqueue(p, st);
    const cudaError_t e = cudaStreamEndCapture(st, &graph);
    if (rc) { if (graph) cudaGraphDestroy(graph); return rc; }
    CTBA_CUDA(p, e);
    cudaGraphExec_t exec = nullptr;
    const cudaError_t ei = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    CTBA_CUDA(p, ei);
    sd->band_graph_exec = exec; sd->band_graph_key = key;
  }
  CTBA_CUDA(p, cudaGraphLaunch((cudaGraphExec_t)sd->band_graph_exec, st));
  return CTBA_OK;
}

}  // namespace ctba

#ifdef CTBA_BAND_TIMING
extern "C" int ctba_debug_band_clocks(long long* out8) {
  return cudaMemcpyFromSymbol(out8, ctba::g_band_clk, sizeof(long long) * 8) == cudaSuccess ? 0 : -2;
}
#endif
