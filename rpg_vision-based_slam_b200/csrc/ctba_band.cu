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
#include "ctba_solver.h"

namespace ctba {

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

constexpr int kLd = kNB + 4;       // shared-memory leading dimension
constexpr int kStepThreads = 256;  // 8 warps: warp w owns tiles (tr = w>>1, tc = 2*(w&1) + {0,1}) of a 32x32 block

// dst(32x32, smem ld kLd) = C - A*Bm^T  (rows < nrows of A / C valid, C row stride ldc; rows >= nrows become 0).  sub=false: dst = C.
// tile_abt is warp-collective (mma.sync): nrows is 16 or 32 and tiles are 8 rows high, so `r < nrows` is warp-uniform.
__device__ __forceinline__ void block_load_minus_abt(double* dst, const double* __restrict__ C, int ldc, int nrows, bool sub,
                                                     const double* __restrict__ A, int lda, const double* __restrict__ Bm, int warp, int lane) {
  const int tr = warp >> 1, g = lane >> 2, t = lane & 3;
  const int r = 8 * tr + g;
  const bool live = (8 * tr) < nrows;
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int tc = 2 * (warp & 1) + h;
    double c0 = 0.0, c1 = 0.0;
    if (live && sub) tile_abt(A, lda, Bm, kNB, tr, tc, lane, c0, c1);
    const int c = 8 * tc + 2 * t;
    double v0 = 0.0, v1 = 0.0;
    if (live) { const double2 cv = *reinterpret_cast<const double2*>(C + (size_t)r * ldc + c); v0 = cv.x - c0; v1 = cv.y - c1; }
    dst[r * kLd + c] = v0; dst[r * kLd + c + 1] = v1;
  }
}

// 32x32 Cholesky by ONE warp, matrix rows in registers (lane = row), right-looking, column broadcast by shuffles.
// On exit sD holds L (lower, zeros above the diagonal) and sRinv[c] = 1 / L[c][c].
__device__ __forceinline__ void potrf32_warp(double* sD, double* sRinv, int lane, int32_t* flag) {
  double a[kNB];
#pragma unroll
  for (int k = 0; k < kNB; ++k) a[k] = sD[lane * kLd + k];
  bool bad = false;
#pragma unroll
  for (int c = 0; c < kNB; ++c) {
    double piv = __shfl_sync(0xffffffffu, a[c], c);
    if (!(piv > 0.0) || !isfinite(piv)) { bad = true; piv = 1.0; }
    const double rinv = rsqrt(piv);
    const double l = a[c] * rinv;
    a[c] = l;
    if (lane == c) sRinv[c] = rinv;
#pragma unroll
    for (int k = c + 1; k < kNB; ++k) { const double lk = __shfl_sync(0xffffffffu, l, k); a[k] -= l * lk; }
  }
  if (bad && lane == 0) atomicExch(flag, 1);
#pragma unroll
  for (int k = 0; k < kNB; ++k) sD[lane * kLd + k] = (k <= lane) ? a[k] : 0.0;
}
// X * L^T = ACC for the 32 rows of sX (lane = row), in place.  L and 1/diag from shared memory (uniform-address broadcasts).
__device__ __forceinline__ void trsm32_warp(double* sX, const double* sD, const double* sRinv, int lane) {
  double x[kNB];
#pragma unroll
  for (int k = 0; k < kNB; ++k) x[k] = sX[lane * kLd + k];
#pragma unroll
  for (int c = 0; c < kNB; ++c) {
    x[c] *= sRinv[c];
#pragma unroll
    for (int k = c + 1; k < kNB; ++k) x[k] -= x[c] * sD[k * kLd + c];
  }
#pragma unroll
  for (int k = 0; k < kNB; ++k) sX[lane * kLd + k] = x[k];
}

// One launch per block column j (right-looking with a one-column look-ahead):
//   panel CTAs  [0, rows+2)      : apply column j-1's update to THEIR block of column j (and to the diagonal block, redundantly),
//                                  factor the diagonal block, triangular-solve their block, store it.  CTA 0 = diagonal (also stores
//                                  L_jj^-1 for the backward substitution), CTA rows+1 = the 16 border rows (15 globals + right-hand side).
//   update CTAs [rows+2, ...)    : remaining trailing update of column j-1:  T(i,k) -= L(i,j-1) L(k,j-1)^T  for j+1 <= k <= i <= j-1+nbw,
//                                  and the border rows of those block columns.
// Pre-condition of launch j: columns < j final, every trailing block carries the updates of columns <= j-2.
__global__ void __launch_bounds__(kStepThreads) k_band_step(BandLayout B, int j, double* __restrict__ dinv, int32_t* __restrict__ flag) {
  __shared__ __align__(16) double sD[kNB * kLd];
  __shared__ __align__(16) double sX[kNB * kLd];
  __shared__ double sRinv[kNB];
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
    block_load_minus_abt(sD, Ajj, kNB, kNB, prev, Lprev + 1024, kNB, Lprev + 1024, warp, lane);
    const bool is_border = (b == rows + 1);
    if (b > 0) {
      if (!is_border) {
        const bool have_prev = prev && (b + 1 <= B.nbw);   // block (j+b, j-1) lies inside the band
        block_load_minus_abt(sX, Ajj + (size_t)b * 1024, kNB, kNB, have_prev, Lprev + (size_t)min(b + 1, B.nbw) * 1024, kNB, Lprev + 1024, warp, lane);
      } else {
        block_load_minus_abt(sX, B.border + (size_t)j * kNB, B.n_pad, kBorder, prev, Bprev, B.n_pad, Lprev + 1024, warp, lane);
      }
    }
    __syncthreads();
    if (warp == 0) {
      potrf32_warp(sD, sRinv, lane, flag);
      __syncwarp();
      if (b > 0) {
        trsm32_warp(sX, sD, sRinv, lane);
      } else {
        // L_jj^-1 for the backward substitution: solve X L^T = I (X = L^-T, upper) and store its transpose (lower, row-major)
        double x[kNB];
#pragma unroll
        for (int k = 0; k < kNB; ++k) x[k] = (k == lane) ? 1.0 : 0.0;
#pragma unroll
        for (int c = 0; c < kNB; ++c) {
          x[c] *= sRinv[c];
#pragma unroll
          for (int k = c + 1; k < kNB; ++k) x[k] -= x[c] * sD[k * kLd + c];
        }
#pragma unroll
        for (int k = 0; k < kNB; ++k) sX[k * kLd + lane] = x[k];   // sX[c][r] = X[r][c] = (L^-1)[c][r]
      }
    }
    __syncthreads();
    if (b == 0) {
      for (int e = tid; e < kNB * kNB; e += kStepThreads) {
        const int r = e >> 5, c = e & 31;
        Ajj[e] = sD[r * kLd + c];
        dinv[(size_t)j * 1024 + e] = sX[r * kLd + c];
      }
    } else if (!is_border) {
      double* Lij = Ajj + (size_t)b * 1024;
      for (int e = tid; e < kNB * kNB; e += kStepThreads) Lij[e] = sX[(e >> 5) * kLd + (e & 31)];
    } else {
      for (int e = tid; e < kBorder * kNB; e += kStepThreads) { const int r = e >> 5, c = e & 31; B.border[(size_t)r * B.n_pad + j * kNB + c] = sX[r * kLd + c]; }
    }
    return;
  }
  // ---------------- trailing update of column j-1 (only launched for j >= 1) ----------------
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

// x = L^-T t, block columns from last to first, single CTA with nbw+1 warps:
//   warp o (1..nbw) : p_o = L(i+o, i)^T x_{i+o}   (block rows streamed with coalesced loads)
//   warp 0          : x_i = L_ii^-T (t_i - sum_o p_o)
// x blocks live in a shared ring of nbw+1 slots.  delta[0, n_pad) holds t on entry and x on exit.
// kPrefetch (<= 16 warps, 128 registers per thread): the next column's 32 values per lane are loaded before the current
// column's products are formed, so the L2 latency of the stream is off the critical path.
template <int kMaxThreads, bool kPrefetch>
__global__ void __launch_bounds__(kMaxThreads) k_band_backsolve(BandLayout B, const double* __restrict__ dinv, double* __restrict__ delta) {
  extern __shared__ double sh[];
  const int nb1 = B.nbw + 1;
  double* xs = sh;                   // nb1 * 32
  double* part = sh + nb1 * kNB;     // nb1 * 32
  double* ts = part + nb1 * kNB;     // 32
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // source of this warp's 32x32 operand for block column i: warp 0 -> (L_ii^-1)[c][lane], warp o -> L(i+o,i)[r][lane]
  auto src = [&](int i) -> const double* {
    if (i < 0) return nullptr;
    if (warp == 0) return dinv + (size_t)i * 1024 + lane;
    if (i + warp < B.nbc) return B.band + ((size_t)i * nb1 + warp) * 1024 + lane;
    return nullptr;
  };
  double cur[kPrefetch ? kNB : 1], nxt[kPrefetch ? kNB : 1];
  if (kPrefetch) {
    const double* s0 = src(B.nbc - 1);
#pragma unroll
    for (int r = 0; r < kNB; ++r) { cur[kPrefetch ? r : 0] = s0 ? s0[r * kNB] : 0.0; }
  }
  // dot of this warp's operand with a 32-vector in shared memory (uniform-address broadcasts)
  auto dot32 = [&](const double* g, const double* v) -> double {
    double p0 = 0.0, p1 = 0.0, p2 = 0.0, p3 = 0.0;
    if (kPrefetch) {
#pragma unroll
      for (int r = 0; r < kNB; r += 4) {
        p0 += cur[kPrefetch ? r : 0] * v[r]; p1 += cur[kPrefetch ? r + 1 : 0] * v[r + 1];
        p2 += cur[kPrefetch ? r + 2 : 0] * v[r + 2]; p3 += cur[kPrefetch ? r + 3 : 0] * v[r + 3];
      }
    } else {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        double w[16];
#pragma unroll
        for (int r = 0; r < 16; ++r) w[r] = g[(16 * h + r) * kNB];
#pragma unroll
        for (int r = 0; r < 16; r += 4) {
          p0 += w[r] * v[16 * h + r]; p1 += w[r + 1] * v[16 * h + r + 1]; p2 += w[r + 2] * v[16 * h + r + 2]; p3 += w[r + 3] * v[16 * h + r + 3];
        }
      }
    }
    return (p0 + p1) + (p2 + p3);
  };
  for (int i = B.nbc - 1; i >= 0; --i) {
    const double* g = src(i);
    if (kPrefetch) {
      const double* gn = src(i - 1);
#pragma unroll
      for (int r = 0; r < kNB; ++r) nxt[kPrefetch ? r : 0] = gn ? gn[r * kNB] : 0.0;
    }
    if (warp > 0) part[warp * kNB + lane] = g ? dot32(g, xs + ((i + warp) % nb1) * kNB) : 0.0;
    __syncthreads();
    if (warp == 0) {
      double t = delta[(size_t)i * kNB + lane];
      for (int o = 1; o < nb1; ++o) t -= part[o * kNB + lane];
      ts[lane] = t;
      __syncwarp();
      const double x = dot32(g, ts);
      xs[(i % nb1) * kNB + lane] = x;
      delta[(size_t)i * kNB + lane] = x;
    }
    __syncthreads();
    if (kPrefetch) {
#pragma unroll
      for (int r = 0; r < kNB; ++r) cur[kPrefetch ? r : 0] = nxt[kPrefetch ? r : 0];
    }
  }
}

int band_factor_solve(ctba_problem* p) {
  SolverData* sd = p->solver;
  cudaStream_t st = p->stream;
  BandLayout& B = sd->Hw;
  if (B.nbw + 1 > 32) { p->err = "band_factor_solve: band wider than 31 blocks of 32 scalars"; return CTBA_E_ARG; }
  CTBA_CUDA(p, sd->dinv.resize((size_t)B.nbc * 1024));
  CTBA_CUDA(p, cudaMemsetAsync(sd->flag.p, 0, 4 * sizeof(int32_t), st));
  for (int j = 0; j < B.nbc; ++j) {
    const int rows = std::min(B.nbw, B.nbc - 1 - j);
    int grid = rows + 2;
    if (j > 0) { const int m = std::min(B.nbw - 1, B.nbc - 1 - j); if (m > 0) grid += m * (m + 1) / 2 + m; }
    k_band_step<<<grid, kStepThreads, 0, st>>>(B, j, sd->dinv.p, sd->flag.p);
  }
  p->launches += B.nbc;
  const int nblk = cdiv(B.n_pad, 256);
  double* partial = sd->Sb.p;  // F*42 doubles are free again after k_pose_to_knots; need nblk*256
  if ((size_t)nblk * 256 > sd->Sb.n) { CTBA_CUDA(p, sd->Sb.resize((size_t)nblk * 256)); partial = sd->Sb.p; }
  k_corner_partial<<<nblk, 256, 0, st>>>(B, partial);
  k_corner_solve<<<1, 256, 0, st>>>(B, nblk, partial, sd->delta.p, sd->flag.p);
  k_back_rhs<<<cdiv(B.n_pad, 256), 256, 0, st>>>(B, sd->delta.p);
  const int nb1 = B.nbw + 1;
  if (nb1 <= 16) k_band_backsolve<512, true><<<1, 32 * nb1, (2 * nb1 + 1) * kNB * sizeof(double), st>>>(B, sd->dinv.p, sd->delta.p);
  else k_band_backsolve<1024, false><<<1, 32 * nb1, (2 * nb1 + 1) * kNB * sizeof(double), st>>>(B, sd->dinv.p, sd->delta.p);
  p->launches += 4;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

}  // namespace ctba
