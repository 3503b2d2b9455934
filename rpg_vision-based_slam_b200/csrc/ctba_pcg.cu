// ctba_pcg.cu — block-Jacobi preconditioned conjugate gradients on the reduced (landmark-eliminated) LM system.
//
// The iterative alternative to the bordered band Cholesky of ctba_band.cu (north_star (3): "block-Jacobi PCG, or banded block Cholesky where
// the band fits"; the reference's counterpart would be ceres ITERATIVE_SCHUR — its shipped drivers use the exact SPARSE_NORMAL_CHOLESKY,
// include/spline/bspline_so3r3.h:1771).  It solves the SAME system  [H Bd^T; Bd G] [x; z] = [y; c]  of ctba_solver.h (H: block band, lower
// blocks stored; Bd: the <= 15 global rows; right-hand side = border row 15 / corner row 15), so the LM loop, the landmark
// back-substitution and the step scalars are untouched: option linear_solver = 1.  Preconditioner: the inverses of the 32 x 32 diagonal
// blocks of H and of the 15 x 15 corner.  Every reduction is a two-stage sum in a fixed order (bit-reproducible).
//   k_pcg_precond_build   L L^T of every diagonal block in shared memory, M_j^-1 = L^-T L^-1
//   k_pcg_matvec_col / k_pcg_matvec_row   q = A p: the transposed products of block column j (coalesced) / the direct products of block row i
//   k_pcg_glob            global rows: q_g = Bd p + G p_g
//   k_pcg_update / k_pcg_dir   x, r, z = M^-1 r, the dot products; p = z + beta p
// The product reads every stored block twice (16 KB per 32 x 32 block and iteration): an HBM-bound kernel, 2 (nbw+1) nbc 8 KB per iteration.
#include <algorithm>
#include <cmath>
#include <initializer_list>
#include <cstdio>
#include <vector>
#include "ctba_solver.h"

namespace ctba {

static inline int cdiv(size_t a, int b) { return (int)((a + b - 1) / b); }

struct PcgData {
  DBuf<double> minv, minv_g, x, r, z, pd, q, part, scal;   // scal: [rz, pAp, rr, bb, alpha, beta, rz_new]
  int nbc = 0;
};

// ---- preconditioner: one warp per diagonal block --------------------------------------------------------------------
__global__ void __launch_bounds__(32) k_pcg_precond_build(BandLayout B, double* __restrict__ minv, int32_t* __restrict__ flag) {
  __shared__ double L[32][33], X[32][33];
  const int j = blockIdx.x, lane = threadIdx.x;
  const double* A = B.band + ((size_t)j * (B.nbw + 1)) * 1024;
  for (int r = 0; r < 32; ++r) L[r][lane] = lane <= r ? A[r * 32 + lane] : 0.0;   // lower triangle as assembled
  __syncwarp();
  bool bad = false;
  for (int c = 0; c < 32; ++c) {   // right-looking Cholesky, lane = row
    const double d = L[c][c];
    if (!(d > 0.0) || !isfinite(d)) bad = true;
    const double rinv = 1.0 / sqrt(bad ? 1.0 : d);
    const double l = lane >= c ? L[lane][c] * rinv : 0.0;
    __syncwarp();
    L[lane][c] = l;
    __syncwarp();
    for (int k = c + 1; k < 32; ++k) if (lane >= k) L[lane][k] -= l * L[k][c];
    __syncwarp();
  }
  // X = L^-1: lane = column of the identity, forward substitution
  double xcol[32];
#pragma unroll
  for (int r = 0; r < 32; ++r) {
    double s = (r == lane) ? 1.0 : 0.0;
    for (int k = 0; k < r; ++k) s -= L[r][k] * xcol[k];
    xcol[r] = s / L[r][r];
  }
#pragma unroll
  for (int r = 0; r < 32; ++r) X[r][lane] = xcol[r];
  __syncwarp();
  // M^-1 = X^T X: entry (a, lane)
  for (int a = 0; a < 32; ++a) {
    double s = 0.0;
    for (int r = 0; r < 32; ++r) s += X[r][a] * X[r][lane];
    minv[(size_t)j * 1024 + a * 32 + lane] = s;
  }
  if (bad && lane == 0) atomicExch(flag, 1);
}
// corner: dense inverse of the 15 x 15 block (unit diagonal on inactive globals) by Gauss-Jordan, one thread
__global__ void k_pcg_precond_corner(BandLayout B, double* __restrict__ minv_g, int32_t* __restrict__ flag) {
  if (threadIdx.x || blockIdx.x) return;
  const int n = GT_SIZE;
  double A[GT_SIZE][GT_SIZE], I[GT_SIZE][GT_SIZE];
  for (int a = 0; a < n; ++a) for (int b = 0; b < n; ++b) { A[a][b] = a >= b ? B.corner[a * kBorder + b] : B.corner[b * kBorder + a]; I[a][b] = a == b ? 1.0 : 0.0; }
  for (int c = 0; c < n; ++c) {
    const double d = A[c][c];
    if (!(d > 0.0) || !isfinite(d)) { atomicExch(flag, 2); return; }
    const double inv = 1.0 / d;
    for (int b = 0; b < n; ++b) { A[c][b] *= inv; I[c][b] *= inv; }
    for (int a = 0; a < n; ++a) if (a != c) { const double f = A[a][c]; for (int b = 0; b < n; ++b) { A[a][b] -= f * A[c][b]; I[a][b] -= f * I[c][b]; } }
  }
  for (int a = 0; a < n; ++a) for (int b = 0; b < n; ++b) minv_g[a * kBorder + b] = I[a][b];
}

// ---- q = A p ------------------------------------------------------------------------------------------------------------
// block column j (the blocks (j + o, j), o = 0 .. nbw, are contiguous): q_j = sum_o Bk_o^T p_{j+o}; the diagonal block holds its lower triangle
__global__ void __launch_bounds__(256) k_pcg_matvec_col(BandLayout B, const double* __restrict__ p, double* __restrict__ q) {
  __shared__ double red[8][32];
  const int j = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nb1 = B.nbw + 1;
  double acc = 0.0;
  for (int o = warp; o < nb1 && j + o < B.nbc; o += 8) {
    const double* Bk = B.band + ((size_t)j * nb1 + o) * 1024;
    const double* pi = p + (size_t)(j + o) * 32;
    if (o == 0) { for (int r = lane; r < 32; ++r) acc += Bk[r * 32 + lane] * pi[r]; }         // column `lane` of the lower triangle: rows >= lane
    else { for (int r = 0; r < 32; ++r) acc += Bk[r * 32 + lane] * pi[r]; }
  }
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) { double s = 0.0; for (int w = 0; w < 8; ++w) s += red[w][lane]; q[(size_t)j * 32 + lane] = s; }
}
// block row i: q_i += sum_{o >= 1} H(i, i - o) p_{i-o} + (strictly lower part of the diagonal block) p_i + Bd^T p_g
__global__ void __launch_bounds__(256) k_pcg_matvec_row(BandLayout B, const double* __restrict__ p, double* __restrict__ q) {
  __shared__ double red[8][32];
  const int i = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nb1 = B.nbw + 1;
  double acc = 0.0;
  for (int o = warp; o <= B.nbw && i - o >= 0; o += 8) {
    const double* Bk = B.band + ((size_t)(i - o) * nb1 + o) * 1024 + lane * 32;   // row `lane` of block (i, i - o)
    const double* pj = p + (size_t)(i - o) * 32;
    if (o == 0) { for (int c = 0; c < lane; ++c) acc += Bk[c] * pj[c]; }           // strictly lower part: the diagonal went with the column pass
    else { for (int c = 0; c < 32; ++c) acc += Bk[c] * pj[c]; }
  }
  if (warp == 7) for (int g = 0; g < GT_SIZE; ++g) acc += B.border[(size_t)g * B.n_pad + (size_t)i * 32 + lane] * p[B.n_pad + g];
  red[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) { double s = 0.0; for (int w = 0; w < 8; ++w) s += red[w][lane]; q[(size_t)i * 32 + lane] += s; }
}
// global rows: q_g = Bd[g] . p + G[g] . p_g   (one CTA per global row, fixed-order tree)
__global__ void __launch_bounds__(256) k_pcg_glob(BandLayout B, const double* __restrict__ p, double* __restrict__ q) {
  __shared__ double red[256];
  const int g = blockIdx.x;
  double s = 0.0;
  for (size_t c = threadIdx.x; c < (size_t)B.n_pad; c += 256) s += B.border[(size_t)g * B.n_pad + c] * p[c];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if (threadIdx.x < w) red[threadIdx.x] += red[threadIdx.x + w]; __syncthreads(); }
  if (threadIdx.x == 0) {
    double t = red[0];
    for (int h = 0; h < GT_SIZE; ++h) t += (g >= h ? B.corner[g * kBorder + h] : B.corner[h * kBorder + g]) * p[B.n_pad + h];
    q[B.n_pad + g] = t;
  }
}

// ---- vector kernels; block `nbc` of every launch handles the 16 global entries ------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// mode 0 (start): r = b, x = 0, z = M^-1 r, p = z, partials of rz and bb.   mode 1: x += alpha p, r -= alpha q, z = M^-1 r, partials of rz_new, rr
__global__ void __launch_bounds__(32) k_pcg_update(BandLayout B, int mode, const double* __restrict__ minv, const double* __restrict__ minv_g, double* __restrict__ x,
                                                    double* __restrict__ r, double* __restrict__ z, double* __restrict__ pd, const double* __restrict__ q,
                                                    const double* __restrict__ scal, double* __restrict__ part) {
  __shared__ double rs[32];
  const int j = blockIdx.x, lane = threadIdx.x;
  const bool glob = j == B.nbc;
  const size_t i = glob ? (size_t)B.n_pad + lane : (size_t)j * 32 + lane;
  const bool live = !glob || lane < GT_SIZE;          // an unknown
  const bool inb = !glob || lane < kBorder;            // inside the vectors (entry n_pad + 15 is the unused right-hand-side slot: kept at 0)
  double rv = 0.0;
  if (mode == 0) {
    rv = !live ? 0.0 : glob ? B.corner[kRhsRow * kBorder + lane] : B.border[(size_t)kRhsRow * B.n_pad + i];
    if (inb) { x[i] = 0.0; r[i] = rv; }
  } else if (live) {
    const double alpha = scal[4];
    x[i] += alpha * pd[i];
    rv = r[i] - alpha * q[i];
    r[i] = rv;
  }
  rs[lane] = rv;
  __syncwarp();
  double zv = 0.0;
  if (glob) { if (live) for (int c = 0; c < GT_SIZE; ++c) zv += minv_g[lane * kBorder + c] * rs[c]; }
  else { const double* M = minv + (size_t)j * 1024 + lane * 32; for (int c = 0; c < 32; ++c) zv += M[c] * rs[c]; }
  if (inb) { z[i] = zv; if (mode == 0) pd[i] = zv; }
  const double a = warp_sum(rv * zv), b = warp_sum(rv * rv);
  if (lane == 0) { part[2 * j] = a; part[2 * j + 1] = b; }
}
// fixed-order sums of the per-block partials; which = 0: after the start (rz, bb), 1: pAp -> alpha, 2: rz_new, rr -> beta
__global__ void __launch_bounds__(256) k_pcg_scalars(int which, int nblk, const double* __restrict__ part, double* __restrict__ scal) {
  __shared__ double sa[256], sb[256];
  double a = 0.0, b = 0.0;
  for (int k = threadIdx.x; k < nblk; k += 256) { a += part[2 * k]; b += part[2 * k + 1]; }
  sa[threadIdx.x] = a; sb[threadIdx.x] = b;
  __syncthreads();
  for (int w = 128; w > 0; w >>= 1) { if (threadIdx.x < w) { sa[threadIdx.x] += sa[threadIdx.x + w]; sb[threadIdx.x] += sb[threadIdx.x + w]; } __syncthreads(); }
  if (threadIdx.x == 0) {
    if (which == 0) { scal[0] = sa[0]; scal[2] = sb[0]; scal[3] = sb[0]; }
    else if (which == 1) { scal[1] = sa[0]; scal[4] = sa[0] > 0.0 ? scal[0] / sa[0] : 0.0; }
    else { scal[5] = scal[0] > 0.0 ? sa[0] / scal[0] : 0.0; scal[0] = sa[0]; scal[2] = sb[0]; }
  }
}
// partials of p . q
__global__ void __launch_bounds__(32) k_pcg_dot_pq(BandLayout B, const double* __restrict__ pd, const double* __restrict__ q, double* __restrict__ part) {
  const int j = blockIdx.x, lane = threadIdx.x;
  const bool glob = j == B.nbc;
  const size_t i = glob ? (size_t)B.n_pad + lane : (size_t)j * 32 + lane;
  const double v = (!glob || lane < GT_SIZE) ? pd[i] * q[i] : 0.0;
  const double a = warp_sum(v);
  if (lane == 0) { part[2 * j] = a; part[2 * j + 1] = 0.0; }
}
__global__ void k_pcg_dir(int n, const double* __restrict__ z, const double* __restrict__ scal, double* __restrict__ pd) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) pd[i] = z[i] + scal[5] * pd[i];
}

void pcg_release(SolverData* sd) {
  if (!sd->pcg) return;
  PcgData* d = (PcgData*)sd->pcg;
  for (DBuf<double>* b : {&d->minv, &d->minv_g, &d->x, &d->r, &d->z, &d->pd, &d->q, &d->part, &d->scal}) b->release();
  delete d;
  sd->pcg = nullptr;
}

// Solves the Hw system into sd->delta.  Stops at |r| <= pcg_tol |b| (checked every 25 iterations) or after pcg_max_iters iterations.
int band_pcg_solve(ctba_problem* p) {
  SolverData* sd = p->solver;
  if (p->sharded) { p->err = "linear_solver = pcg: not available for a problem sharded over ranks"; return CTBA_E_STATE; }
  cudaStream_t st = p->stream;
  const BandLayout& B = sd->Hw;
  if (!sd->pcg) sd->pcg = new PcgData();
  PcgData& d = *(PcgData*)sd->pcg;
  const int n = B.n_pad + kBorder, nblk = B.nbc + 1;
  CTBA_CUDA(p, d.minv.resize((size_t)B.nbc * 1024)); CTBA_CUDA(p, d.minv_g.resize(kBorder * kBorder));
  for (DBuf<double>* b : {&d.x, &d.r, &d.z, &d.pd, &d.q}) CTBA_CUDA(p, b->resize(n));
  CTBA_CUDA(p, d.part.resize(2 * (size_t)nblk)); CTBA_CUDA(p, d.scal.resize(8));
  CTBA_CUDA(p, cudaMemsetAsync(sd->flag.p, 0, 4 * sizeof(int32_t), st));
  CTBA_CUDA(p, cudaMemsetAsync(d.q.p, 0, n * sizeof(double), st));
  k_pcg_precond_build<<<B.nbc, 32, 0, st>>>(B, d.minv.p, sd->flag.p);
  k_pcg_precond_corner<<<1, 32, 0, st>>>(B, d.minv_g.p, sd->flag.p);
  k_pcg_update<<<nblk, 32, 0, st>>>(B, 0, d.minv.p, d.minv_g.p, d.x.p, d.r.p, d.z.p, d.pd.p, d.q.p, d.scal.p, d.part.p);
  k_pcg_scalars<<<1, 256, 0, st>>>(0, nblk, d.part.p, d.scal.p);
  int64_t nl = 4;
  double h[8];
  const double tol2 = p->pcg_tol * p->pcg_tol;
  int it = 0;
  sd->pcg_iterations = 0;
  while (it < p->pcg_max_iters) {
    const int burst = std::min(25, p->pcg_max_iters - it);
    for (int b = 0; b < burst; ++b) {
      k_pcg_matvec_col<<<B.nbc, 256, 0, st>>>(B, d.pd.p, d.q.p);
      k_pcg_matvec_row<<<B.nbc, 256, 0, st>>>(B, d.pd.p, d.q.p);
      k_pcg_glob<<<GT_SIZE, 256, 0, st>>>(B, d.pd.p, d.q.p);
      k_pcg_dot_pq<<<nblk, 32, 0, st>>>(B, d.pd.p, d.q.p, d.part.p);
      k_pcg_scalars<<<1, 256, 0, st>>>(1, nblk, d.part.p, d.scal.p);
      k_pcg_update<<<nblk, 32, 0, st>>>(B, 1, d.minv.p, d.minv_g.p, d.x.p, d.r.p, d.z.p, d.pd.p, d.q.p, d.scal.p, d.part.p);
      k_pcg_scalars<<<1, 256, 0, st>>>(2, nblk, d.part.p, d.scal.p);
      k_pcg_dir<<<cdiv(n, 256), 256, 0, st>>>(n, d.z.p, d.scal.p, d.pd.p);
      nl += 8;
    }
    it += burst;
    CTBA_CUDA(p, cudaMemcpyAsync(h, d.scal.p, sizeof(h), cudaMemcpyDeviceToHost, st));
    CTBA_CUDA(p, cudaStreamSynchronize(st));
    if (!(h[2] == h[2]) || !(h[3] > 0.0) || h[2] <= tol2 * h[3]) break;
  }
  sd->pcg_iterations = it;
  sd->pcg_rel_residual = (h[3] > 0.0 && h[2] == h[2]) ? std::sqrt(h[2] / h[3]) : -1.0;
  if (!(h[2] == h[2])) { int32_t one = 1; CTBA_CUDA(p, cudaMemcpyAsync(sd->flag.p, &one, sizeof(one), cudaMemcpyHostToDevice, st)); }
  CTBA_CUDA(p, cudaMemcpyAsync(sd->delta.p, d.x.p, n * sizeof(double), cudaMemcpyDeviceToDevice, st));
  p->launches += nl;
  CTBA_CUDA(p, cudaGetLastError());
  return CTBA_OK;
}

}  // namespace ctba
