// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// The same trust-region Levenberg-Marquardt as lm.h (ceres::Solve with the reference's options, include/spline/bspline_so3r3.h:1768-1782
// [Ceres, recalled — SURVEY.md Appendix B]) with a SPARSE exact linear solve on host threads, so that whole BASELINE configs can be run on
// the CPU: it is what `bench.py --impl reference` / `cpu_baseline` time, and it gives LM traces at sizes the dense solver of lm.h cannot reach.
//
// SPARSE_NORMAL_CHOLESKY factors  J^T J + D^T D  exactly (CHOLMOD / Eigen; neither is in this image).  Exactness is all that parity needs, so
// the elimination order is ours: landmarks first (3x3 blocks, Schur complement), then the spline / bias knots in time order (a symmetric band,
// because every residual touches N consecutive knots), the few global parameters last (dense border).  This is FAVOURABLE to the CPU: the
// reference's frame blocks are structurally dense over all landmarks of a frame, which a general sparse Cholesky has to carry as fill.
//
// Residuals and Jacobians come from the same MiniProblem blocks as lm.h (Jet autodiff with Ceres' DynamicAutoDiff schedule), evaluated on
// `num_threads` threads (ceres::Solver::Options::num_threads, :1773).
#pragma once
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstring>
#include <functional>
#include <thread>
#include <vector>
#include "lm.h"

namespace oracle {

struct SparseTimings {
  double eval_jac_s = 0, eval_res_s = 0, build_s = 0, factor_s = 0, total_s = 0;
  int jac_evals = 0, res_evals = 0, solves = 0;
  int n_band = 0, half_bandwidth = 0, n_globals = 0, n_landmarks = 0;
};

namespace sparse_detail {

inline double now_s() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

template <class F>
inline void parallel_for(int nthreads, size_t n, F&& fn) {   // fn(begin, end, tid) over contiguous ranges
  if (nthreads <= 1 || n < 2) { fn((size_t)0, n, 0); return; }
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back([&, t] { fn(n * t / nthreads, n * (t + 1) / nthreads, t); });
  for (auto& t : th) t.join();
}
template <class F>
inline void parallel_dynamic(int nthreads, size_t n, F&& fn) {  // fn(index, tid), work stealing through one counter
  if (nthreads <= 1 || n < 2) { for (size_t i = 0; i < n; ++i) fn(i, 0); return; }
  std::atomic<size_t> next{0};
  std::vector<std::thread> th;
  for (int t = 0; t < nthreads; ++t) th.emplace_back([&, t] { for (size_t i; (i = next.fetch_add(1)) < n;) fn(i, t); });
  for (auto& t : th) t.join();
}

struct SpinBarrier {
  std::atomic<int> count{0}; std::atomic<int> sense{0}; int n;
  explicit SpinBarrier(int nn) : n(nn) {}
  void wait() {
    const int s = sense.load(std::memory_order_acquire);
    if (count.fetch_add(1, std::memory_order_acq_rel) == n - 1) { count.store(0, std::memory_order_relaxed); sense.store(s ^ 1, std::memory_order_release); }
    else { int spins = 0; while (sense.load(std::memory_order_acquire) == s) if (++spins > 2000) std::this_thread::yield(); }
  }
};

// One residual block after the loss corrector, columns restricted to ACTIVE parameters.  `cols` are MiniProblem tangent columns of the
// band / global parameters; the landmark part of a frame block is kept per observation (2 x 3 and the landmark's tangent column).
struct CompactBlock {
  int nres = 0, nc = 0;
  std::vector<int> cols;
  std::vector<double> J, r;
  std::vector<int> lmcol;      // frames: per observation, first tangent column of its landmark (-1: constant)
  std::vector<double> Jl;      // frames: per observation 2 x 3 (row-major)
  double cost = 0;
};

// C[k] -= a * P[k]: the one hot loop of the band factorisation
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target_clones("avx512f", "avx2,fma", "default")))
#endif
inline void axpy_neg(double* __restrict__ c, const double* __restrict__ p, double a, int n) {
  for (int k = 0; k < n; ++k) c[k] -= a * p[k];
}

// Bordered band system:  [ H  Br^T ; Br  Cn ],  H symmetric n x n with half bandwidth hb, stored by rows: entry (i, c), i-hb <= c <= i, at
// band[i*(hb+1) + c - i + hb];  border nb rows x n (row-major; the LAST border row carries the right-hand side, so that forward substitution
// is part of the factorisation);  corner nb x nb (lower part).
struct BandSystem {
  int n = 0, hb = 0, nb = 0;   // nb = globals + 1
  std::vector<double> band, border, corner;
  void resize(int n_, int hb_, int nb_) { n = n_; hb = hb_; nb = nb_; band.assign((size_t)n * (hb + 1), 0.0); border.assign((size_t)nb * n, 0.0); corner.assign((size_t)nb * nb, 0.0); }
  void add(int i, int j, double v) {   // unified index space: [0, n) band, n + g border
    const int r = i > j ? i : j, c = i > j ? j : i;
    if (r < n) band[(size_t)r * (hb + 1) + (c - r + hb)] += v;
    else if (c < n) border[(size_t)(r - n) * n + c] += v;
    else corner[(size_t)(r - n) * nb + (c - n)] += v;
  }
};

// In-place factorisation + solve.  On exit x (n + nb - 1) holds the solution of [H Br^T; Br Cn] x = rhs, rhs = last border row / corner row.
inline bool band_solve(BandSystem& S, std::vector<double>& x, int nthreads) {
  const int n = S.n, hb = S.hb, nb = S.nb, ld = hb + 1, W = 64;
  bool ok = true;
  const int T = std::max(1, std::min(nthreads, 64));
  SpinBarrier bar(T);
  std::vector<double> PT((size_t)W * (hb + W));   // transposed panel: PT[m][row - (J0 + wJ)]
  auto worker = [&](int tid) {
    for (int J0 = 0; J0 < n; J0 += W) {
      const int wJ = std::min(W, n - J0), r0 = J0 + wJ, r1 = std::min(n, r0 + hb);   // rows below the diagonal block that see it: [r0, r1)
      if (tid == 0) {   // diagonal block, unblocked; a row only stores columns >= row - hb
        for (int j = J0; j < J0 + wJ; ++j) {
          double* rj = &S.band[(size_t)j * ld + hb];   // rj[c - j]
          double d = rj[0];
          for (int m = std::max(J0, j - hb); m < j; ++m) d -= rj[m - j] * rj[m - j];
          if (!(d > 0.0) || !std::isfinite(d)) { ok = false; d = 1.0; }
          d = std::sqrt(d); rj[0] = d;
          for (int i = j + 1; i < J0 + wJ && i - hb <= j; ++i) {   // rows of the block that reach column j
            double* ri = &S.band[(size_t)i * ld + hb];
            double s = ri[j - i];
            for (int m = std::max(J0, i - hb); m < j; ++m) s -= ri[m - i] * rj[m - j];
            ri[j - i] = s / d;
          }
        }
      }
      bar.wait();
      // triangular solves of the rows below (and of the border rows) against the diagonal block
      const int nrows = (r1 - r0) + nb;
      for (int q = tid; q < nrows; q += T) {
        double* row; int first = J0;
        if (q < r1 - r0) { const int i = r0 + q; row = &S.band[(size_t)i * ld + hb] - i; first = std::max(J0, i - hb); }   // row[c]
        else row = &S.border[(size_t)(q - (r1 - r0)) * n];
        for (int c = first; c < J0 + wJ; ++c) {
          const double* rc = &S.band[(size_t)c * ld + hb] - c;
          double s = row[c];
          for (int m = std::max(first, c - hb); m < c; ++m) s -= row[m] * rc[m];
          row[c] = s / rc[c];
        }
      }
      bar.wait();
      if (tid == 0)   // transposed copy of the panel rows [r0, r1): PT[m][k - r0] = L[k][J0 + m] (zero left of the band)
        for (int k = r0; k < r1; ++k) {
          const double* rk = &S.band[(size_t)k * ld + hb] - k;
          for (int m = 0; m < wJ; ++m) PT[(size_t)m * (hb + W) + (k - r0)] = (J0 + m >= k - hb) ? rk[J0 + m] : 0.0;
        }
      bar.wait();
      // trailing update: band rows i in [r0, r1): columns [r0, i]; border rows: columns [r0, r1)
      for (int q = tid; q < nrows; q += T) {
        if (q < r1 - r0) {
          const int i = r0 + q;
          double* ri = &S.band[(size_t)i * ld + hb] - i;
          const int mfirst = std::max(J0, i - hb);
          for (int m = mfirst; m < J0 + wJ; ++m) axpy_neg(ri + r0, &PT[(size_t)(m - J0) * (hb + W)], ri[m], i - r0 + 1);
        } else {
          double* rg = &S.border[(size_t)(q - (r1 - r0)) * n];
          for (int m = J0; m < J0 + wJ; ++m) axpy_neg(rg + r0, &PT[(size_t)(m - J0) * (hb + W)], rg[m], r1 - r0);
        }
      }
      bar.wait();
    }
  };
  if (T == 1) worker(0);
  else { std::vector<std::thread> th; for (int t = 0; t < T; ++t) th.emplace_back(worker, t); for (auto& t : th) t.join(); }
  if (!ok) return false;
  // corner: Cn - Lb Lb^T (the last row carries the right-hand side)
  const int ng = nb - 1;
  std::vector<double> C((size_t)nb * nb);
  for (int a = 0; a < nb; ++a)
    for (int b = 0; b <= a; ++b) {
      double s = S.corner[(size_t)a * nb + b];
      const double* ra = &S.border[(size_t)a * n]; const double* rb = &S.border[(size_t)b * n];
      for (int c = 0; c < n; ++c) s -= ra[c] * rb[c];
      C[(size_t)a * nb + b] = s;
    }
  for (int j = 0; j < ng; ++j) {   // Cholesky of the ng x ng part, the right-hand side row rides along
    double d = C[(size_t)j * nb + j];
    for (int m = 0; m < j; ++m) d -= C[(size_t)j * nb + m] * C[(size_t)j * nb + m];
    if (!(d > 0.0) || !std::isfinite(d)) return false;
    d = std::sqrt(d); C[(size_t)j * nb + j] = d;
    for (int i = j + 1; i < nb; ++i) {
      double s = C[(size_t)i * nb + j];
      for (int m = 0; m < j; ++m) s -= C[(size_t)i * nb + m] * C[(size_t)j * nb + m];
      C[(size_t)i * nb + j] = s / d;
    }
  }
  x.assign((size_t)n + ng, 0.0);
  for (int i = ng - 1; i >= 0; --i) {
    double s = C[(size_t)ng * nb + i];
    for (int k = i + 1; k < ng; ++k) s -= C[(size_t)k * nb + i] * x[n + k];
    x[n + i] = s / C[(size_t)i * nb + i];
  }
  // band back substitution: L^T x = y - Lb^T z, y = last border row
  std::vector<double> y(n);
  const double* yr = &S.border[(size_t)ng * n];
  for (int c = 0; c < n; ++c) {
    double s = yr[c];
    for (int g = 0; g < ng; ++g) s -= S.border[(size_t)g * n + c] * x[n + g];
    y[c] = s;
  }
  for (int i = n - 1; i >= 0; --i) {
    const double* ri = &S.band[(size_t)i * ld + hb] - i;
    const double xi = y[i] / ri[i];
    x[i] = xi;
    const int c0 = std::max(0, i - hb);
    for (int c = c0; c < i; ++c) y[c] -= ri[c] * xi;
  }
  return true;
}

}  // namespace sparse_detail

// Where each parameter block lives in the elimination order: landmarks (eliminated first), band blocks in time order, the rest = globals.
struct SparseSpec {
  int n_band = 0;
  const std::vector<double>* lm = nullptr;            // landmark table (3 doubles per landmark)
  std::function<int(const double*)> band_base;        // >= 0: first band index of the block at this address; -1: not a band block
};

// Sparse LM on a MiniProblem with the given elimination order.  Same constants, control flow and iteration log as solveLM (lm.h).
inline SolveSummary solveLMSparseSpec(MiniProblem& mp, const SparseSpec& spec, int max_num_iterations, int num_threads, SparseTimings* tm_out = nullptr,
                                      bool no_termination = false) {
  using namespace sparse_detail;
  const double initial_radius = 1e4, max_radius = 1e16, min_radius = 1e-32;
  const double min_relative_decrease = 1e-3, min_lm_diagonal = 1e-6, max_lm_diagonal = 1e32;
  const double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  const int max_consecutive_invalid = 5;
  SparseTimings tm;
  const double t_begin = now_s();
  const int ncol = mp.num_cols;
  const int T = std::max(1, num_threads);
  const size_t nblocks = mp.blocks.size();

  // ---- structure: tangent column -> unified index (band | border) or landmark ----
  static const std::vector<double> no_lm;
  const std::vector<double>& lmv = spec.lm ? *spec.lm : no_lm;
  const size_t L = lmv.size() / 3;
  auto in = [](const double* p, const std::vector<double>& v) { return !v.empty() && p >= v.data() && p < v.data() + v.size(); };
  tm.n_band = spec.n_band;
  const int n_band = tm.n_band;
  std::vector<int> uidx(ncol, -1);       // band / border index; <= -2 for landmark columns
  std::vector<int> lm_first_col(L, -1);
  int ng = 0;
  for (const auto& p : mp.params) {
    if (p.col < 0) continue;
    int base = spec.band_base(p.ptr);
    if (base < 0) {
      if (in(p.ptr, lmv)) { const size_t l = (p.ptr - lmv.data()) / 3; lm_first_col[l] = p.col; for (int k = 0; k < 3; ++k) uidx[p.col + k] = -2 - 3 * (int)l - k; continue; }
      base = n_band + ng; ng += p.lsize();
    }
    for (int k = 0; k < p.lsize(); ++k) uidx[p.col + k] = base + k;
  }
  tm.n_globals = ng; tm.n_landmarks = (int)L;
  std::vector<int> lm_of_col(ncol, -1);
  for (size_t l = 0; l < L; ++l) if (lm_first_col[l] >= 0) for (int k = 0; k < 3; ++k) lm_of_col[lm_first_col[l] + k] = (int)l;

  // per block: local column -> tangent column, split into dense (band / global) and landmark parts
  struct BlockMap { std::vector<int> dense_local, dense_col; std::vector<int> lm_local, lm_col; };
  std::vector<BlockMap> bmap(nblocks);
  std::vector<int> blk_lo(nblocks, 0), blk_hi(nblocks, 0);   // band column span of the block's dense part
  int hb = 0;
  std::vector<int> lm_lo(L, n_band), lm_hi(L, -1);
  for (size_t bi = 0; bi < nblocks; ++bi) {
    const auto& b = mp.blocks[bi];
    BlockMap& m = bmap[bi];
    int l0 = 0, lo = n_band, hi = -1;
    for (int id : b.pb) {
      const auto& p = mp.params[id];
      if (p.col >= 0) {
        if (uidx[p.col] <= -2) { m.lm_local.push_back(l0); m.lm_col.push_back(p.col); }
        else for (int k = 0; k < p.lsize(); ++k) {
          m.dense_local.push_back(l0 + k); m.dense_col.push_back(p.col + k);
          const int u = uidx[p.col + k];
          if (u < n_band) { lo = std::min(lo, u); hi = std::max(hi, u); }
        }
      } else if (b.kind == 0 && in(p.ptr, lmv)) { m.lm_local.push_back(l0); m.lm_col.push_back(-1); }   // keeps observation index == list position
      l0 += p.lsize();
    }
    blk_lo[bi] = lo; blk_hi[bi] = hi;
    if (hi >= lo) hb = std::max(hb, hi - lo);
    for (int c : m.lm_col) if (c >= 0) { const int l = lm_of_col[c]; lm_lo[l] = std::min(lm_lo[l], lo); lm_hi[l] = std::max(lm_hi[l], hi); }
  }
  for (size_t l = 0; l < L; ++l) if (lm_hi[l] >= lm_lo[l]) hb = std::max(hb, lm_hi[l] - lm_lo[l]);
  tm.half_bandwidth = hb;
  std::vector<std::vector<std::pair<int, int>>> lm_obs(L);   // landmark -> (block, observation)
  for (size_t bi = 0; bi < nblocks; ++bi)
    for (size_t o = 0; o < bmap[bi].lm_col.size(); ++o) if (bmap[bi].lm_col[o] >= 0) lm_obs[lm_of_col[bmap[bi].lm_col[o]]].push_back({(int)bi, (int)o});

  // ---- evaluation ----
  std::vector<CompactBlock> cb(nblocks);
  auto evaluate = [&](bool jac) -> double {
    const double t0 = now_s();
    parallel_dynamic(T, nblocks, [&](size_t bi, int) {
      const auto& b = mp.blocks[bi];
      EvalBlockOut o;
      mp.evalBlock(b, jac, &o);
      double c = 0.5 * o.sq_norm, sr = 1.0;
      if (b.cauchy_a > 0) { double rho[3]; cauchy(b.cauchy_a, o.sq_norm, rho); c = 0.5 * rho[0]; sr = std::sqrt(rho[1]); }
      CompactBlock& C = cb[bi];
      C.cost = c;
      if (!jac) return;
      const BlockMap& m = bmap[bi];
      C.nres = b.nres; C.nc = (int)m.dense_col.size(); C.cols = m.dense_col;
      C.r.resize(b.nres); for (int r = 0; r < b.nres; ++r) C.r[r] = sr * o.r[r];
      C.J.resize((size_t)b.nres * C.nc);
      for (int r = 0; r < b.nres; ++r) for (int k = 0; k < C.nc; ++k) C.J[(size_t)r * C.nc + k] = sr * o.jl[(size_t)r * o.ncols_local + m.dense_local[k]];
      const size_t no = m.lm_col.size();
      C.lmcol = m.lm_col; C.Jl.resize(no * 6);
      for (size_t ob = 0; ob < no; ++ob)
        for (int rr = 0; rr < 2; ++rr) for (int k = 0; k < 3; ++k) C.Jl[ob * 6 + rr * 3 + k] = sr * o.jl[(size_t)(2 * ob + rr) * o.ncols_local + m.lm_local[ob] + k];
    });
    double cost = 0; for (const auto& C : cb) cost += C.cost;
    (jac ? tm.eval_jac_s : tm.eval_res_s) += now_s() - t0; (jac ? tm.jac_evals : tm.res_evals)++;
    return cost;
  };
  auto grad_diag = [&](std::vector<double>& g, std::vector<double>& dg) {   // gradient and diag(J^T J) over all tangent columns
    g.assign(ncol, 0.0); dg.assign(ncol, 0.0);
    for (const auto& C : cb) {
      for (int r = 0; r < C.nres; ++r) for (int k = 0; k < C.nc; ++k) { const double v = C.J[(size_t)r * C.nc + k]; g[C.cols[k]] += v * C.r[r]; dg[C.cols[k]] += v * v; }
      for (size_t ob = 0; ob < C.lmcol.size(); ++ob) if (C.lmcol[ob] >= 0)
        for (int rr = 0; rr < 2; ++rr) for (int k = 0; k < 3; ++k) { const double v = C.Jl[ob * 6 + rr * 3 + k]; g[C.lmcol[ob] + k] += v * C.r[2 * ob + rr]; dg[C.lmcol[ob] + k] += v * v; }
    }
  };

  SolveSummary sum;
  double x_cost = evaluate(true);
  sum.initial_cost = x_cost;
  std::vector<double> g, dg;
  grad_diag(g, dg);
  std::vector<double> scale(ncol);
  for (int j = 0; j < ncol; ++j) scale[j] = 1.0 / (1.0 + std::sqrt(dg[j]));   // jacobi scaling, fixed for the whole Solve
  double x_norm = xNorm(mp), gmax = gradientMaxNorm(mp, g);
  sum.iterations.push_back({x_cost, 0.0, gmax, 0.0, 0.0, initial_radius, 1});
  auto finish = [&]() { sum.final_cost = x_cost; tm.total_s = now_s() - t_begin; if (tm_out) *tm_out = tm; return sum; };
  if (!no_termination && gmax <= gradient_tolerance) { sum.termination_type = 0; return finish(); }

  BandSystem S;
  const int nb = ng + 1;
  // chunks of the band for conflict-free parallel accumulation: a unit (block / landmark) whose span starts in chunk c touches chunks c, c+1 only
  const int chunk = std::max(hb + 1, (n_band + 4 * T - 1) / std::max(1, 4 * T));
  const int nchunks = std::max(1, (n_band + chunk - 1) / chunk);
  std::vector<std::vector<int>> chunk_blocks(nchunks), chunk_lms(nchunks);
  std::vector<int> nb_blocks;
  for (size_t bi = 0; bi < nblocks; ++bi) { if (blk_hi[bi] >= blk_lo[bi]) chunk_blocks[blk_lo[bi] / chunk].push_back((int)bi); else nb_blocks.push_back((int)bi); }
  for (size_t l = 0; l < L; ++l) if (!lm_obs[l].empty() && lm_hi[l] >= lm_lo[l]) chunk_lms[lm_lo[l] / chunk].push_back((int)l);
  std::vector<double> Vinv(L * 9), gl(L * 3);   // per landmark: (V + D)^-1, J_l^T r (scaled)
  std::vector<std::vector<int>> slot_scratch(T, std::vector<int>((size_t)n_band + ng, -1)), u_scratch(T);
  std::vector<std::vector<double>> w_scratch(T), y_scratch(T);

  double radius = initial_radius, decrease_factor = 2.0;
  int invalid = 0, iter = 0;
  while (true) {
    if (iter >= max_num_iterations) { sum.termination_type = 1; break; }
    if (radius < min_radius) { if (no_termination) radius = initial_radius; else { sum.termination_type = 0; break; } }
    ++iter;
    // ---- reduced system  (scaled J;  D = clamp(diag) / radius) ----
    double t0 = now_s();
    S.resize(n_band, hb, nb);
    std::vector<double> corner_part((size_t)T * nb * nb, 0.0);
    parallel_for(T, L, [&](size_t l0, size_t l1, int) {   // landmarks: V = sum J_l^T J_l + D, g_l
      for (size_t l = l0; l < l1; ++l) {
        if (lm_obs[l].empty()) continue;
        double V[9] = {0}, gv[3] = {0};
        const int c0 = lm_first_col[l];
        for (auto& bo : lm_obs[l]) {
          const CompactBlock& C = cb[bo.first];
          const double* J = &C.Jl[(size_t)bo.second * 6];
          for (int rr = 0; rr < 2; ++rr) {
            double js[3]; for (int k = 0; k < 3; ++k) js[k] = J[rr * 3 + k] * scale[c0 + k];
            for (int a = 0; a < 3; ++a) { gv[a] += js[a] * C.r[2 * bo.second + rr]; for (int b2 = 0; b2 < 3; ++b2) V[3 * a + b2] += js[a] * js[b2]; }
          }
        }
        for (int a = 0; a < 3; ++a) { V[4 * a] += std::min(std::max(V[4 * a], min_lm_diagonal), max_lm_diagonal) / radius; gl[3 * l + a] = gv[a]; }
        const double a = V[0], b = V[1], c = V[2], d = V[4], e = V[5], f = V[8];   // inverse of the symmetric 3x3
        const double co0 = d * f - e * e, co1 = c * e - b * f, co2 = b * e - c * d;
        const double id = 1.0 / (a * co0 + b * co1 + c * co2);
        double* I = &Vinv[9 * l];
        I[0] = co0 * id; I[1] = co1 * id; I[2] = co2 * id; I[3] = I[1]; I[4] = (a * f - c * c) * id; I[5] = (b * c - a * e) * id; I[6] = I[2]; I[7] = I[5]; I[8] = (a * d - b * b) * id;
      }
    });
    auto add_sys = [&](int tid, int ui, int uj, double v) {
      if (ui >= n_band && uj >= n_band) { const int r = std::max(ui, uj) - n_band, c = std::min(ui, uj) - n_band; corner_part[(size_t)tid * nb * nb + (size_t)r * nb + c] += v; }
      else S.add(ui, uj, v);
    };
    auto block_direct = [&](int bi, int tid) {   // J^T J and -J^T r of the band / global columns of one block
      const CompactBlock& C = cb[bi];
      std::vector<double> js(C.nc);
      for (int r = 0; r < C.nres; ++r) {
        for (int k = 0; k < C.nc; ++k) js[k] = C.J[(size_t)r * C.nc + k] * scale[C.cols[k]];
        for (int k = 0; k < C.nc; ++k) {
          if (js[k] == 0.0) continue;
          const int uk = uidx[C.cols[k]];
          if (uk < n_band) S.border[(size_t)ng * n_band + uk] -= js[k] * C.r[r]; else corner_part[(size_t)tid * nb * nb + (size_t)ng * nb + (uk - n_band)] -= js[k] * C.r[r];
          for (int k2 = 0; k2 <= k; ++k2) if (js[k2] != 0.0) add_sys(tid, uk, uidx[C.cols[k2]], js[k] * js[k2]);
        }
      }
    };
    auto landmark_schur = [&](int l, int tid) {
      // W = J_dense^T J_l (scaled) over the distinct band / global unknowns the landmark's observations touch, kept sorted by unified index
      // so that a row of the Schur update lands in one band row
      std::vector<int>& slot = slot_scratch[tid];
      std::vector<int>& U = u_scratch[tid]; U.clear();
      const int c0 = lm_first_col[l];
      for (auto& bo : lm_obs[l]) for (int col : cb[bo.first].cols) { const int u = uidx[col]; if (slot[u] < 0) { slot[u] = 0; U.push_back(u); } }
      std::sort(U.begin(), U.end());
      const size_t m = U.size();
      for (size_t q = 0; q < m; ++q) slot[U[q]] = (int)q;
      std::vector<double>& Wm = w_scratch[tid]; Wm.assign(m * 3, 0.0);
      for (auto& bo : lm_obs[l]) {
        const CompactBlock& C = cb[bo.first];
        const double* Jl = &C.Jl[(size_t)bo.second * 6];
        for (int rr = 0; rr < 2; ++rr) {
          const double* Jr = &C.J[(size_t)(2 * bo.second + rr) * C.nc];
          double jl[3]; for (int k = 0; k < 3; ++k) jl[k] = Jl[rr * 3 + k] * scale[c0 + k];
          for (int k = 0; k < C.nc; ++k) {
            const double v = Jr[k] * scale[C.cols[k]];
            if (v != 0.0) { double* w = &Wm[(size_t)slot[uidx[C.cols[k]]] * 3]; w[0] += v * jl[0]; w[1] += v * jl[1]; w[2] += v * jl[2]; }
          }
        }
      }
      const double* I = &Vinv[9 * (size_t)l];
      std::vector<double>& Y = y_scratch[tid]; Y.resize(m * 3);
      for (size_t q = 0; q < m; ++q) for (int a = 0; a < 3; ++a) Y[q * 3 + a] = Wm[q * 3] * I[a] + Wm[q * 3 + 1] * I[3 + a] + Wm[q * 3 + 2] * I[6 + a];
      size_t mb = 0; while (mb < m && U[mb] < n_band) ++mb;     // band unknowns first, globals after
      for (size_t q = 0; q < m; ++q) {
        const int uq = U[q];
        const double y0 = Y[q * 3], y1 = Y[q * 3 + 1], y2 = Y[q * 3 + 2];
        const double yg = y0 * gl[3 * l] + y1 * gl[3 * l + 1] + y2 * gl[3 * l + 2];   // rhs: -(g_p - W V^-1 g_l)
        if (uq < n_band) {
          S.border[(size_t)ng * n_band + uq] += yg;
          double* row = &S.band[(size_t)uq * (hb + 1) + hb] - uq;     // row[u2], u2 <= uq
          for (size_t q2 = 0; q2 <= q; ++q2) row[U[q2]] -= y0 * Wm[q2 * 3] + y1 * Wm[q2 * 3 + 1] + y2 * Wm[q2 * 3 + 2];
        } else {
          double* cp = &corner_part[(size_t)tid * nb * nb];
          cp[(size_t)ng * nb + (uq - n_band)] += yg;
          double* brow = &S.border[(size_t)(uq - n_band) * n_band];
          for (size_t q2 = 0; q2 < mb; ++q2) brow[U[q2]] -= y0 * Wm[q2 * 3] + y1 * Wm[q2 * 3 + 1] + y2 * Wm[q2 * 3 + 2];
          for (size_t q2 = mb; q2 <= q; ++q2) cp[(size_t)(uq - n_band) * nb + (U[q2] - n_band)] -= y0 * Wm[q2 * 3] + y1 * Wm[q2 * 3 + 1] + y2 * Wm[q2 * 3 + 2];
        }
      }
      for (size_t q = 0; q < m; ++q) slot[U[q]] = -1;
    };
    for (int colour = 0; colour < 2; ++colour) {   // two colours of chunks: units of one colour never touch the same band rows / border columns
      std::vector<int> mine;
      for (int c = colour; c < nchunks; c += 2) mine.push_back(c);
      parallel_dynamic(T, mine.size(), [&](size_t q, int tid) {
        const int c = mine[q];
        for (int bi : chunk_blocks[c]) block_direct(bi, tid);
        for (int l : chunk_lms[c]) landmark_schur(l, tid);
      });
    }
    for (int bi : nb_blocks) block_direct(bi, 0);   // blocks without band columns (gravity norm)
    for (int t = 0; t < T; ++t) for (int e = 0; e < nb * nb; ++e) S.corner[e] += corner_part[(size_t)t * nb * nb + e];
    // LM diagonal on band and global columns (D = clamp(diag(J_s^T J_s)) / radius, the Schur term is not part of it); band nodes that no
    // active parameter maps to (constant or unreferenced knots) get a unit diagonal so that the factorisation stays defined
    {
      std::vector<char> active(n_band + ng, 0);
      for (int j = 0; j < ncol; ++j) if (uidx[j] >= 0) active[uidx[j]] = 1;
      for (int u = 0; u < n_band; ++u) if (!active[u]) S.band[(size_t)u * (hb + 1) + hb] = 1.0;
      for (int j = 0; j < ncol; ++j) {
        const int u = uidx[j];
        if (u < 0) continue;
        const double d = std::min(std::max(dg[j] * scale[j] * scale[j], min_lm_diagonal), max_lm_diagonal) / radius;
        if (u < n_band) S.band[(size_t)u * (hb + 1) + hb] += d; else S.corner[(size_t)(u - n_band) * nb + (u - n_band)] += d;
      }
    }
    tm.build_s += now_s() - t0;
    t0 = now_s();
    std::vector<double> xs;
    bool ok = band_solve(S, xs, T);
    tm.factor_s += now_s() - t0; tm.solves++;
    std::vector<double> delta(ncol, 0.0);   // scaled step over all tangent columns
    double model_cost_change = 0.0;
    if (ok) {
      for (int j = 0; j < ncol; ++j) if (uidx[j] >= 0) delta[j] = xs[uidx[j]];
      parallel_for(T, L, [&](size_t l0, size_t l1, int) {   // landmarks: delta_l = V^-1 (-g_l - W^T delta_p)
        for (size_t l = l0; l < l1; ++l) {
          if (lm_obs[l].empty()) continue;
          const int c0 = lm_first_col[l];
          double rhs[3] = {-gl[3 * l], -gl[3 * l + 1], -gl[3 * l + 2]};
          for (auto& bo : lm_obs[l]) {
            const CompactBlock& C = cb[bo.first];
            const double* Jl = &C.Jl[(size_t)bo.second * 6];
            for (int rr = 0; rr < 2; ++rr) {
              const double* Jr = &C.J[(size_t)(2 * bo.second + rr) * C.nc];
              double jd = 0.0;
              for (int k = 0; k < C.nc; ++k) jd += Jr[k] * scale[C.cols[k]] * delta[C.cols[k]];
              for (int a = 0; a < 3; ++a) rhs[a] -= Jl[rr * 3 + a] * scale[c0 + a] * jd;
            }
          }
          const double* I = &Vinv[9 * l];
          for (int a = 0; a < 3; ++a) delta[c0 + a] = I[3 * a] * rhs[0] + I[3 * a + 1] * rhs[1] + I[3 * a + 2] * rhs[2];
        }
      });
      // model_cost_change = -delta^T (g_s + 1/2 H_s delta) = -sum_rows ((J_s delta) r + 1/2 (J_s delta)^2)
      double acc = 0.0;
      for (const auto& C : cb) {
        for (int r = 0; r < C.nres; ++r) {
          double jd = 0.0;
          for (int k = 0; k < C.nc; ++k) jd += C.J[(size_t)r * C.nc + k] * scale[C.cols[k]] * delta[C.cols[k]];
          if (!C.lmcol.empty()) { const int ob = r / 2, rr = r % 2, c0 = C.lmcol[ob]; if (c0 >= 0) for (int k = 0; k < 3; ++k) jd += C.Jl[(size_t)ob * 6 + rr * 3 + k] * scale[c0 + k] * delta[c0 + k]; }
          acc += jd * C.r[r] + 0.5 * jd * jd;
        }
      }
      model_cost_change = -acc;
      for (int j = 0; j < ncol; ++j) if (!std::isfinite(delta[j])) ok = false;
    }
    if (!ok || model_cost_change <= 0.0) {
      if (++invalid >= max_consecutive_invalid && !no_termination) { sum.termination_type = 2; break; }
      radius /= decrease_factor; decrease_factor *= 2.0;
      sum.num_unsuccessful_steps++;
      sum.iterations.push_back({x_cost, 0.0, gmax, 0.0, 0.0, radius, 0});
      continue;
    }
    invalid = 0;
    std::vector<double> du(ncol);
    for (int j = 0; j < ncol; ++j) du[j] = delta[j] * scale[j];
    std::vector<std::vector<double>> cand;
    plusAll(mp, du, &cand);
    double step_norm = 0;
    for (size_t i = 0; i < mp.params.size(); ++i) {
      const auto& p = mp.params[i];
      if (p.col < 0) continue;
      for (int k = 0; k < p.gsize; ++k) { const double d = cand[i][k] - p.ptr[k]; step_norm += d * d; }
    }
    step_norm = std::sqrt(step_norm);
    StateSwap sw(mp, cand);
    const double cand_cost = evaluate(false);
    const double cost_change = x_cost - cand_cost;
    const double ratio = cost_change / model_cost_change;
    if (!no_termination && step_norm <= parameter_tolerance * (x_norm + parameter_tolerance)) {
      sw.restore();
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
      sum.termination_type = 0; break;
    }
    if (!no_termination && std::fabs(cost_change) <= function_tolerance * x_cost) {
      sw.restore();
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
      sum.termination_type = 0; break;
    }
    if (ratio > min_relative_decrease) {
      x_cost = evaluate(true);
      grad_diag(g, dg);
      x_norm = xNorm(mp); gmax = gradientMaxNorm(mp, g);
      radius = std::min(max_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3)));
      decrease_factor = 2.0;
      sum.num_successful_steps++;
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 1});
      if (!no_termination && gmax <= gradient_tolerance) { sum.termination_type = 0; break; }
    } else {
      sw.restore();
      radius /= decrease_factor; decrease_factor *= 2.0;
      sum.num_unsuccessful_steps++;
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
    }
  }
  return finish();
}


// Sparse LM on a SplineProblem: spline / bias knots in time order form the band.
inline SolveSummary solveLMSparse(SplineProblem& sp, int max_num_iterations, int num_threads, SparseTimings* tm_out = nullptr, bool no_termination = false) {
  sp.build();
  MiniProblem& mp = sp.mp;
  const size_t K = sp.numKnots(), Kb = sp.numBiasKnots();
  auto in = [](const double* p, const std::vector<double>& v) { return !v.empty() && p >= v.data() && p < v.data() + v.size(); };
  // nodes in time order: spline knot i at i * dt, bias knot j at j * dt_bias (ties: spline knot first)
  std::vector<int> node_of_knot(K), node_of_bias(Kb);
  SparseSpec spec;
  {
    bool use_bias = false;
    for (const auto& p : mp.params) if (p.col >= 0 && (in(p.ptr, sp.acc_bias) || in(p.ptr, sp.gyr_bias))) use_bias = true;
    const size_t Kb_use = use_bias ? Kb : 0;
    size_t i = 0, j = 0; int node = 0;
    while (i < K || j < Kb_use) {
      const double ti = i < K ? (double)i * (double)sp.dt_ns : 1e300, tj = j < Kb_use ? (double)j * (double)sp.dt_ns_bias : 1e300;
      if (ti <= tj) node_of_knot[i++] = node++; else node_of_bias[j++] = node++;
    }
    spec.n_band = 6 * node;
  }
  spec.lm = &sp.lm;
  spec.band_base = [&](const double* ptr) -> int {
    if (in(ptr, sp.so3)) return 6 * node_of_knot[(ptr - sp.so3.data()) / 4];
    if (in(ptr, sp.r3)) return 6 * node_of_knot[(ptr - sp.r3.data()) / 3] + 3;
    if (in(ptr, sp.acc_bias)) return 6 * node_of_bias[(ptr - sp.acc_bias.data()) / 3];
    if (in(ptr, sp.gyr_bias)) return 6 * node_of_bias[(ptr - sp.gyr_bias.data()) / 3] + 3;
    return -1;
  };
  return solveLMSparseSpec(mp, spec, max_num_iterations, num_threads, tm_out, no_termination);
}

}  // namespace oracle
