// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// Trust-region Levenberg-Marquardt restating the behaviour of ceres::Solve() with the options the
// reference sets (include/spline/bspline_so3r3.h:1768-1782: SPARSE_NORMAL_CHOLESKY, max_num_iterations,
// everything else default).  [Ceres, recalled — see SURVEY.md Appendix B; Ceres is not in
// /root/reference, its version is unpinned (CMakeLists.txt:68) => "parity unpinned" at this boundary.]
// The normal equations are solved exactly (dense Cholesky here; SPARSE_NORMAL_CHOLESKY is also exact,
// so the step is the same up to rounding).
#pragma once
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>
#include "problem.h"

namespace oracle {

struct IterationLog { double cost, cost_change, gradient_max_norm, step_norm, tr_ratio, tr_radius; int successful; };

struct SolveSummary {
  int termination_type = 1;  // 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE (ceres::TerminationType)
  double initial_cost = 0, final_cost = 0;
  int num_successful_steps = 0, num_unsuccessful_steps = 0;
  std::vector<IterationLog> iterations;
};

struct DenseSystem {
  int n = 0;
  std::vector<double> H;  // n x n (lower+upper), J^T J with loss-corrected J
  std::vector<double> g;  // J^T r
  double cost = 0;
};

// Plus for every active block: SO3 -> q * exp(delta) (lie_local_parametrization.h:34-41), else x + delta
inline void plusAll(const MiniProblem& mp, const std::vector<double>& delta, std::vector<std::vector<double>>* out) {
  out->resize(mp.params.size());
  for (size_t i = 0; i < mp.params.size(); ++i) {
    const auto& p = mp.params[i];
    auto& o = (*out)[i];
    o.assign(p.ptr, p.ptr + p.gsize);
    if (p.col < 0) continue;
    if (p.so3) {
      SO3<double> q{p.ptr[0], p.ptr[1], p.ptr[2], p.ptr[3]};
      SO3<double> e = SO3<double>::exp(V3<double>{delta[p.col], delta[p.col + 1], delta[p.col + 2]});
      SO3<double> r = q * e;
      o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = r.w;
    } else if (p.se3) {
      se3_plus_fwd(p.ptr, &delta[p.col], o.data());
    } else {
      for (int k = 0; k < p.gsize; ++k) o[k] = p.ptr[k] + delta[p.col + k];
    }
  }
}

struct StateSwap {
  // temporarily writes candidate values into the user-owned arrays
  const MiniProblem& mp; std::vector<std::vector<double>> saved;
  StateSwap(const MiniProblem& m, const std::vector<std::vector<double>>& cand) : mp(m) {
    saved.resize(mp.params.size());
    for (size_t i = 0; i < mp.params.size(); ++i) {
      const auto& p = mp.params[i];
      saved[i].assign(p.ptr, p.ptr + p.gsize);
      std::memcpy(p.ptr, cand[i].data(), sizeof(double) * p.gsize);
    }
  }
  void restore() { for (size_t i = 0; i < mp.params.size(); ++i) std::memcpy(mp.params[i].ptr, saved[i].data(), sizeof(double) * mp.params[i].gsize); }
};

// cost = 1/2 sum rho(||r_b||^2); optionally the dense normal equations with the corrector applied.
inline double evaluateSystem(const MiniProblem& mp, DenseSystem* sys, int num_threads = 1) {
  const int n = mp.num_cols;
  const size_t nb = mp.blocks.size();
  if (sys) { sys->n = n; sys->H.assign((size_t)n * n, 0.0); sys->g.assign(n, 0.0); }
  std::vector<double> costs(nb, 0.0);
  std::vector<EvalBlockOut> outs(sys ? nb : 0);
  auto work = [&](size_t b0, size_t b1) {
    EvalBlockOut tmp;
    for (size_t bi = b0; bi < b1; ++bi) {
      const auto& b = mp.blocks[bi];
      EvalBlockOut& o = sys ? outs[bi] : tmp;
      mp.evalBlock(b, sys != nullptr, &o);
      double s = o.sq_norm;
      if (b.cauchy_a > 0) {
        double rho[3]; cauchy(b.cauchy_a, s, rho);
        costs[bi] = 0.5 * rho[0];
        if (sys) {
          const double sr = std::sqrt(rho[1]);  // rho'' < 0 => residual and Jacobian scaled by sqrt(rho')
          for (auto& v : o.r) v *= sr;
          for (auto& v : o.jl) v *= sr;
        }
      } else {
        costs[bi] = 0.5 * s;
      }
    }
  };
  if (num_threads <= 1 || nb < 8) {
    work(0, nb);
  } else {
    std::vector<std::thread> th;
    for (int t = 0; t < num_threads; ++t) th.emplace_back(work, nb * t / num_threads, nb * (t + 1) / num_threads);
    for (auto& t : th) t.join();
  }
  double cost = 0;
  for (double c : costs) cost += c;
  if (sys) {
    sys->cost = cost;
    for (size_t bi = 0; bi < nb; ++bi) {
      const auto& b = mp.blocks[bi];
      const auto& o = outs[bi];
      // local column map of this block
      std::vector<int> cols(o.ncols_local, -1);
      int l0 = 0;
      for (int id : b.pb) {
        const auto& p = mp.params[id];
        for (int k = 0; k < p.lsize(); ++k) cols[l0 + k] = p.col < 0 ? -1 : p.col + k;
        l0 += p.lsize();
      }
      // compress to active, non-zero columns
      std::vector<int> act;
      for (int c = 0; c < o.ncols_local; ++c) if (cols[c] >= 0) act.push_back(c);
      for (int r = 0; r < b.nres; ++r) {
        const double* row = &o.jl[(size_t)r * o.ncols_local];
        std::vector<int> nz;
        for (int c : act) if (row[c] != 0.0) nz.push_back(c);
        for (int c1 : nz) {
          sys->g[cols[c1]] += row[c1] * o.r[r];
          double* Hrow = &sys->H[(size_t)cols[c1] * n];
          for (int c2 : nz) Hrow[cols[c2]] += row[c1] * row[c2];
        }
      }
    }
  }
  return cost;
}

// dense in-place lower Cholesky; returns false if not positive definite
inline bool cholesky(std::vector<double>& A, int n) {
  for (int j = 0; j < n; ++j) {
    double d = A[(size_t)j * n + j];
    for (int k = 0; k < j; ++k) d -= A[(size_t)j * n + k] * A[(size_t)j * n + k];
    if (!(d > 0.0)) return false;
    d = std::sqrt(d);
    A[(size_t)j * n + j] = d;
    for (int i = j + 1; i < n; ++i) {
      double s = A[(size_t)i * n + j];
      const double* ai = &A[(size_t)i * n];
      const double* aj = &A[(size_t)j * n];
      for (int k = 0; k < j; ++k) s -= ai[k] * aj[k];
      A[(size_t)i * n + j] = s / d;
    }
  }
  return true;
}
inline void cholSolve(const std::vector<double>& L, int n, std::vector<double>& b) {
  for (int i = 0; i < n; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= L[(size_t)i * n + k] * b[k]; b[i] = s / L[(size_t)i * n + i]; }
  for (int i = n - 1; i >= 0; --i) { double s = b[i]; for (int k = i + 1; k < n; ++k) s -= L[(size_t)k * n + i] * b[k]; b[i] = s / L[(size_t)i * n + i]; }
}

inline double xNorm(const MiniProblem& mp) {
  double s = 0;
  for (const auto& p : mp.params) if (p.col >= 0) for (int k = 0; k < p.gsize; ++k) s += p.ptr[k] * p.ptr[k];
  return std::sqrt(s);
}
inline double gradientMaxNorm(const MiniProblem& mp, const std::vector<double>& g) {
  // max | x - Plus(x, -g) | over the ambient coordinates
  std::vector<double> neg(g.size());
  for (size_t i = 0; i < g.size(); ++i) neg[i] = -g[i];
  std::vector<std::vector<double>> xp;
  plusAll(mp, neg, &xp);
  double m = 0;
  for (size_t i = 0; i < mp.params.size(); ++i) {
    const auto& p = mp.params[i];
    if (p.col < 0) continue;
    for (int k = 0; k < p.gsize; ++k) m = std::max(m, std::fabs(xp[i][k] - p.ptr[k]));
  }
  return m;
}

inline SolveSummary solveLM(MiniProblem& mp, int max_num_iterations, int num_threads = 1) {
  // ceres::Solver::Options defaults [Ceres, recalled]
  const double initial_radius = 1e4, max_radius = 1e16, min_radius = 1e-32;
  const double min_relative_decrease = 1e-3, min_lm_diagonal = 1e-6, max_lm_diagonal = 1e32;
  const double function_tolerance = 1e-6, gradient_tolerance = 1e-10, parameter_tolerance = 1e-8;
  const int max_consecutive_invalid = 5;

  SolveSummary sum;
  const int n = mp.num_cols;
  DenseSystem sys;
  double x_cost = evaluateSystem(mp, &sys, num_threads);
  sum.initial_cost = x_cost;
  // jacobi scaling, fixed for the whole Solve
  std::vector<double> scale(n);
  for (int j = 0; j < n; ++j) scale[j] = 1.0 / (1.0 + std::sqrt(sys.H[(size_t)j * n + j]));
  double x_norm = xNorm(mp);
  double gmax = gradientMaxNorm(mp, sys.g);
  sum.iterations.push_back({x_cost, 0.0, gmax, 0.0, 0.0, initial_radius, 1});
  if (gmax <= gradient_tolerance) { sum.termination_type = 0; sum.final_cost = x_cost; return sum; }

  double radius = initial_radius, decrease_factor = 2.0;
  int invalid = 0;
  int iter = 0;
  while (true) {
    if (iter >= max_num_iterations) { sum.termination_type = 1; break; }
    if (radius < min_radius) { sum.termination_type = 0; break; }
    ++iter;
    // scaled system: Hs = S H S, gs = S g
    std::vector<double> A((size_t)n * n), rhs(n);
    for (int i = 0; i < n; ++i) {
      for (int j = 0; j < n; ++j) A[(size_t)i * n + j] = sys.H[(size_t)i * n + j] * scale[i] * scale[j];
      rhs[i] = -sys.g[i] * scale[i];
    }
    std::vector<double> Hs_diag(n);
    for (int i = 0; i < n; ++i) Hs_diag[i] = A[(size_t)i * n + i];
    std::vector<double> As = A;  // keep Hs for the model cost
    for (int i = 0; i < n; ++i) {
      double d = std::min(std::max(Hs_diag[i], min_lm_diagonal), max_lm_diagonal);
      A[(size_t)i * n + i] += d / radius;
    }
    bool ok = cholesky(A, n);
    std::vector<double> delta = rhs;
    double model_cost_change = 0;
    if (ok) {
      cholSolve(A, n, delta);
      // model_cost_change = -delta^T (gs + 1/2 Hs delta)
      double acc = 0;
      for (int i = 0; i < n; ++i) {
        double hd = 0;
        const double* row = &As[(size_t)i * n];
        for (int j = 0; j < n; ++j) hd += row[j] * delta[j];
        acc += delta[i] * (-rhs[i] + 0.5 * hd);
      }
      model_cost_change = -acc;
      for (int i = 0; i < n; ++i) if (!std::isfinite(delta[i])) ok = false;
    }
    if (!ok || model_cost_change <= 0.0) {
      // HandleInvalidStep [Ceres, recalled]: the max_num_consecutive_invalid_steps-th invalid step in a row ends the solve with FAILURE
      if (++invalid >= max_consecutive_invalid) { sum.termination_type = 2; break; }
      radius /= decrease_factor; decrease_factor *= 2.0;
      sum.num_unsuccessful_steps++;
      sum.iterations.push_back({x_cost, 0.0, gmax, 0.0, 0.0, radius, 0});
      continue;
    }
    invalid = 0;
    std::vector<double> du(n);
    for (int i = 0; i < n; ++i) du[i] = delta[i] * scale[i];
    std::vector<std::vector<double>> cand;
    plusAll(mp, du, &cand);
    double step_norm = 0;
    for (size_t i = 0; i < mp.params.size(); ++i) {
      const auto& p = mp.params[i];
      if (p.col < 0) continue;
      for (int k = 0; k < p.gsize; ++k) { double d = cand[i][k] - p.ptr[k]; step_norm += d * d; }
    }
    step_norm = std::sqrt(step_norm);
    StateSwap sw(mp, cand);
    double cand_cost = evaluateSystem(mp, nullptr, num_threads);
    const double cost_change = x_cost - cand_cost;
    const double ratio = cost_change / model_cost_change;
    // ParameterToleranceReached / FunctionToleranceReached are tested BEFORE accept/reject (Ceres 2.x order)
    if (step_norm <= parameter_tolerance * (x_norm + parameter_tolerance)) {
      sw.restore();
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
      sum.termination_type = 0; break;
    }
    if (std::fabs(cost_change) <= function_tolerance * x_cost) {
      sw.restore();
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
      sum.termination_type = 0; break;
    }
    if (ratio > min_relative_decrease) {
      // accepted: candidate stays in the user arrays
      x_cost = evaluateSystem(mp, &sys, num_threads);
      x_norm = xNorm(mp);
      gmax = gradientMaxNorm(mp, sys.g);
      radius = std::min(max_radius, radius / std::max(1.0 / 3.0, 1.0 - std::pow(2.0 * ratio - 1.0, 3)));
      decrease_factor = 2.0;
      sum.num_successful_steps++;
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 1});
      if (gmax <= gradient_tolerance) { sum.termination_type = 0; break; }
    } else {
      sw.restore();
      radius /= decrease_factor; decrease_factor *= 2.0;
      sum.num_unsuccessful_steps++;
      sum.iterations.push_back({x_cost, cost_change, gmax, step_norm, ratio, radius, 0});
    }
  }
  sum.final_cost = x_cost;
  return sum;
}

}  // namespace oracle
