// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// (1) MiniProblem: the subset of ceres::Problem semantics the reference relies on
//     [Ceres, recalled — Ceres is absent from /root/reference; parity with real Ceres is unpinned]:
//     parameter blocks identified by raw pointer, optional SO3 local parameterisation
//     (include/optimization/lie_local_parametrization.h:34-55), SetParameterBlockConstant,
//     residual blocks with optional CauchyLoss, stride-4 DynamicAutoDiff evaluation.
// (2) SplineProblem: restatement of BSplineSO3R3<N> problem assembly
//     (include/spline/bspline_so3r3.h:75-117, 355-399, 523-1046, 1213-1227, 1784-1875).
#pragma once
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <functional>
#include <map>
#include <memory>
#include <string>
#include <thread>
#include <vector>
#include "functors.h"

namespace oracle {

// SE3 plus / plus-Jacobian on the Sophus layout (defined in dt_problem.h)
void se3_plus_fwd(const double* T, const double* d, double* out);
void se3_plus_jacobian_fwd(const double* T, double J[7][6]);

struct ParamBlock {
  double* ptr;
  int gsize;
  bool so3;
  bool constant;
  int lsize() const { return so3 ? 3 : (se3 ? 6 : gsize); }
  int col = -1;  // tangent-space column offset among ACTIVE blocks (assigned by finalize)
  bool se3 = false;   // Sophus SE3 block [q, t] with LieLocalParameterization<SE3d> (discrete-time estimator, dt_problem.h)
};

struct ResidualBlock {
  int kind;   // 0 reproj frame, 1 accel, 2 gyro, 3 gp, 4 bias, 5 gravity, 6 pose prior, 7 world alignment
  int index;  // index inside its kind
  int nres;
  double cauchy_a;  // 0 => no loss
  std::vector<int> pb;
  // (blocks, residuals, jac_global or null)
  std::function<void(const std::vector<const double*>&, const std::vector<int>&, double*, double*)> eval;
};

struct EvalBlockOut {
  std::vector<double> r;          // nres (raw, before loss correction)
  std::vector<double> jl;         // nres x sum(lsize over ALL blocks of the residual, incl. constant) row-major
  int ncols_local = 0;
  double sq_norm = 0.0;
};

struct MiniProblem {
  std::vector<ParamBlock> params;
  std::map<const double*, int> by_ptr;
  std::vector<ResidualBlock> blocks;
  int num_cols = 0;

  int AddParameterBlock(double* ptr, int gsize, bool so3 = false) {
    auto it = by_ptr.find(ptr);
    if (it != by_ptr.end()) {
      if (so3) params[it->second].so3 = true;
      return it->second;
    }
    int id = (int)params.size();
    params.push_back({ptr, gsize, so3, false});
    by_ptr[ptr] = id;
    return id;
  }
  int AddParameterBlockSE3(double* ptr) {
    const int id = AddParameterBlock(ptr, 7, false);
    params[id].se3 = true;
    return id;
  }
  void SetParameterBlockConstant(double* ptr) { params[by_ptr.at(ptr)].constant = true; }

  // assign tangent columns to the non-constant blocks that appear in at least one residual block,
  // in order of first appearance (Ceres orders by program construction; any fixed order is fine
  // because the normal equations are solved exactly).
  void finalize() {
    for (auto& p : params) p.col = -1;
    num_cols = 0;
    for (auto& b : blocks)
      for (int id : b.pb) {
        auto& p = params[id];
        if (!p.constant && p.col < 0) { p.col = num_cols; num_cols += p.lsize(); }
      }
  }

  // raw residuals + local Jacobian of one residual block (all its parameter blocks, in block order)
  void evalBlock(const ResidualBlock& b, bool want_jac, EvalBlockOut* out) const {
    std::vector<const double*> ptrs;
    std::vector<int> sizes;
    int gtot = 0, ltot = 0;
    for (int id : b.pb) { ptrs.push_back(params[id].ptr); sizes.push_back(params[id].gsize); gtot += params[id].gsize; ltot += params[id].lsize(); }
    out->r.assign(b.nres, 0.0);
    out->ncols_local = ltot;
    if (!want_jac) {
      b.eval(ptrs, sizes, out->r.data(), nullptr);
    } else {
      std::vector<double> jg((size_t)b.nres * gtot);
      b.eval(ptrs, sizes, out->r.data(), jg.data());
      out->jl.assign((size_t)b.nres * ltot, 0.0);
      int g0 = 0, l0 = 0;
      for (int id : b.pb) {
        const auto& p = params[id];
        if (p.so3) {
          // J_local = J_global (r x 4) * Dx_this_mul_exp_x_at_0 (4 x 3)
          SO3<double> q{p.ptr[0], p.ptr[1], p.ptr[2], p.ptr[3]};
          double PJ[4][3];
          q.Dx_this_mul_exp_x_at_0(PJ);
          for (int r = 0; r < b.nres; ++r)
            for (int c = 0; c < 3; ++c) {
              double s = 0;
              for (int k = 0; k < 4; ++k) s += jg[(size_t)r * gtot + g0 + k] * PJ[k][c];
              out->jl[(size_t)r * ltot + l0 + c] = s;
            }
        } else if (p.se3) {
          // J_local = J_global (r x 7) * SE3::Dx_this_mul_exp_x_at_0 (7 x 6), se3.hpp:135-204
          double PJ[7][6];
          se3_plus_jacobian_fwd(p.ptr, PJ);
          for (int r = 0; r < b.nres; ++r)
            for (int c = 0; c < 6; ++c) {
              double s = 0;
              for (int k = 0; k < 7; ++k) s += jg[(size_t)r * gtot + g0 + k] * PJ[k][c];
              out->jl[(size_t)r * ltot + l0 + c] = s;
            }
        } else {
          for (int r = 0; r < b.nres; ++r)
            for (int c = 0; c < p.gsize; ++c) out->jl[(size_t)r * ltot + l0 + c] = jg[(size_t)r * gtot + g0 + c];
        }
        g0 += p.gsize; l0 += p.lsize();
      }
    }
    double s = 0;
    for (double v : out->r) s += v * v;
    out->sq_norm = s;
  }
};

// CauchyLoss(a) [Ceres, recalled]: rho(s) = b log(1 + s/b), b = a^2.
inline void cauchy(double a, double s, double rho[3]) {
  const double b = a * a, c = 1.0 / b;
  const double sum = 1.0 + s * c, inv = 1.0 / sum;
  rho[0] = b * std::log(sum);
  rho[1] = std::max(std::numeric_limits<double>::min(), inv);
  rho[2] = -c * (inv * inv);
}

// ---------------------------------------------------------------------------------
struct FrameMeas {
  int64_t s; double u; double inv_std;
  std::vector<double> uv; std::vector<int> lm;
  int64_t obs_offset;  // offset of the frame's first observation in the caller's flat arrays
};
struct ImuMeas { int64_t s; double st_s, k; int64_t s_bias; double u_bias; double meas[3]; double inv_std; };
struct GpMeas { int64_t s; double st_s, k; double meas[3]; double inv_std; };
struct PriorMeas { int64_t s; double u; int which; double meas[4]; double inv_std; };   // which: 0 position, 1 orientation
struct BiasMeas { int64_t s0, s1; double u0, u1, inv_std; int which; /*0 acc 1 gyro*/ };

struct SplineProblem {
  int N;
  static constexpr int Nb = 4;
  int64_t dt_ns, start_t_ns;
  double dt_s, inv_dt;
  // state (bspline_so3r3.h:1936-1973)
  std::vector<double> so3, r3;            // 4K, 3K
  std::vector<double> acc_bias, gyr_bias; // 3Kb each
  std::vector<double> lm;                 // 3L
  double g[3] = {0, 0, 0}, p_bp[3] = {0, 0, 0};
  double R_bc_corr[4] = {0, 0, 0, 1}, p_bc_corr[3] = {0, 0, 0};
  double t_offset_cam_imu_corr_s = 0.0, t_s_gp_corr_s = 0.0, scale_corr = 0.0;
  double scale = 0.0, t_offset_cam_imu_s = 0.0, t_s_gp_s = 0.0, t_s_gp_init_s = 0.0;
  bool optimize_p_bp = true, optimize_gravity = false, optimize_scale = false,
       optimize_camimu_extrinsics = true, optimize_t_offset_s_gs = true;  // ctor defaults :94-98
  int64_t dt_ns_bias = 1000000000; double inv_dt_bias = 1.0;
  CameraModel cam; double cam_t_offset_cam_imu = 0.0;
  double cauchy_a = 3.0;  // bspline_so3r3.h:1032

  std::vector<FrameMeas> frames;
  std::vector<ImuMeas> accel, gyro;
  std::vector<GpMeas> gp;
  std::vector<PriorMeas> priors;
  std::vector<GpMeas> align;              // addWorldColmapAlignmentMeasurement
  double T_wg[7] = {0, 0, 0, 1, 0, 0, 0};   // T_w_g_ (q x y z w | p), :1968
  double R_wg_corr[4] = {0, 0, 0, 1}, p_wg_corr[3] = {0, 0, 0};
  double align_cauchy_a = 0.10;           // :1189
  std::vector<BiasMeas> bias;
  bool gravity_norm = false;

  MiniProblem mp;
  bool built = false;

  SplineProblem(int order, int64_t dt, int64_t start) : N(order), dt_ns(dt), start_t_ns(start) {
    dt_s = dt_ns * 1e-9;     // :78
    inv_dt = 1e9 / dt_ns;    // :79
  }
  size_t numKnots() const { return so3.size() / 4; }
  size_t numBiasKnots() const { return acc_bias.size() / 3; }

  // :355-369
  void initializeBiases(double dt_s_bias, const double* a0, const double* g0) {
    dt_ns_bias = (int64_t) static_cast<uint64_t>(dt_s_bias / 1e-9);
    inv_dt_bias = 1e9 / dt_ns_bias;
    double d = dt_s * (double)(r3.size() / 3);
    int num_knots = (int)(d / dt_s_bias);
    acc_bias.resize(3 * (size_t)num_knots); gyr_bias.resize(3 * (size_t)num_knots);
    for (int i = 0; i < num_knots; ++i) for (int c = 0; c < 3; ++c) { acc_bias[3 * i + c] = a0[c]; gyr_bias[3 * i + c] = g0[c]; }
  }
  // :1870-1875 (N_, not Nb_)
  int64_t getEndTnsBiasSpline() const {
    int a = (int)numBiasKnots() - N + 1;
    int64_t t_end_ns = (int64_t)a * dt_ns_bias + start_t_ns;
    return (int64_t)((double)t_end_ns - 1e6);   // `t_end_ns - 1e6` is evaluated in double and converted back (:1874)
  }

  // :523-624 / :626-702 (window mapping only; the residual blocks are wired in build())
  bool mapImu(int64_t time_ns, ImuMeas* m) const {
    int64_t t_offset_imu_cam_ns = static_cast<int64_t>(-t_offset_cam_imu_s * 1e9);
    int64_t st_ns = (time_ns + t_offset_imu_cam_ns - start_t_ns);
    double st_s = st_ns * 1e-9;
    if (st_ns < 0) return false;
    int64_t s = st_ns / dt_ns;
    int k_int = (int)(st_ns / dt_ns);
    double k = static_cast<double>(k_int);
    if (size_t(s + N) > numKnots()) return false;
    int64_t st_bias_ns = (time_ns - start_t_ns);
    int64_t s_bias = st_bias_ns / dt_ns_bias;
    double u_bias = double(st_bias_ns % dt_ns_bias) / double(dt_ns_bias);
    if (size_t(s_bias + N) > numBiasKnots()) {
      st_bias_ns = getEndTnsBiasSpline() - start_t_ns;
      s_bias = st_bias_ns / dt_ns_bias;
      u_bias = double(st_bias_ns % dt_ns_bias) / double(dt_ns_bias);
    }
    m->s = s; m->st_s = st_s; m->k = k; m->s_bias = s_bias; m->u_bias = u_bias;
    return true;
  }
  int addImu(int n, const int64_t* t_ns, const double* acc, const double* gyr, double acc_nd, double gyr_nd) {
    int added = 0;
    for (int i = 0; i < n; ++i) {
      ImuMeas m;
      if (!mapImu(t_ns[i], &m)) continue;
      ImuMeas a = m, w = m;
      for (int c = 0; c < 3; ++c) { a.meas[c] = acc[3 * i + c]; w.meas[c] = gyr[3 * i + c]; }
      a.inv_std = 1.0 / acc_nd; w.inv_std = 1.0 / gyr_nd;
      accel.push_back(a); gyro.push_back(w); ++added;
    }
    built = false;
    return added;
  }
  // :880-958
  int addGp(int n, const int64_t* t_ns, const double* pos, double noise) {
    int added = 0;
    for (int i = 0; i < n; ++i) {
      int64_t t_s_gp_init_ns = static_cast<int64_t>(t_s_gp_s * 1e9);
      int64_t st_ns = (t_ns[i] - start_t_ns - t_s_gp_init_ns);
      double st_s = st_ns * 1e-9;
      if (st_ns < 0) continue;
      int64_t s = st_ns / dt_ns;
      int k_int = (int)(st_ns / dt_ns);
      if (size_t(s + N) > numKnots()) continue;
      GpMeas m; m.s = s; m.st_s = st_s; m.k = (double)k_int; m.inv_std = 1.0 / noise;
      for (int c = 0; c < 3; ++c) m.meas[c] = pos[3 * i + c];
      gp.push_back(m); ++added;
    }
    built = false;
    return added;
  }
  // :960-1046
  int addFrames(int F, const int64_t* t_ns, const int32_t* obs_begin, const double* uv, const int32_t* lmi, double noise) {
    int added = 0;
    for (int f = 0; f < F; ++f) {
      int64_t st_ns = (t_ns[f] - start_t_ns);
      if (st_ns < 0) continue;
      int n = obs_begin[f + 1] - obs_begin[f];
      if (n == 0) continue;
      int64_t s = st_ns / dt_ns;
      double u = double(st_ns % dt_ns) / double(dt_ns);
      if (size_t(s + N) > numKnots()) continue;
      FrameMeas m; m.s = s; m.u = u; m.inv_std = 1.0 / noise; m.obs_offset = obs_begin[f];
      m.uv.assign(uv + 2 * (size_t)obs_begin[f], uv + 2 * (size_t)obs_begin[f + 1]);
      m.lm.assign(lmi + obs_begin[f], lmi + obs_begin[f + 1]);
      frames.push_back(std::move(m)); ++added;
    }
    built = false;
    return added;
  }
  // :706-792
  int addBiasCost(double freq_hz, double bias_err_dt_s, double acc_rw, double gyr_rw) {
    double err_freq_s = 1.0 / freq_hz;
    double err_freq_ns = (double) static_cast<uint64_t>(err_freq_s / 1e-9);
    int64_t t_start_ns = start_t_ns;
    int64_t t_end_ns = getEndTnsBiasSpline();
    int64_t bias_err_dt_ns = static_cast<int64_t>(bias_err_dt_s * 1e9);
    int added = 0;
    for (int64_t t_ns = t_start_ns; t_ns < t_end_ns; t_ns += err_freq_ns) {
      int64_t st_ns_0 = (t_ns - start_t_ns);
      int64_t st_ns_1 = (t_ns + bias_err_dt_ns - start_t_ns);
      int64_t s_0 = st_ns_0 / dt_ns_bias, s_1 = st_ns_1 / dt_ns_bias;
      double u_0 = double(st_ns_0 % dt_ns_bias) / double(dt_ns_bias);
      double u_1 = double(st_ns_1 % dt_ns_bias) / double(dt_ns_bias);
      if (size_t(s_0 + Nb) > numBiasKnots()) continue;
      if (size_t(s_1 + Nb) > numBiasKnots()) continue;
      bias.push_back({s_0, s_1, u_0, u_1, 1.0 / (0.001 * acc_rw), 0});
      bias.push_back({s_0, s_1, u_0, u_1, 1.0 / (0.001 * gyr_rw), 1});
      ++added;
    }
    built = false;
    return added;
  }

  // addPositionMeasurement :1048-1085 / addOrientationMeasurement :1087-1123
  int addPriors(int which, int n, const int64_t* t_ns, const double* meas, double noise) {
    int added = 0;
    const int w = which ? 4 : 3;
    for (int i = 0; i < n; ++i) {
      int64_t st_ns = (t_ns[i] - start_t_ns);
      if (st_ns < 0) continue;
      int64_t s = st_ns / dt_ns;
      double u = double(st_ns % dt_ns) / double(dt_ns);
      if (size_t(s + N) > numKnots()) continue;
      PriorMeas m; m.s = s; m.u = u; m.which = which; m.inv_std = 1.0 / noise;
      for (int c = 0; c < 4; ++c) m.meas[c] = c < w ? meas[(size_t)w * i + c] : 0.0;
      priors.push_back(m); ++added;
    }
    built = false;
    return added;
  }

  // addWorldColmapAlignmentMeasurement :1121-1211 (same window arithmetic as addGlobalPosMeasurement)
  int addAlign(int n, const int64_t* t_ns, const double* pos, double noise) {
    int added = 0;
    for (int i = 0; i < n; ++i) {
      int64_t t_s_gp_init_ns = static_cast<int64_t>(t_s_gp_s * 1e9);
      int64_t st_ns = (t_ns[i] - start_t_ns - t_s_gp_init_ns);
      double st_s = st_ns * 1e-9;
      if (st_ns < 0) continue;
      int64_t s = st_ns / dt_ns;
      int k_int = (int)(st_ns / dt_ns);
      if (size_t(s + N) > numKnots()) continue;
      GpMeas m; m.s = s; m.st_s = st_s; m.k = (double)k_int; m.inv_std = 1.0 / noise;
      for (int c = 0; c < 3; ++c) m.meas[c] = pos[3 * i + c];
      align.push_back(m); ++added;
    }
    built = false;
    return added;
  }

  // Wire everything into the MiniProblem in the reference driver's order
  // (src/optimize_continuous_time.cpp:497, 526, 567-601): gravity norm, SO3 parameterisations,
  // GP, accel+gyro interleaved, frames, bias cost.
  template <int NN>
  void buildT() {
    mp = MiniProblem();
    const size_t K = numKnots();
    if (gravity_norm) {
      ResidualBlock b; b.kind = 5; b.index = 0; b.nres = 1; b.cauchy_a = 0;
      b.pb = {mp.AddParameterBlock(g, 3)};
      GravityNormCostFunctor f;
      b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 1, r, j); };
      mp.blocks.push_back(std::move(b));
    }
    for (size_t i = 0; i < K; ++i) mp.AddParameterBlock(&so3[4 * i], 4, true);
    mp.AddParameterBlock(R_bc_corr, 4, true);
    for (size_t i = 0; i < priors.size(); ++i) {   // fit_spline_to_colmap.cpp:182-192: after the parameterisations
      const PriorMeas& m = priors[i];
      ResidualBlock b; b.kind = 6; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
      if (m.which == 0) {
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&r3[3 * (m.s + j)], 3));
        PositionCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2]}, m.u, inv_dt, m.inv_std};
        b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      } else {
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
        OrientationCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2], m.meas[3]}, m.u, inv_dt, m.inv_std};
        b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      }
      mp.blocks.push_back(std::move(b));
    }
    if (!align.empty()) mp.AddParameterBlock(R_wg_corr, 4, true);   // addRotationLocalParametrization (align_spline_to_global_frame.cpp:194)
    for (size_t i = 0; i < align.size(); ++i) {
      const GpMeas& m = align[i];
      ResidualBlock b; b.kind = 7; b.index = (int)i; b.nres = 3; b.cauchy_a = align_cauchy_a;
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&r3[3 * (m.s + j)], 3));
      b.pb.push_back(mp.AddParameterBlock(&t_s_gp_corr_s, 1));
      b.pb.push_back(mp.AddParameterBlock(&scale_corr, 1));
      b.pb.push_back(mp.AddParameterBlock(R_wg_corr, 4, true));
      b.pb.push_back(mp.AddParameterBlock(p_wg_corr, 3));
      b.pb.push_back(mp.AddParameterBlock(p_bp, 3));
      GlobalPositionSensorWithAlignmentCorrectionCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2]}, {T_wg[0], T_wg[1], T_wg[2], T_wg[3]},
                                                                   {T_wg[4], T_wg[5], T_wg[6]}, scale, m.st_s, m.k, dt_s, inv_dt, m.inv_std};
      b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      mp.blocks.push_back(std::move(b));
      for (int j = 0; j < NN; ++j) { mp.SetParameterBlockConstant(&so3[4 * (m.s + j)]); mp.SetParameterBlockConstant(&r3[3 * (m.s + j)]); }   // :1191-1199
      if (!optimize_p_bp) mp.SetParameterBlockConstant(p_bp);
      if (!optimize_t_offset_s_gs) mp.SetParameterBlockConstant(&t_s_gp_corr_s);
    }
    for (size_t i = 0; i < gp.size(); ++i) {
      const GpMeas& m = gp[i];
      ResidualBlock b; b.kind = 3; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&r3[3 * (m.s + j)], 3));
      b.pb.push_back(mp.AddParameterBlock(p_bp, 3));
      b.pb.push_back(mp.AddParameterBlock(&t_s_gp_corr_s, 1));
      b.pb.push_back(mp.AddParameterBlock(&scale_corr, 1));
      GlobalPositionSensorCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2]}, scale, m.st_s, m.k, dt_s, inv_dt, m.inv_std};
      b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      mp.blocks.push_back(std::move(b));
      if (!optimize_scale) mp.SetParameterBlockConstant(&scale_corr);
      if (!optimize_p_bp) mp.SetParameterBlockConstant(p_bp);
      if (!optimize_t_offset_s_gs) mp.SetParameterBlockConstant(&t_s_gp_corr_s);
    }
    for (size_t i = 0; i < accel.size(); ++i) {
      {
        const ImuMeas& m = accel[i];
        ResidualBlock b; b.kind = 1; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&r3[3 * (m.s + j)], 3));
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&acc_bias[3 * (m.s_bias + j)], 3));  // N_ blocks, 4 read (:585-588)
        b.pb.push_back(mp.AddParameterBlock(g, 3));
        b.pb.push_back(mp.AddParameterBlock(&t_offset_cam_imu_corr_s, 1));
        AccelerometerCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2]}, m.st_s, m.k, dt_s, inv_dt, m.u_bias, inv_dt_bias, m.inv_std};
        b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
        mp.blocks.push_back(std::move(b));
        if (!optimize_gravity) mp.SetParameterBlockConstant(g);
        if (!optimize_camimu_extrinsics) mp.SetParameterBlockConstant(&t_offset_cam_imu_corr_s);
      }
      {
        const ImuMeas& m = gyro[i];
        ResidualBlock b; b.kind = 2; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
        for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&gyr_bias[3 * (m.s_bias + j)], 3));
        b.pb.push_back(mp.AddParameterBlock(&t_offset_cam_imu_corr_s, 1));
        GyroCostFunctor<NN> f{{m.meas[0], m.meas[1], m.meas[2]}, m.st_s, m.k, dt_s, inv_dt, m.u_bias, inv_dt_bias, m.inv_std};
        b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
        mp.blocks.push_back(std::move(b));
        if (!optimize_camimu_extrinsics) mp.SetParameterBlockConstant(&t_offset_cam_imu_corr_s);
      }
    }
    for (size_t i = 0; i < frames.size(); ++i) {
      const FrameMeas& m = frames[i];
      const int n = (int)m.lm.size();
      ResidualBlock b; b.kind = 0; b.index = (int)i; b.nres = 2 * n; b.cauchy_a = cauchy_a;
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&so3[4 * (m.s + j)], 4, true));
      for (int j = 0; j < NN; ++j) b.pb.push_back(mp.AddParameterBlock(&r3[3 * (m.s + j)], 3));
      b.pb.push_back(mp.AddParameterBlock(&scale_corr, 1));
      b.pb.push_back(mp.AddParameterBlock(R_bc_corr, 4, true));
      b.pb.push_back(mp.AddParameterBlock(p_bc_corr, 3));
      for (int j = 0; j < n; ++j) b.pb.push_back(mp.AddParameterBlock(&lm[3 * (size_t)m.lm[j]], 3));
      ReprojectionCostFunctor<NN> f{&cam, scale, m.u, inv_dt, m.inv_std, n, m.uv.data()};
      b.eval = [f, n](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 2 * n, r, j); };
      mp.blocks.push_back(std::move(b));
      if (!optimize_scale) mp.SetParameterBlockConstant(&scale_corr);
      if (!optimize_camimu_extrinsics) { mp.SetParameterBlockConstant(R_bc_corr); mp.SetParameterBlockConstant(p_bc_corr); }
    }
    for (size_t i = 0; i < bias.size(); ++i) {
      const BiasMeas& m = bias[i];
      std::vector<double>& B = m.which == 0 ? acc_bias : gyr_bias;
      ResidualBlock b; b.kind = 4; b.index = (int)i; b.nres = 3; b.cauchy_a = 0;
      for (int j = 0; j < Nb; ++j) b.pb.push_back(mp.AddParameterBlock(&B[3 * (m.s0 + j)], 3));
      for (int j = 0; j < Nb; ++j) b.pb.push_back(mp.AddParameterBlock(&B[3 * (m.s1 + j)], 3));
      BiasCostFunctor f{m.u0, m.u1, inv_dt_bias, m.inv_std};
      b.eval = [f](const std::vector<const double*>& p, const std::vector<int>& sz, double* r, double* j) { dynamic_autodiff(f, p, sz, 3, r, j); };
      mp.blocks.push_back(std::move(b));
    }
    mp.finalize();
    built = true;
  }
  void build() {
    if (built) return;
    if (N == 4) buildT<4>(); else if (N == 6) buildT<6>(); else if (N == 5) buildT<5>();
  }

  // :1807-1811, :1853-1857 — bookkeeping only; the residual blocks keep their captured st_s, k
  void relinearize(int mask = 3) {   // bit 0: relinearizeSplineImuTimeOffset, bit 1: relinearizeSplineGlobSensTimeOffset
    if (mask & 1) { t_offset_cam_imu_s += t_offset_cam_imu_corr_s; t_offset_cam_imu_corr_s = 0.0; }
    if (mask & 2) { t_s_gp_s += t_s_gp_corr_s; t_s_gp_corr_s = 0.0; }
  }
};

}  // namespace oracle
