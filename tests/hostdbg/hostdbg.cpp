// Host-side DEBUG harness (tests only): runs the CTBA_HD per-element evaluators of
// rpg_vision-based_slam_b200/csrc/ctba_terms.h on the CPU so the analytic Jacobians can be
// checked against the oracle on a GPU-less machine.  Not part of the product library.
#include <cstring>
#include "../../rpg_vision-based_slam_b200/csrc/ctba_terms.h"
#include "../../rpg_vision-based_slam_b200/csrc/ctba_align.h"
using namespace ctba;

static Globals make_globals(int N, const double* cfg) {
  // cfg: [model, fx, fy, cx, cy, d0..d3, q_bc0(4), p_bc0(3), dt_s, dt_bias_s, scale]
  Globals G;
  fill_tables(N, &G.tb); fill_tables(4, &G.tb4);
  G.cam.model = (int)cfg[0]; G.cam.fx = cfg[1]; G.cam.fy = cfg[2]; G.cam.cx = cfg[3]; G.cam.cy = cfg[4];
  for (int i = 0; i < 4; ++i) { G.cam.d[i] = cfg[5 + i]; G.q_bc0[i] = cfg[9 + i]; }
  for (int i = 0; i < 3; ++i) G.p_bc0[i] = cfg[13 + i];
  G.dt_s = cfg[16]; G.inv_dt = 1.0 / cfg[16]; G.inv_dt_bias = 1.0 / cfg[17]; G.scale = cfg[18]; G.gate_sq = 100.0;
  return G;
}

template <int N>
static void frame_reproj(const Globals& G, const double* glb, const double* kq, const double* kp, double u, int n, const double* pw,
                         const double* uv, double inv_std, double* r, double* J) {
  FramePose<N> fp;
  frame_pose_eval<N, true>(G, glb, kq, kp, u, &fp);
  for (int i = 0; i < n; ++i) {
    double Jl[6];
    v3 p{pw[3 * i], pw[3 * i + 1], pw[3 * i + 2]};
    reproj_obs_eval<true>(G, fp.R_cw, fp.t_cw, p, uv[2 * i], uv[2 * i + 1], inv_std, r + 2 * i, Jl);
    reproj_expand<N>(G, glb, fp, p, Jl, J + (size_t)i * 2 * (6 * N + 10));
  }
}

// Serial host backend of align_lm_loop (ctba_align.h): the same per-sample accumulation and LM step functions as the kernels of
// ctba_align.cu, summed in measurement order.
template <int N>
struct HostAlignBackend {
  Globals G; const double* T_wg; const double *kq, *kp; int n; const int32_t* s; const double *st_s, *k, *meas, *inv_std;
  double cauchy_b; unsigned mask;
  double x[AL_STATE], cand[AL_STATE], sys[2][AS_SIZE], jscale[AT_SIZE]; int cur = 0;
  int evaluate(bool candidate, double* cost) {
    double* S = sys[candidate ? 1 - cur : cur];
    for (int i = 0; i < AS_SIZE; ++i) S[i] = 0.0;
    for (int i = 0; i < n; ++i)
      align_sample_accumulate<N>(G, T_wg, candidate ? cand : x, kq + 4 * (size_t)s[i], kp + 3 * (size_t)s[i], st_s[i], k[i], meas + 3 * (size_t)i,
                                 inv_std[i], cauchy_b, true, S);
    *cost = S[AS_COST];
    return 0;
  }
  int step(double radius, bool init_jscale, double* sc) { align_lm_step(sys[cur], x, mask, radius, init_jscale, jscale, cand, sc); return 0; }
  void accept() { std::memcpy(x, cand, sizeof(x)); cur = 1 - cur; }
};
template <int N>
static int align_lm_host(const Globals& G, const double* T_wg, double* x12, unsigned mask, const double* kq, const double* kp, int n, const int32_t* s,
                         const double* st_s, const double* k, const double* meas, const double* inv_std, double cauchy_a, int max_iters,
                         ctba_summary* out, double* sys_out) {
  HostAlignBackend<N> be;
  be.G = G; be.T_wg = T_wg; be.kq = kq; be.kp = kp; be.n = n; be.s = s; be.st_s = st_s; be.k = k; be.meas = meas; be.inv_std = inv_std;
  be.cauchy_b = cauchy_a * cauchy_a; be.mask = mask;
  std::memcpy(be.x, x12, sizeof(be.x));
  int rc = align_lm_loop(be, max_iters, false, out);
  std::memcpy(x12, be.x, sizeof(be.x));
  if (sys_out) std::memcpy(sys_out, be.sys[be.cur], sizeof(double) * AS_SIZE);
  return rc;
}

extern "C" {
int dbg_align_lm(int N, const double* cfg, const double* T_wg, double* x12, unsigned mask, const double* kq, const double* kp, int n, const int32_t* s,
                 const double* st_s, const double* k, const double* meas, const double* inv_std, double cauchy_a, int max_iters,
                 ctba_summary* out, double* sys_out) {
  Globals G = make_globals(N, cfg);
  if (N == 4) return align_lm_host<4>(G, T_wg, x12, mask, kq, kp, n, s, st_s, k, meas, inv_std, cauchy_a, max_iters, out, sys_out);
  if (N == 6) return align_lm_host<6>(G, T_wg, x12, mask, kq, kp, n, s, st_s, k, meas, inv_std, cauchy_a, max_iters, out, sys_out);
  return -1;
}
int dbg_frame_reproj(int N, const double* cfg, const double* glb, const double* kq, const double* kp, double u, int n, const double* pw,
                     const double* uv, double inv_std, double* r, double* J) {
  Globals G = make_globals(N, cfg);
  if (N == 4) frame_reproj<4>(G, glb, kq, kp, u, n, pw, uv, inv_std, r, J);
  else if (N == 6) frame_reproj<6>(G, glb, kq, kp, u, n, pw, uv, inv_std, r, J);
  else return -1;
  return 0;
}
int dbg_imu(int N, const double* cfg, const double* glb, const double* kq, const double* kp, const double* ba, const double* bg, double st_s,
            double k, double u_bias, const double* acc, const double* gyr, double isa, double isg, double* ra, double* rg, double* Ja, double* Jg) {
  Globals G = make_globals(N, cfg);
  if (N == 4) imu_sample_eval<4, true>(G, glb, kq, kp, ba, bg, st_s, k, u_bias, acc, gyr, isa, isg, ra, rg, Ja, Jg);
  else if (N == 6) imu_sample_eval<6, true>(G, glb, kq, kp, ba, bg, st_s, k, u_bias, acc, gyr, isa, isg, ra, rg, Ja, Jg);
  else return -1;
  return 0;
}
int dbg_gp(int N, const double* cfg, const double* glb, const double* kq, const double* kp, double st_s, double k, const double* meas, double is,
           double* r, double* J) {
  Globals G = make_globals(N, cfg);
  if (N == 4) gp_sample_eval<4, true>(G, glb, kq, kp, st_s, k, meas, is, r, J);
  else if (N == 6) gp_sample_eval<6, true>(G, glb, kq, kp, st_s, k, meas, is, r, J);
  else return -1;
  return 0;
}
int dbg_prior(int N, const double* cfg, const double* kq, const double* kp, int which, double u, const double* meas4, double is, double* r, double* J) {
  Globals G = make_globals(N, cfg);
  if (N == 4) prior_eval<4, true>(G, kq, kp, which, u, meas4, is, r, J);
  else if (N == 6) prior_eval<6, true>(G, kq, kp, which, u, meas4, is, r, J);
  else return -1;
  return 0;
}
int dbg_align(int N, const double* cfg, const double* T_wg, const double* x12, const double* kq, const double* kp, double st_s, double k,
              const double* meas, double* e, double* J) {
  Globals G = make_globals(N, cfg);
  if (N == 4) align_eval<4, true>(G, T_wg, x12, kq, kp, st_s, k, meas, e, J);
  else if (N == 6) align_eval<6, true>(G, T_wg, x12, kq, kp, st_s, k, meas, e, J);
  else return -1;
  return 0;
}
}
