// Host-side DEBUG harness (tests only): runs the CTBA_HD per-element evaluators of
// rpg_vision-based_slam_b200/csrc/ctba_terms.h on the CPU so the analytic Jacobians can be
// checked against the oracle on a GPU-less machine.  Not part of the product library.
#include <cstring>
#include "../../rpg_vision-based_slam_b200/csrc/ctba_terms.h"
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

extern "C" {
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
