// Host-side DEBUG harness (tests only): runs the CTBA_HD evaluators of rpg_vision-based_slam_b200/csrc/ctba_dt_math.h (discrete-time
// estimator) on the CPU so that their formulas can be checked against the oracle on a GPU-less machine.  Not part of the product library.
#include <cstring>
#include <vector>
#include "../../rpg_vision-based_slam_b200/csrc/ctba_dt_math.h"
using namespace ctba;

// cfg: [model, fx, fy, cx, cy, d0..d3, q_bc0(4), p_bc0(3), width, height, t_offset_cam_imu, scale, t_offset_gs_imu,
//       gyro_noise_density, acc_noise_density, gyro_random_walk, acc_random_walk, g_max, a_max, g_W(3), cauchy_a]   (31 doubles)
static DtConst make_const(const double* cfg) {
  DtConst K;
  K.cam.model = (int)cfg[0]; K.cam.fx = cfg[1]; K.cam.fy = cfg[2]; K.cam.cx = cfg[3]; K.cam.cy = cfg[4];
  for (int i = 0; i < 4; ++i) { K.cam.d[i] = cfg[5 + i]; K.q_bc0[i] = cfg[9 + i]; }
  for (int i = 0; i < 3; ++i) K.p_bc0[i] = cfg[13 + i];
  K.width = cfg[16]; K.height = cfg[17]; K.t_offset_cam_imu = cfg[18]; K.scale = cfg[19]; K.t_offset_gs_imu = cfg[20];
  K.gyro_noise_density = cfg[21]; K.acc_noise_density = cfg[22]; K.gyro_random_walk = cfg[23]; K.acc_random_walk = cfg[24]; K.g_max = cfg[25]; K.a_max = cfg[26];
  for (int i = 0; i < 3; ++i) K.g_W[i] = cfg[27 + i];
  K.cauchy_a = cfg[30];
  return K;
}

extern "C" {
void hdt_se3_plus(const double* T, const double* d, double* out) { dt_se3_plus(T, d, out); }
void hdt_se3_plus_jacobian(const double* T, double* J) { dt_se3_plus_jacobian(T, J); }
// one frame: n valid reprojections -> r (2n), J (2n x 17)
void hdt_reproj(const double* cfg, const double* pose, const double* glb, int n, const double* pw, const double* uv, const double* vel, double inv_std, double* r,
                double* J) {
  const DtConst K = make_const(cfg);
  DtFrameCam fc;
  dt_reproj_frame(K, pose, glb, &fc);
  for (int i = 0; i < n; ++i) dt_reproj_point<true>(K, fc, v3{pw[3 * i], pw[3 * i + 1], pw[3 * i + 2]}, uv + 2 * i, vel + 2 * i, inv_std, r + 2 * i, J + 34 * (size_t)i);
}
void hdt_gp(const double* cfg, const double* pose0, const double* pose1, const double* glb, double t, const double* z, double tk0, double tk1, double inv_std, double* r,
            double* J) {
  const DtConst K = make_const(cfg);
  dt_gp_eval<true>(K, pose0, pose1, glb, t, z, tk0, tk1, inv_std, r, J);
}
// a persistent factor cache (the caller keeps the 288-double block between calls; zero-initialised block with redo = 1 on the first call)
int hdt_imu_cache_size() { return (int)sizeof(DtImuCache); }
void hdt_imu_cache_init(void* c) { std::memset(c, 0, sizeof(DtImuCache)); ((DtImuCache*)c)->redo = 1; ((DtImuCache*)c)->Delta_q[3] = 1.0; }
int hdt_imu_redo_count(const void* c) { return ((const DtImuCache*)c)->redo_counter; }
int hdt_imu(const double* cfg, void* cache, int n, const double* ts, const double* gyr, const double* acc, double t0, double t1, const double* T0, const double* sb0,
            const double* T1, const double* sb1, double* r, double* J) {
  const DtConst K = make_const(cfg);
  std::vector<double> scratch(3 * 225);
  return dt_imu_eval(K, (DtImuCache*)cache, ts, gyr, acc, n, t0, t1, T0, sb0, T1, sb1, r, J, scratch.data());
}
}
