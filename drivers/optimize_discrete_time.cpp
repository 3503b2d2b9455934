// optimize_discrete_time — the reference driver of the same name (BASELINE configs[3]: one SE(3) state + [v, b_g, b_a] per frame) re-hosted on
// the C ABI of include/ctba_dt.h:
//     /root/reference/src/optimize_discrete_time.cpp:195-726
// Same YAML keys (experiments/*/*/full_batch_optimization/discrete_time/*.yaml), same inputs (EuRoC / Kalibr calibration, imu and global
// position text files, COLMAP cameras.bin / images.bin / points3D.bin, the COLMAP camera trajectory, T_wg.txt / scale.txt of the alignment
// step), the same preparation on the host — frames sorted by image name (:320-337), states T_WB = T_WG (s) T_GC T_CB inside [start_t, end_t]
// (:339-404), 3-D points to the world frame, velocities by finite differences, computeValidReprojections, computeFeatureVelocity
// (estimator.h:293-325), the IMU samples containing both edges of every consecutive pair (sensor/imu.h:86-135), the left-closest state of
// every global position (sensor/global_position_sensor.h:80-96) — and the same result files.  Everything the reference does inside
// optimize() (estimator.h:699-713 -> ceres::Solve) runs on the GPU (ctba_dt_solve).
// Differences on the host side: points3D rows are in points3D.bin order (the reference writes its unordered_map's iteration order), and the
// reference saves the INITIAL 3-D points twice (save3DPoints(..., false) after the solve, :716): reproduced, so "points3D.txt" is written in
// addition, not instead.
//   usage: optimize_discrete_time <config.yaml> [cuda device]
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <numeric>
#include <sys/stat.h>
#include "ctba_colmap.hpp"
#include "ctba_dt.h"
#include "ctba_io.hpp"

using namespace ctba;
using namespace ctba::io;

static void require(bool ok, const std::string& msg) {   // GVI_FUSION_ASSERT_STREAM (util/assert.h:65-70)
  if (!ok) { std::cerr << msg << "\n"; std::exit(2); }
}
struct DtHandle {   // owns the estimator; every call checked (the reference aborts where these return an error code)
  ctba_dt_problem* h = nullptr;
  explicit DtHandle(int device) {
    const int rc = ctba_dt_create(device, &h);
    if (rc != CTBA_OK) throw Error(rc, "ctba_dt_create failed: no usable CUDA device (the library is CUDA-only, no CPU fallback)");
  }
  ~DtHandle() { if (h) ctba_dt_destroy(h); }
  int chk(int rc) const { if (rc < 0) throw Error(rc, ctba_dt_last_error(h)); return rc; }
};
static void save1D(const std::string& dir, const std::string& fn, const std::vector<double>& v) {   // trace1DVector, precision 9 (:127-136)
  std::ofstream f; createTrace(dir, fn, &f); f.precision(9);
  for (double e : v) f << e << std::endl;
}
static void savePoints(const std::string& dir, const std::string& fn, const std::vector<double>& xyz) {   // tracePoints3D, precision 9 (:139-156)
  std::ofstream f; createTrace(dir, fn, &f); f.precision(9);
  f << "# x y z" << std::endl;
  for (size_t l = 0; l + 2 < xyz.size(); l += 3) f << xyz[l] << " " << xyz[l + 1] << " " << xyz[l + 2] << std::endl;
}
// saveTrajectoryVelocityBiases (:44-107): traceTimeAndPose / traceTimeAndPosition (util/save.h:220-242), precision 15; velocity = sb[0..2],
// accelerometer bias = the TAIL (sb[6..8]), gyroscope bias = the middle segment (sb[3..5])
static void saveStates(const std::string& dir, DtHandle& est, const std::vector<double>& ts_imu, bool optimized) {
  const size_t F = ts_imu.size();
  std::vector<double> poses(7 * F), sb(9 * F);
  est.chk(ctba_dt_get_states(est.h, poses.data(), sb.data()));
  const std::string pre = optimized ? "" : "initial_";
  std::ofstream traj, vel, ab, gb;
  createTrace(dir, pre + "trajectory.txt", &traj); createTrace(dir, pre + "velocity.txt", &vel);
  createTrace(dir, pre + "acc_bias.txt", &ab); createTrace(dir, pre + "gyro_bias.txt", &gb);
  traj.precision(15); vel.precision(15); ab.precision(15); gb.precision(15);
  traj << "# ts x y z qx qy qz qw" << std::endl;
  vel << "# ts x y z" << std::endl; ab << "# ts x y z" << std::endl; gb << "# ts x y z" << std::endl;
  for (size_t i = 0; i < F; ++i) {
    const double* T = &poses[7 * i]; const double* x = &sb[9 * i];
    traj << ts_imu[i] << " " << T[4] << " " << T[5] << " " << T[6] << " " << T[0] << " " << T[1] << " " << T[2] << " " << T[3] << std::endl;
    vel << ts_imu[i] << " " << x[0] << " " << x[1] << " " << x[2] << std::endl;
    ab << ts_imu[i] << " " << x[6] << " " << x[7] << " " << x[8] << std::endl;
    gb << ts_imu[i] << " " << x[3] << " " << x[4] << " " << x[5] << std::endl;
  }
}
static double meanAbove(const std::vector<double>& v, double thr, int* n) { double s = 0; *n = 0; for (double e : v) if (e > thr) { s += e; ++*n; } return *n ? s / *n : 0.0; }

static int run(const Yaml& config, int device) {
  const int cam_id = config.as<int>("cam_id");
  const std::string imu_fn = config.as<std::string>("imu_fn"), gp_fn = config.as<std::string>("gp_fn");
  const std::string colmap_dir = config.as<std::string>("colmap_dir"), colmap_fn = config.as<std::string>("colmap_fn");
  const std::string alignment_dir = config.as<std::string>("alignment_dir");
  std::string trace_dir = config.as<std::string>("full_batch_optimization_dir"); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/discrete_time"; mkdir(trace_dir.c_str(), 0777);
  const std::string dataset_fn = config.as<std::string>("dataset");
  std::cout << "Dataset: " << dataset_fn << "\nresult dir: " << trace_dir << "\n\n";
  const double start_ts = config.as<double>("start_t"), end_ts = config.as<double>("end_t");

  SimpleCamImuCalib calib;
  require(loadCalibForDataset(config, dataset_fn, cam_id, calib), "Camera-Imu calibration could not be loaded");
  const double acc_noise_density = config.as<double>("sigma_acc_c"), acc_random_walk = config.as<double>("sigma_acc_bias_c");
  const double gyro_noise_density = config.as<double>("sigma_omega_c"), gyro_random_walk = config.as<double>("sigma_omega_bias_c");
  std::vector<ImuMeasurement> imu;
  require(loadImuMeasurementsFromFile(imu_fn, imu), "Could not load imu");
  std::cout << "Loaded " << imu.size() << " imu measurements.\n\n";
  std::vector<GlobalPosMeasurement> gp_meas;
  require(loadGpMeasurements(gp_fn, gp_meas), "Could not load global measurements.");
  std::cout << "Loaded " << gp_meas.size() << " global measurements.\n\n";

  std::vector<ColmapCamera> cameras;
  readCamerasBinary(colmap_dir + "cameras.bin", cameras);
  require(!cameras.empty(), "no cameras");
  std::cout << "Loaded: " << cameras.size() << " cameras.\n\n";
  const Pose T_cb = calib.T_cam_imu_, T_bc = poseInverse(T_cb);                      // :299-302
  const double t_offset_cam_imu = calib.timeshift_cam_imu_;
  ColmapScene sc;
  const int grid_cell_size = config.as<int>("grid_cell_size");
  readImagesAndPoints3DBinary(colmap_dir + "images.bin", colmap_dir + "points3D.bin", cameras, sc);
  std::cout << "Full trajectory contains: " << sc.frames.size() << " images.\n\nFull 3D map contains: " << sc.lm_id.size() << " 3D points.\n\n";
  for (int32_t l : sc.lm_index) require(l >= 0, "images.bin refers to a 3D point that points3D.bin does not hold");

  // frames in the order of their image names (argSort = std::stable_sort on the strings, util/utils.h:16-31)
  std::vector<size_t> order(sc.frames.size());
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](size_t a, size_t b) { return sc.frames[a].name < sc.frames[b].name; });

  std::vector<Quat> q_gc_full; std::vector<Vec3> p_gc_full; std::vector<double> cam_ts_full;
  require(loadTrajectoryFromFile(colmap_fn, q_gc_full, p_gc_full, cam_ts_full), "Could not load COLMAP camera poses.");
  require(sc.frames.size() == cam_ts_full.size(),
          "frames_full_traj.size()= " + std::to_string(sc.frames.size()) + " cam_ts_full_traj.size()= " + std::to_string(cam_ts_full.size()));
  Pose T_wg; double scale = 0;
  require(loadPose(alignment_dir + "/T_wg.txt", T_wg), "Could not T_WG");
  require(loadScalar<double>(alignment_dir + "/scale.txt", scale), "Could not scale");
  std::cout << "Loaded scale: " << scale << "\n\n";

  // states: T_WB = (R_WG R_GB, s R_WG p_GB + p_WG), T_GB = T_GC T_CB, for the frames inside [start_t, end_t]        (:339-404)
  std::vector<size_t> sel;                 // index into the sorted full trajectory
  for (size_t i = 0; i < cam_ts_full.size(); i++) if (cam_ts_full[i] >= start_ts && cam_ts_full[i] <= end_ts) sel.push_back(i);
  const int F = (int)sel.size();
  require(F >= 2, "fewer than two frames between start_t and end_t");
  std::vector<double> poses(7 * (size_t)F), cam_ts(F), ts_imu(F);
  for (int f = 0; f < F; ++f) {
    const size_t i = sel[f];
    const Pose T_gb = poseMul(Pose{p_gc_full[i], q_gc_full[i]}, T_cb);
    const Quat q_wb = quatMul(T_wg.q, T_gb.q);
    const Vec3 r = rotate(T_wg.q, T_gb.p);
    for (int c = 0; c < 4; ++c) poses[7 * f + c] = q_wb[c];
    for (int c = 0; c < 3; ++c) poses[7 * f + 4 + c] = scale * r[c] + T_wg.p[c];
    cam_ts[f] = cam_ts_full[i];
    ts_imu[f] = cam_ts[f] + t_offset_cam_imu;                                      // State::ts_imu_ (estimator.h:96-99)
  }
  std::vector<double> points_w(sc.lm_xyz.size());
  for (size_t l = 0; l + 2 < points_w.size(); l += 3) {
    const Vec3 r = rotate(T_wg.q, Vec3{{sc.lm_xyz[l], sc.lm_xyz[l + 1], sc.lm_xyz[l + 2]}});
    for (int c = 0; c < 3; ++c) points_w[l + c] = scale * r[c] + T_wg.p[c];
  }
  // the selected frames' 2-D points as one CSR (frame f = the sel[f]-th frame in name order)
  std::vector<int32_t> pt_begin(F + 1, 0), lm_index; std::vector<double> uv;
  for (int f = 0; f < F; ++f) {
    const size_t src = order[sel[f]];
    for (int q = sc.pt_begin[src]; q < sc.pt_begin[src + 1]; ++q) { uv.push_back(sc.uv[2 * q]); uv.push_back(sc.uv[2 * q + 1]); lm_index.push_back(sc.lm_index[q]); }
    pt_begin[f + 1] = (int32_t)lm_index.size();
  }

  DtHandle est(device);
  const ColmapCamera& c0 = cameras[0];
  const double intr[4] = {c0.fx, c0.fy, c0.cx, c0.cy}, dist[4] = {c0.dist_coeffs[0], c0.dist_coeffs[1], c0.dist_coeffs[2], c0.dist_coeffs[3]};
  const double g_W[3] = {0.0, 0.0, 9.80665};                                         // ImuCalibration::g_W_ (sensor/imu_calib.h:32-35)
  est.chk(ctba_dt_set_imu(est.h, gyro_noise_density, acc_noise_density, gyro_random_walk, acc_random_walk, 7.8, 150.0, g_W));
  est.chk(ctba_dt_set_camera(est.h, c0.model_id, intr, dist, T_bc.q.data(), T_bc.p.data(), t_offset_cam_imu, (double)c0.w, (double)c0.h));
  est.chk(ctba_dt_initialize_states(est.h, F, poses.data(), cam_ts.data()));
  est.chk(ctba_dt_initialize_linear_velocity(est.h));
  const double t_cam_gp_init = config.as<double>("t_offset_cam_gp_init");
  std::printf("Loaded camera - global sensor time offset: %.15g\n\n", t_cam_gp_init);
  const double t_gp_imu_init = t_offset_cam_imu - t_cam_gp_init;                    // :447-449
  const std::vector<double> p_BP = config.asList("p_BP");
  require(p_BP.size() == 3, "p_BP must hold three values");
  est.chk(ctba_dt_set_globals(est.h, scale, p_BP.data(), t_gp_imu_init));
  // optimize_camimu_extrinsics stays at the constructor default (true, estimator.h:64-65): the driver never sets it
  est.chk(ctba_dt_set_flags(est.h, config.as<bool>("optimize_scale"), config.as<bool>("optimize_pBP"), 1, config.as<bool>("optimize_t_offset_globsens_imu")));
  est.chk(ctba_dt_set_landmarks(est.h, (int)(points_w.size() / 3), points_w.data()));

  // computeValidReprojections (:467-481), then computeFeatureVelocity of the valid points against the NEXT frame's points (any of them)
  std::vector<uint8_t> valid(std::max<size_t>(lm_index.size(), 1)); std::vector<int32_t> n_valid(F);
  est.chk(ctba_dt_compute_valid_reprojections(est.h, F, pt_begin.data(), uv.data(), lm_index.data(), (int)c0.w, (int)c0.h, grid_cell_size, valid.data(),
                                              n_valid.data()));
  std::vector<double> vel(uv.size(), 0.0);
  for (int f = 0; f + 1 < F; ++f) {
    const double dt = cam_ts[f + 1] - cam_ts[f];
    std::unordered_map<int32_t, int32_t> next;                                      // landmark -> first point of the next frame that shows it
    for (int q = pt_begin[f + 2] - 1; q >= pt_begin[f + 1]; --q) next[lm_index[q]] = q;
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) {
      if (!valid[q]) continue;
      auto it = next.find(lm_index[q]);
      if (it == next.end()) continue;
      vel[2 * q] = (uv[2 * it->second] - uv[2 * q]) / dt; vel[2 * q + 1] = (uv[2 * it->second + 1] - uv[2 * q + 1]) / dt;
    }
  }

  // IMU factors: the samples containing both edges of every consecutive pair (getMeasurementsContainingEdges, remove_measurements = false)
  int n_imu_factors = 0;
  for (int f = 1; f < F; ++f) {
    const double t0 = ts_imu[f - 1], t1 = ts_imu[f];
    if (imu.empty() || imu.front().t > t0 || imu.back().t < t1) continue;
    size_t k0 = 0;
    for (; k0 < imu.size(); ++k0) if (imu[k0].t > t0) { --k0; break; }             // the first sample is not newer than t0 (checked above)
    size_t k1 = k0;
    for (; k1 < imu.size(); ++k1) if (imu[k1].t > t1) break;
    if (k1 >= imu.size()) k1 = imu.size() - 1;                                      // it_k_p_1 == end(): the reference reads one past the end; clamped
    std::vector<double> t, gyr, acc;
    for (size_t k = k0; k <= k1; ++k) {
      t.push_back(imu[k].t);
      for (int c = 0; c < 3; ++c) { gyr.push_back(imu[k].ang_vel[c]); acc.push_back(imu[k].lin_acc[c]); }
    }
    est.chk(ctba_dt_add_imu(est.h, f - 1, (int)t.size(), t.data(), gyr.data(), acc.data()));
    ++n_imu_factors;
  }
  // vision: the valid reprojections of every frame
  const double sigma_reproj_px = config.as<double>("sigma_reproj_px");
  std::cout << "Sigma reprojection errors: " << sigma_reproj_px << " [pixel]\n\n";
  {
    std::vector<int32_t> state_id(F), ob(F + 1, 0), lm; std::vector<double> uvv, velv;
    for (int f = 0; f < F; ++f) {
      state_id[f] = f;
      for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q)
        if (valid[q]) { uvv.push_back(uv[2 * q]); uvv.push_back(uv[2 * q + 1]); velv.push_back(vel[2 * q]); velv.push_back(vel[2 * q + 1]); lm.push_back(lm_index[q]); }
      ob[f + 1] = (int32_t)lm.size();
    }
    const int rc = est.chk(ctba_dt_add_frames(est.h, F, state_id.data(), ob.data(), uvv.data(), velv.data(), lm.data(), sigma_reproj_px));
    std::cout << "Frames in the problem: " << rc << " of " << F << ", valid reprojections: " << lm.size() << ", IMU factors: " << n_imu_factors << "\n\n";
  }
  // global positions: valid when a state lies to their left (getLeftClosestStateId(state_ts, t + t_gp_imu_init) >= 0, :493-502)
  const double gp_meas_std = gpStdForDataset(config, dataset_fn);
  std::cout << "Sigma global measurements: " << gp_meas_std << " [m]\n\n";
  {
    std::vector<double> t, z; std::vector<int32_t> f0;
    for (const GlobalPosMeasurement& m : gp_meas) {
      const double tm = m.t + t_gp_imu_init;
      int sid = -1;
      for (int i = 0; i < F; ++i) if (ts_imu[i] > tm) { sid = i - 1; break; }
      if (sid < 0) continue;
      t.push_back(m.t); f0.push_back(sid);
      for (int c = 0; c < 3; ++c) z.push_back(m.pos[c]);
    }
    est.chk(ctba_dt_add_gp(est.h, (int)t.size(), t.data(), z.data(), f0.data(), gp_meas_std));
    std::cout << "Global positions in the problem: " << t.size() << " of " << gp_meas.size() << "\n\n";
  }
  int64_t counts[5];
  est.chk(ctba_dt_counts(est.h, counts));

  auto errors = [&](int kind, size_t n) { std::vector<double> e(std::max<size_t>(n, 1)); est.chk(ctba_dt_evaluate_errors(est.h, kind, e.data())); e.resize(n); return e; };
  auto globals = [&]() { std::vector<double> g(13); est.chk(ctba_dt_get_globals(est.h, g.data())); return g; };
  auto T_BC_now = [&](const std::vector<double>& g) { return poseMul(T_bc, Pose{{{g[5], g[6], g[7]}}, Quat{{g[1], g[2], g[3], g[4]}}}); };   // getImuCamRelTransformation (:754-758)
  auto saveGlobals = [&](bool optimized) {
    const std::vector<double> g = globals();
    const std::string pre = optimized ? "" : "initial_";
    saveScalar(trace_dir, pre + "scale.txt", scale + g[0]);
    savePosition(trace_dir, pre + "p_BP.txt", Vec3{{g[10], g[11], g[12]}});
    savePose(trace_dir, pre + "T_BC.txt", T_BC_now(g));
    saveScalar(trace_dir, pre + "cam_imu_time_offset.txt", t_offset_cam_imu + g[8]);
    savePosition(trace_dir, pre + "gravity.txt", Vec3{{g_W[0], g_W[1], g_W[2]}});
    saveScalar(trace_dir, pre + "globalsensor_imu_time_offset.txt", t_gp_imu_init + g[9]);
  };

  std::cout << "+-+-+-+-+- Initial Errors +-+-+-+-+-\n\n";
  int n_err = 0;
  std::vector<double> init_gp_errors = errors(CTBA_DT_GP, (size_t)counts[3]), init_reproj_errors = errors(CTBA_DT_REPROJ, (size_t)counts[1]);
  std::cout << "Number of global position errors: " << init_gp_errors.size() << "\nInitial global position error mean: " << meanAbove(init_gp_errors, -1.0, &n_err) << "\n\n";
  const double m0 = meanAbove(init_reproj_errors, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nInitial reprojection error mean: " << m0 << "\n\n";
  std::cout << "Saving initial values ...\n\n";
  saveStates(trace_dir, est, ts_imu, false);
  save1D(trace_dir, "initial_global_position_errors.txt", init_gp_errors); save1D(trace_dir, "initial_reprojection_errors.txt", init_reproj_errors);
  saveGlobals(false);
  savePoints(trace_dir, "initial_points3D.txt", points_w);

  const int n_max_iters = config.as<int>("n_max_iters");
  std::cout << "\nLoaded optimization parameters: \nmax number of iterations = " << n_max_iters << "\nnumber of threads = " << config.as<int>("n_threads")
            << " (the solve runs on the GPU)\n\n\n\nOptimizing ...\n\n";
  std::vector<ctba_iteration> its((size_t)n_max_iters + 2);
  ctba_summary summary{};
  summary.iterations = its.data(); summary.iterations_capacity = (int)its.size();
  est.chk(ctba_dt_solve(est.h, n_max_iters, &summary));
  std::printf("solve: iterations %d (successful %d, unsuccessful %d), cost %.12g -> %.12g, termination %d, %.3f s, %lld kernel launches\n", summary.num_iterations,
              summary.num_successful_steps, summary.num_unsuccessful_steps, summary.initial_cost, summary.final_cost, summary.termination_type, summary.time_total_s,
              (long long)summary.kernel_launches);

  std::cout << "+-+-+-+-+- Final Errors +-+-+-+-+-\n\n";
  std::vector<double> opti_gp_errors = errors(CTBA_DT_GP, (size_t)counts[3]), opti_reproj_errors = errors(CTBA_DT_REPROJ, (size_t)counts[1]);
  std::cout << "Number of global position errors: " << opti_gp_errors.size() << "\nOptimized global position error mean: " << meanAbove(opti_gp_errors, -1.0, &n_err) << "\n\n";
  const double m1 = meanAbove(opti_reproj_errors, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nOptimized reprojection error mean: " << m1 << "\n\n";
  {
    const std::vector<double> g = globals();
    const Pose T = T_BC_now(g);
    double R[9]; matrixFromQuat(Quat{{g[1], g[2], g[3], g[4]}}, R);
    const double rot_angle_deg = std::acos((R[0] + R[4] + R[8] - 1.0 - 1e-6) / 2.0) * 180.0 / 3.14159;   // :676-677
    std::cout << "Optimized imu - camera transformation (= T_BC) : q = [" << T.q[0] << " " << T.q[1] << " " << T.q[2] << " " << T.q[3] << "], p = [" << T.p[0] << " "
              << T.p[1] << " " << T.p[2] << "]\nrotation angle correction was: " << rot_angle_deg << " [deg]\ntranslation norm correction was: "
              << std::sqrt(g[5] * g[5] + g[6] * g[6] + g[7] * g[7]) << " [m]\n\n";
    std::cout << "Optimized cam - imu time offset: " << t_offset_cam_imu + g[8] << " [s]\nTime offset correction was: " << g[8] << " [s]\n\n";
    std::cout << "Optimized global sensor - imu time offset: " << t_gp_imu_init + g[9] << " [s]\nTime offset correction was: " << g[9] << " [s]\n\n";
    std::cout << "Optimized p_bp:\n" << g[10] << "\n" << g[11] << "\n" << g[12] << "\nnorm correction was: "
              << std::sqrt((g[10] - p_BP[0]) * (g[10] - p_BP[0]) + (g[11] - p_BP[1]) * (g[11] - p_BP[1]) + (g[12] - p_BP[2]) * (g[12] - p_BP[2])) << " [m]\n\n";
    std::cout << "Optimized colmap - global reference scale: " << scale + g[0] << "\nScale correction was: " << g[0] << "\n\n";
  }
  std::cout << "Saving optimized values ...\n\n";
  saveStates(trace_dir, est, ts_imu, true);
  save1D(trace_dir, "global_position_errors.txt", opti_gp_errors); save1D(trace_dir, "reprojection_errors.txt", opti_reproj_errors);
  saveGlobals(true);
  savePoints(trace_dir, "initial_points3D.txt", points_w);   // as the reference does (:716); the optimised points follow
  {
    std::vector<double> lm(points_w.size());
    est.chk(ctba_dt_get_landmarks(est.h, lm.data()));
    savePoints(trace_dir, "points3D.txt", lm);
  }
  return 0;
}

int main(int argc, char** argv) {
  require(argc > 1, "Please, pass the .yaml filename");
  try {
    const Yaml config = Yaml::LoadFile(argv[1]);
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    return run(config, device);
  } catch (const ctba::Error& e) {
    std::cerr << "ctba error " << e.code << ": " << e.what() << "\n";
    return 3;
  } catch (const std::exception& e) {
    std::cerr << e.what() << "\n";
    return 2;
  }
  return 0;
}
