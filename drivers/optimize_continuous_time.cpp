// optimize_continuous_time — the reference driver of the same name (visual + inertial + global position, BASELINE config 1) re-hosted on
// the C ABI:     /root/reference/src/optimize_continuous_time.cpp:233-780
// Same YAML keys (experiments/*/*/full_batch_optimization/continuous_time/*.yaml), same inputs (EuRoC-format calibration YAML, imu.txt, global
// position measurements, the COLMAP export, the fitted spline_knots.txt, and T_wg.txt / scale.txt / spline_globalsensor_time_offset.txt /
// p_bp.txt written by align_spline_to_global_frame), the same call order, outer loop (both time offsets re-linearised between solves) and
// result files.  optimize() runs on the GPU (ctba_solve).  Host-side difference: points3D.txt rows are in
// points3D.bin order.
//   usage: optimize_continuous_time <config.yaml> [cuda device] [order 4|6]
#include <cstdio>
#include <cstdlib>
#include <sys/stat.h>
#include "ctba_colmap.hpp"
#include "ctba_io.hpp"

using namespace ctba;
using namespace ctba::io;

static void require(bool ok, const std::string& msg) {   // GVI_FUSION_ASSERT_STREAM (util/assert.h:65-70)
  if (!ok) { std::cerr << msg << "\n"; std::exit(2); }
}
static void save1D(const std::string& dir, const std::string& fn, const std::vector<double>& v) {   // trace1DVector, precision 9 (:88-115)
  std::ofstream f; createTrace(dir, fn, &f); f.precision(9);
  for (double e : v) f << e << std::endl;
}
static void savePoints(const std::string& dir, const std::string& fn, const std::vector<Vec3>& pts) {   // tracePoints3D / tracePoints3DArray
  std::ofstream f; createTrace(dir, fn, &f); f.precision(9);
  f << "# x y z" << std::endl;
  for (const Vec3& p : pts) f << p[0] << " " << p[1] << " " << p[2] << std::endl;
}
template <class Spline>
static void saveSplineAt(const std::string& dir, Spline& spline, double rate_hz, bool optimized) {   // :41-61
  std::ofstream f; createTrace(dir, optimized ? "spline.txt" : "initial_spline.txt", &f); f.precision(15);
  traceSpline(spline, static_cast<int64_t>((1.0 / rate_hz) * 1e9), f);
}
static double meanAbove(const std::vector<double>& v, double thr, int* n) { double s = 0; *n = 0; for (double e : v) if (e > thr) { s += e; ++*n; } return *n ? s / *n : 0.0; }

template <int order>
static int run(const Yaml& config, int device) {
  const int cam_id = config.as<int>("cam_id");
  const std::string imu_fn = config.as<std::string>("imu_fn"), gp_fn = config.as<std::string>("gp_fn"), colmap_dir = config.as<std::string>("colmap_dir");
  const double spline_knots_dt_s = config.as<double>("spline_control_nodes_dt_s"), sampling_spline_rate_hz = config.as<double>("sampling_spline_rate");
  const int spline_knots_dt_ms = spline_knots_dt_s * 1000;
  const std::string sub = "/order_" + std::to_string(order) + "/dt_" + std::to_string(spline_knots_dt_ms) + "_ms";
  const std::string init_spline_fn = config.as<std::string>("colmap_spline_dir") + sub + "/spline_knots.txt";
  const std::string alignment_dir = config.as<std::string>("spline_global_ref_alignment_dir") + sub;
  std::string trace_dir = config.as<std::string>("full_batch_optimization_dir"); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/continuous_time"; mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/order_" + std::to_string(order); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/dt_" + std::to_string(spline_knots_dt_ms) + "_ms"; mkdir(trace_dir.c_str(), 0777);
  const std::string dataset_fn = config.as<std::string>("dataset");
  std::cout << "Dataset: " << dataset_fn << "\nresult dir: " << trace_dir << "\n\n";

  SimpleCamImuCalib calib;
  require(loadCalibForDataset(config, dataset_fn, cam_id, calib), "Camera-Imu calibration could not be loaded");   // UZH_FPV: Kalibr camchain + imu.yaml; EuRoC / Custom: EuRoC format
  ImuNoise noise;                                                                   // imu noise from the config file (:307-311)
  noise.acc_noise_density = config.as<double>("sigma_acc_c"); noise.acc_random_walk = config.as<double>("sigma_acc_bias_c");
  noise.gyro_noise_density = config.as<double>("sigma_omega_c"); noise.gyro_random_walk = config.as<double>("sigma_omega_bias_c");
  std::vector<ImuMeasurement> imu_meas;
  require(loadImuMeasurementsFromFile(imu_fn, imu_meas), "Could not imu");
  std::cout << "Loaded " << imu_meas.size() << " imu measurements.\n\n";
  std::vector<GlobalPosMeasurement> gp_meas;
  require(loadGpMeasurements(gp_fn, gp_meas), "Could not load global measurements.");
  std::cout << "Loaded " << gp_meas.size() << " global measurements.\n\n";

  std::vector<ColmapCamera> cameras;
  readCamerasBinary(colmap_dir + "cameras.bin", cameras);
  require(!cameras.empty(), "no cameras");
  std::cout << "Loaded: " << cameras.size() << " cameras.\n\n";
  const Pose T_bc = poseInverse(calib.T_cam_imu_);
  ColmapScene sc;
  const int grid_cell_size = config.as<int>("grid_cell_size");
  readImagesAndPoints3DBinary(colmap_dir + "images.bin", colmap_dir + "points3D.bin", cameras, sc);
  std::cout << "Loaded: " << sc.frames.size() << " images.\n\nLoaded: " << sc.lm_id.size() << " 3D points.\n\n";
  for (int32_t l : sc.lm_index) require(l >= 0, "images.bin refers to a 3D point that points3D.bin does not hold");

  std::vector<Quat> knots_q_gb; std::vector<Vec3> knots_r_gb; std::vector<double> knots_ts;
  require(loadSplineFromFile(init_spline_fn, knots_q_gb, knots_r_gb, knots_ts) && !knots_ts.empty(), "Could not load spline.");
  std::cout << "Loaded " << knots_q_gb.size() << " initial spline knots.\n\n";
  Pose T_wg; double scale = 0;                                                      // the results of align_spline_to_global_frame (:414-434)
  require(loadPose(alignment_dir + "/T_wg.txt", T_wg), "Could not T_wg");
  require(loadScalar<double>(alignment_dir + "/scale.txt", scale), "Could not scale");
  std::cout << "Loaded scale: " << scale << "\n\n";
  std::vector<Quat> knots_q_wb(knots_q_gb.size()); std::vector<Vec3> knots_r_wb(knots_r_gb.size());
  for (size_t i = 0; i < knots_q_gb.size(); i++) {
    knots_q_wb[i] = quatMul(T_wg.q, knots_q_gb[i]);
    const Vec3 r = rotate(T_wg.q, knots_r_gb[i]);
    knots_r_wb[i] = Vec3{{scale * r[0] + T_wg.p[0], scale * r[1] + T_wg.p[1], scale * r[2] + T_wg.p[2]}};
  }
  std::vector<Vec3> points_w(sc.lm_id.size());
  for (size_t l = 0; l < points_w.size(); ++l) {
    const Vec3 r = rotate(T_wg.q, Vec3{{sc.lm_xyz[3 * l], sc.lm_xyz[3 * l + 1], sc.lm_xyz[3 * l + 2]}});
    points_w[l] = Vec3{{scale * r[0] + T_wg.p[0], scale * r[1] + T_wg.p[1], scale * r[2] + T_wg.p[2]}};
  }
  std::vector<double> frames_ts;
  require(loadFrameTimes(colmap_dir + "colmap_cam_estimates.txt", frames_ts), "Could not load colmap frame timestamps.");
  require(sc.frames.size() == frames_ts.size(), "frames.size()= " + std::to_string(sc.frames.size()) + " frames_ts.size()= " + std::to_string(frames_ts.size()));

  const int64_t dt_ns = static_cast<int64_t>(spline_knots_dt_s * 1e9);
  const int64_t start_t_ns = knots_ts.front() * 1e9;
  BSplineSO3R3<order> spline_wb(dt_ns, start_t_ns, device);
  const ColmapCamera& c0 = cameras[0];
  spline_wb.setCamera(c0.model_id, {{c0.fx, c0.fy, c0.cx, c0.cy}}, {{c0.dist_coeffs[0], c0.dist_coeffs[1], c0.dist_coeffs[2], c0.dist_coeffs[3]}}, T_bc,
                      calib.timeshift_cam_imu_);
  spline_wb.setImu(noise);
  spline_wb.initializeFromNumTraj(knots_q_wb, knots_r_wb);
  spline_wb.setOptimizeCamImuExtrinsics(config.as<bool>("optimize_camimu_extrinsics"));
  spline_wb.setOptimizeTimeOffsetSplineGlobalSensor(config.as<bool>("optimize_spline_globsens_time_offset"));
  spline_wb.setGravity(Vec3{{0.0, 0.0, 9.80665}});                                  // ImuCalibration::g_W_ (sensor/imu_calib.h:33-41)
  const bool optimize_gravity = config.as<bool>("optimize_gravity");
  spline_wb.setOptimizeGravity(optimize_gravity);
  if (optimize_gravity) spline_wb.addGravityNormError();
  double t_cam_gp_init = 0;
  require(loadScalar<double>(alignment_dir + "/spline_globalsensor_time_offset.txt", t_cam_gp_init), "Could not load spline global sensor time offset");
  spline_wb.setInitSplineGlobSensTimeOffset(t_cam_gp_init);
  Vec3 p_bp;
  require(loadVector3d(alignment_dir + "/p_bp.txt", p_bp), "Could not load p_bp");
  spline_wb.setPbp(p_bp);
  spline_wb.setOptimizePbp(config.as<bool>("optimize_pBP"));
  spline_wb.setScale(scale);
  spline_wb.setOptimizeScale(config.as<bool>("optimize_scale"));
  spline_wb.addRotationLocalParametrization();
  spline_wb.set3DPoints(points_w);
  const double dt_bias_spline_s = config.as<double>("bias_spline_control_nodes_dt_s");
  const std::vector<double> ab = config.asList("init_acc_bias"), gb = config.asList("init_gyro_bias");
  require(ab.size() == 3 && gb.size() == 3, "init_acc_bias / init_gyro_bias must hold 3 values");
  spline_wb.initializeBiases(dt_bias_spline_s, Vec3{{ab[0], ab[1], ab[2]}}, Vec3{{gb[0], gb[1], gb[2]}});

  const int F = (int)sc.frames.size();
  std::vector<int64_t> frame_t_ns(F);
  for (int f = 0; f < F; ++f) frame_t_ns[f] = static_cast<int64_t>(frames_ts[f] * 1e9);
  std::vector<uint8_t> valid(std::max<size_t>(sc.lm_index.size(), 1)); std::vector<int32_t> n_valid(std::max(F, 1));
  int rc = ctba_compute_valid_reprojections(spline_wb.handle(), F, frame_t_ns.data(), sc.pt_begin.data(), sc.uv.data(), sc.lm_index.data(), (int)sc.frames[0].w,
                                            (int)sc.frames[0].h, grid_cell_size, valid.data(), n_valid.data());
  if (rc < 0) throw Error(rc, ctba_last_error(spline_wb.handle()));
  const double gp_meas_std = gpStdForDataset(config, dataset_fn);
  std::cout << "Sigma global measurements: " << gp_meas_std << " [m]\n\n";
  {
    std::vector<int64_t> t(gp_meas.size()); std::vector<double> z(3 * gp_meas.size());
    for (size_t i = 0; i < gp_meas.size(); i++) { t[i] = static_cast<int64_t>(gp_meas[i].t * 1e9); for (int c = 0; c < 3; ++c) z[3 * i + c] = gp_meas[i].pos[c]; }
    rc = ctba_add_gp(spline_wb.handle(), (int)t.size(), t.data(), z.data(), gp_meas_std);
    if (rc < 0) throw Error(rc, ctba_last_error(spline_wb.handle()));
    std::cout << "Global position samples in the problem: " << rc << " of " << t.size() << "\n\n";
  }
  {
    std::vector<int64_t> t(imu_meas.size()); std::vector<Vec3> a(imu_meas.size()), w(imu_meas.size());
    for (size_t i = 0; i < imu_meas.size(); i++) { t[i] = static_cast<int64_t>(imu_meas[i].t * 1e9); a[i] = imu_meas[i].lin_acc; w[i] = imu_meas[i].ang_vel; }
    std::cout << "IMU samples in the problem: " << spline_wb.addImuMeasurements(t, a, w) << " of " << t.size() << "\n\n";
  }
  const double sigma_reproj_px = config.as<double>("sigma_reproj_px");
  std::cout << "Sigma reprojection errors: " << sigma_reproj_px << " [pixel]\n\n";
  {
    std::vector<int32_t> ob(F + 1, 0), lm; std::vector<double> uv;
    for (int f = 0; f < F; ++f) {
      for (int q = sc.pt_begin[f]; q < sc.pt_begin[f + 1]; ++q)
        if (valid[q]) { uv.push_back(sc.uv[2 * q]); uv.push_back(sc.uv[2 * q + 1]); lm.push_back(sc.lm_index[q]); }
      ob[f + 1] = (int32_t)lm.size();
    }
    rc = ctba_add_frames(spline_wb.handle(), F, frame_t_ns.data(), ob.data(), uv.data(), lm.data(), sigma_reproj_px);
    if (rc < 0) throw Error(rc, ctba_last_error(spline_wb.handle()));
    std::cout << "Frames in the problem: " << rc << " of " << F << ", valid reprojections: " << lm.size() << "\n\n";
  }
  const double bias_err_freq_hz = config.as<double>("bias_err_freq_hz"), bias_err_dt_s = config.as<double>("bias_err_dt_s");
  require(bias_err_dt_s >= 4 * dt_bias_spline_s, "Increase bias error dt such that bias_err_dt_s >= 4 * dt_bias_spline_s.");
  spline_wb.addImuBiasCost(bias_err_freq_hz, bias_err_dt_s);

  auto savePredictedImu = [&](bool optimized) {                                     // savePredictedImuMeasurements (:86-127) as two device calls
    const int64_t t0 = spline_wb.getStartTns(), t1 = spline_wb.getEndTns();
    const double off = spline_wb.getSplineImuTimeOffset();
    std::vector<int64_t> ts; std::vector<double> stamps;
    for (const ImuMeasurement& m : imu_meas) {
      const int64_t t_ns = static_cast<int64_t>((m.t - off) * 1e9);
      if (t_ns < t0) continue;
      if (t_ns > t1) break;
      ts.push_back(t_ns); stamps.push_back(m.t);
    }
    std::vector<double> acc(3 * ts.size()), gyr(3 * ts.size());
    if (!ts.empty()) {
      int r = ctba_sample(spline_wb.handle(), CTBA_SAMPLE_ACCEL_BODY, (int64_t)ts.size(), ts.data(), acc.data());
      if (r == CTBA_OK) r = ctba_sample(spline_wb.handle(), CTBA_SAMPLE_GYRO_BODY, (int64_t)ts.size(), ts.data(), gyr.data());
      if (r < 0) throw Error(r, ctba_last_error(spline_wb.handle()));
    }
    std::ofstream f; createTrace(trace_dir, optimized ? "predicted_imu.txt" : "initial_predicted_imu.txt", &f); f.precision(15);
    f << "# timestamp wx wy wz ax ay az" << std::endl;
    for (size_t i = 0; i < ts.size(); ++i)
      f << std::setprecision(15) << stamps[i] << " " << gyr[3 * i] << " " << gyr[3 * i + 1] << " " << gyr[3 * i + 2] << " " << acc[3 * i] << " " << acc[3 * i + 1] << " " << acc[3 * i + 2] << std::endl;
  };
  auto saveBiasSplines = [&]() {                                                    // :63-85, 10 Hz up to getEndTnsBiasSpline (:1870-1875)
    const int64_t dt_ns_bias = static_cast<int64_t>(dt_bias_spline_s * 1e9);
    const int64_t t_end = (int64_t)((double)(((int64_t)spline_wb.numBiasKnots() - order + 1) * dt_ns_bias + spline_wb.getStartTns()) - 1e6);   // getEndTnsBiasSpline: through double
    std::vector<int64_t> ts;
    for (int64_t t = spline_wb.getStartTns(); t < t_end; t += 100000000) ts.push_back(t);
    for (int which = 0; which < 2; ++which) {
      std::vector<double> b(3 * ts.size());
      if (!ts.empty()) { const int r = ctba_sample(spline_wb.handle(), which ? CTBA_SAMPLE_GYRO_BIAS : CTBA_SAMPLE_ACCEL_BIAS, (int64_t)ts.size(), ts.data(), b.data()); if (r < 0) throw Error(r, ctba_last_error(spline_wb.handle())); }
      std::ofstream f; createTrace(trace_dir, which ? "gyro_bias_spline.txt" : "acc_bias_spline.txt", &f); f.precision(9);
      f << "# timestamp x y z" << std::endl;
      for (size_t i = 0; i < ts.size(); ++i) f << std::setprecision(15) << ts[i] * 1e-9 << " " << b[3 * i] << " " << b[3 * i + 1] << " " << b[3 * i + 2] << std::endl;
    }
  };

  auto imuCamRelTransformation = [&]() {                                            // getImuCamRelTransformation (:1813-1817): T_b_c * (R_b_c_corr, p_b_c_corr)
    std::array<double, 19> gl{};
    const int r = ctba_get_globals(spline_wb.handle(), gl.data());
    if (r < 0) throw Error(r, ctba_last_error(spline_wb.handle()));
    const Vec3 pc = rotate(T_bc.q, Vec3{{gl[5], gl[6], gl[7]}});
    return Pose{{{T_bc.p[0] + pc[0], T_bc.p[1] + pc[1], T_bc.p[2] + pc[2]}}, quatMul(T_bc.q, Quat{{gl[1], gl[2], gl[3], gl[4]}})};
  };

  std::cout << "+-+-+-+-+- Initial Errors +-+-+-+-+-\n\n";
  int n_err = 0;
  std::vector<double> e_acc = spline_wb.evaluateErrors(CTBA_ACCEL), e_gyr = spline_wb.evaluateErrors(CTBA_GYRO), e_gp = spline_wb.evaluateErrors(CTBA_GP),
                      e_rep = spline_wb.evaluateErrors(CTBA_REPROJ);
  std::cout << "Number of global position errors: " << e_gp.size() << "\nInitial global position error mean: " << meanAbove(e_gp, -1.0, &n_err) << "\n\n";
  std::cout << "Number of imu errors: " << e_acc.size() << "\nInitial accelerometer error mean: " << meanAbove(e_acc, -1.0, &n_err) << "\nInitial gyro error mean: " << meanAbove(e_gyr, -1.0, &n_err) << "\n\n";
  const double m0 = meanAbove(e_rep, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nInitial reprojection error mean: " << m0 << "\n\n";
  std::cout << "Saving initial values ...\n\n";
  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, false);
  savePredictedImu(false);
  save1D(trace_dir, "initial_acc_errors.txt", e_acc); save1D(trace_dir, "initial_gyro_errors.txt", e_gyr); save1D(trace_dir, "initial_reprojection_errors.txt", e_rep);
  save1D(trace_dir, "initial_global_position_errors.txt", e_gp);
  savePoints(trace_dir, "initial_points3D.txt", points_w);
  saveScalar(trace_dir, "initial_spline_globalsensor_time_offset.txt", spline_wb.getSplineGlobSensTimeOffset());
  saveScalar(trace_dir, "initial_scale.txt", spline_wb.getScale());
  savePosition(trace_dir, "initial_p_bp.txt", spline_wb.getPbp());
  savePose(trace_dir, "initial_T_bc.txt", imuCamRelTransformation());
  saveScalar(trace_dir, "initial_cam_imu_time_offset.txt", spline_wb.getSplineImuTimeOffset());
  savePosition(trace_dir, "initial_gravity.txt", spline_wb.getGravity());

  const int n_max_outer_iters = config.as<int>("n_max_outer_iters"), n_max_inner_iters = config.as<int>("n_max_inner_iters");
  spline_wb.loadOptimizationParams(n_max_outer_iters, n_max_inner_iters, config.as<int>("n_threads"));
  std::cout << "\n\nOptimizing ...\n\n";
  Summary summary = spline_wb.optimize(1);
  spline_wb.relinearizeSplineGlobSensTimeOffset(); spline_wb.relinearizeSplineImuTimeOffset();
  std::printf("solve 0: iterations %zu, cost %.12g -> %.12g\n", summary.iterations.size(), summary.initial_cost, summary.final_cost);
  for (int i = 0; i < n_max_outer_iters; i++) {
    summary = spline_wb.optimize(n_max_inner_iters);
    spline_wb.relinearizeSplineGlobSensTimeOffset(); spline_wb.relinearizeSplineImuTimeOffset();
    std::printf("solve %d: iterations %zu, cost %.12g -> %.12g, termination %d\n", i + 1, summary.iterations.size(), summary.initial_cost, summary.final_cost, summary.termination_type);
    if (summary.termination_type == 0) { std::cout << "\n===== Solver converged =====\n\nnumber of outer iterations = " << i + 1 << "\n"; break; }
    if (i == n_max_outer_iters - 1) std::cout << "\n\n===== Solver did not converge =====\n\n";
  }
  spline_wb.relinearizeSplineGlobSensTimeOffset(); spline_wb.relinearizeSplineImuTimeOffset();

  std::cout << "+-+-+-+-+- Final Errors +-+-+-+-+-\n\n";
  e_acc = spline_wb.evaluateErrors(CTBA_ACCEL); e_gyr = spline_wb.evaluateErrors(CTBA_GYRO); e_rep = spline_wb.evaluateErrors(CTBA_REPROJ);
  e_gp = spline_wb.evaluateErrors(CTBA_GP);
  std::cout << "Number of global position errors: " << e_gp.size() << "\nGlobal position error mean: " << meanAbove(e_gp, -1.0, &n_err) << "\n\n";
  std::cout << "Number of imu errors: " << e_acc.size() << "\nAccelerometer error mean: " << meanAbove(e_acc, -1.0, &n_err) << "\nGyro error mean: " << meanAbove(e_gyr, -1.0, &n_err) << "\n\n";
  const double m1 = meanAbove(e_rep, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nReprojection error mean: " << m1 << "\n\n";
  const Pose T_bc_opt = imuCamRelTransformation();
  std::cout << "Optimized spline - imu time offset: " << spline_wb.getSplineImuTimeOffset() << " [s]\nOptimized scale: " << spline_wb.getScale() << "\n\n";

  std::cout << "Saving optimized values ...\n\n";
  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, true);
  savePredictedImu(true);
  saveBiasSplines();
  save1D(trace_dir, "acc_errors.txt", e_acc); save1D(trace_dir, "gyro_errors.txt", e_gyr); save1D(trace_dir, "reprojection_errors.txt", e_rep);
  save1D(trace_dir, "global_position_errors.txt", e_gp);
  savePoints(trace_dir, "points3D.txt", spline_wb.get3DPoints());
  saveScalar(trace_dir, "spline_globalsensor_time_offset.txt", spline_wb.getSplineGlobSensTimeOffset());
  saveScalar(trace_dir, "scale.txt", spline_wb.getScale());
  savePosition(trace_dir, "p_bp.txt", spline_wb.getPbp());
  savePose(trace_dir, "T_bc.txt", T_bc_opt);
  saveScalar(trace_dir, "cam_imu_time_offset.txt", spline_wb.getSplineImuTimeOffset());
  savePosition(trace_dir, "gravity.txt", spline_wb.getGravity());
  return 0;
}

int main(int argc, char** argv) {
  require(argc > 1, "Please, pass the .yaml filename");
  try {
    const Yaml config = Yaml::LoadFile(argv[1]);
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    const int order = argc > 3 ? std::atoi(argv[3]) : 6;   // the reference hard-codes `const int order = 6`
    std::cout << "Spline order: " << order << "\n\n";
    if (order == 6) return run<6>(config, device);
    if (order == 4) return run<4>(config, device);
    require(false, "spline order must be 4 or 6");
  } catch (const ctba::Error& e) {
    std::cerr << "ctba error " << e.code << ": " << e.what() << "\n";
    return 3;
  } catch (const std::exception& e) {
    std::cerr << e.what() << "\n";
    return 2;
  }
  return 0;
}
