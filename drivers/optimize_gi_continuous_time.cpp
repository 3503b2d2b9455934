// optimize_gi_continuous_time — the reference driver of the same name (global position + inertial, no vision) re-hosted on the C ABI:
//     /root/reference/src/optimize_gi_continuous_time.cpp:196-499
// Same YAML keys (experiments/*/*/full_batch_optimization_gi/continuous_time/*.yaml), same inputs (imu.txt, global position measurements,
// the spline_knots.txt fitted to the global positions by fit_spline_to_gp_measurements), the same call order, outer loop and result files.
// As in the reference no camera is set (cam-imu time offset 0), p_bp starts at zero and scale_ keeps its constructor value 0
// (bspline_so3r3.h:101), so the antenna lever arm does not enter the global-position residual; only `dataset: EuRoC` is accepted (:340-362).
// optimize() runs on the GPU (ctba_solve).   usage: optimize_gi_continuous_time <config.yaml> [cuda device] [order 4|6]
#include <cstdio>
#include <cstdlib>
#include <sys/stat.h>
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
template <class Spline>
static void saveSplineAt(const std::string& dir, Spline& spline, double rate_hz, bool optimized) {   // :41-61
  std::ofstream f; createTrace(dir, optimized ? "spline.txt" : "initial_spline.txt", &f); f.precision(15);
  traceSpline(spline, static_cast<int64_t>((1.0 / rate_hz) * 1e9), f);
}
static double meanAbove(const std::vector<double>& v, double thr, int* n) { double s = 0; *n = 0; for (double e : v) if (e > thr) { s += e; ++*n; } return *n ? s / *n : 0.0; }

template <int order>
static int run(const Yaml& config, int device) {
  const std::string imu_fn = config.as<std::string>("imu_fn"), gp_fn = config.as<std::string>("gp_fn");
  const double spline_knots_dt_s = config.as<double>("spline_control_nodes_dt_s"), sampling_spline_rate_hz = config.as<double>("sampling_spline_rate");
  const int spline_knots_dt_ms = spline_knots_dt_s * 1000;
  const std::string sub = "/order_" + std::to_string(order) + "/dt_" + std::to_string(spline_knots_dt_ms) + "_ms";
  const std::string init_spline_fn = config.as<std::string>("spline_dir") + sub + "/spline_knots.txt";
  std::string trace_dir = config.as<std::string>("full_batch_optimization_dir"); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/continuous_time"; mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/order_" + std::to_string(order); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/dt_" + std::to_string(spline_knots_dt_ms) + "_ms"; mkdir(trace_dir.c_str(), 0777);
  const std::string dataset_fn = config.as<std::string>("dataset");
  std::cout << "Dataset: " << dataset_fn << "\nresult dir: " << trace_dir << "\n\n";
  require(dataset_fn == "EuRoC", dataset_fn == "UZH_FPV" ? "GI fusion not implemented on this dataset!" : "Unknown dataset");

  ImuNoise noise;
  noise.acc_noise_density = config.as<double>("sigma_acc_c"); noise.acc_random_walk = config.as<double>("sigma_acc_bias_c");
  noise.gyro_noise_density = config.as<double>("sigma_omega_c"); noise.gyro_random_walk = config.as<double>("sigma_omega_bias_c");
  std::vector<ImuMeasurement> imu_meas;
  require(loadImuMeasurementsFromFile(imu_fn, imu_meas), "Could not imu");
  std::cout << "Loaded " << imu_meas.size() << " imu measurements.\n\n";
  std::vector<GlobalPosMeasurement> gp_meas;
  require(loadGpMeasurements(gp_fn, gp_meas), "Could not load global measurements.");
  std::cout << "Loaded " << gp_meas.size() << " global measurements.\n\n";
  std::vector<Quat> knots_q_wb; std::vector<Vec3> knots_r_wb; std::vector<double> knots_ts;
  require(loadSplineFromFile(init_spline_fn, knots_q_wb, knots_r_wb, knots_ts) && !knots_ts.empty(), "Could not load spline.");
  std::cout << "Loaded " << knots_q_wb.size() << " initial spline knots.\n\n";

  const int64_t dt_ns = static_cast<int64_t>(spline_knots_dt_s * 1e9);
  const int64_t start_t_ns = knots_ts.front() * 1e9;
  BSplineSO3R3<order> spline_wb(dt_ns, start_t_ns, device);
  spline_wb.setImu(noise);
  spline_wb.initializeFromNumTraj(knots_q_wb, knots_r_wb);
  spline_wb.setGravity(Vec3{{0.0, 0.0, 9.80665}});                                  // ImuCalibration::g_W_ (sensor/imu_calib.h:33-41)
  spline_wb.setOptimizeGravity(config.as<bool>("optimize_gravity"));               // no gravity-norm residual in this driver (:309-311)
  spline_wb.setOptimizePbp(config.as<bool>("optimize_pBP"));
  spline_wb.setOptimizeTimeOffsetSplineGlobalSensor(config.as<bool>("optimize_spline_globsens_time_offset"));
  spline_wb.addRotationLocalParametrization();
  const double dt_bias_spline_s = config.as<double>("bias_spline_control_nodes_dt_s");
  const std::vector<double> ab = config.asList("init_acc_bias"), gb = config.asList("init_gyro_bias");
  require(ab.size() == 3 && gb.size() == 3, "init_acc_bias / init_gyro_bias must hold 3 values");
  spline_wb.initializeBiases(dt_bias_spline_s, Vec3{{ab[0], ab[1], ab[2]}}, Vec3{{gb[0], gb[1], gb[2]}});

  const double gp_meas_std = config.as<double>("gp_std");
  std::cout << "Sigma global measurements: " << gp_meas_std << " [m]\n\n";
  int rc = 0;
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

  std::cout << "+-+-+-+-+- Initial Errors +-+-+-+-+-\n\n";
  int n_err = 0;
  std::vector<double> e_acc = spline_wb.evaluateErrors(CTBA_ACCEL), e_gyr = spline_wb.evaluateErrors(CTBA_GYRO), e_gp = spline_wb.evaluateErrors(CTBA_GP);
  std::cout << "Number of imu errors: " << e_acc.size() << "\nInitial accelerometer error mean: " << meanAbove(e_acc, -1.0, &n_err) << "\nInitial gyro error mean: " << meanAbove(e_gyr, -1.0, &n_err) << "\n\n";
  std::cout << "Number of global position errors: " << e_gp.size() << "\nInitial global position error mean: " << meanAbove(e_gp, -1.0, &n_err) << "\n\n";
  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, false);
  savePredictedImu(false);
  save1D(trace_dir, "initial_acc_errors.txt", e_acc); save1D(trace_dir, "initial_gyro_errors.txt", e_gyr); save1D(trace_dir, "initial_global_position_errors.txt", e_gp);
  saveScalar(trace_dir, "initial_spline_globalsensor_time_offset.txt", spline_wb.getSplineGlobSensTimeOffset());
  savePosition(trace_dir, "initial_p_bp.txt", spline_wb.getPbp());
  savePosition(trace_dir, "initial_gravity.txt", spline_wb.getGravity());

  const int n_max_outer_iters = config.as<int>("n_max_outer_iters"), n_max_inner_iters = config.as<int>("n_max_inner_iters");
  spline_wb.loadOptimizationParams(n_max_outer_iters, n_max_inner_iters, config.as<int>("n_threads"));
  std::cout << "\n\nOptimizing ...\n\n";
  // only the spline-imu offset is re-linearised between the solves (:425-451); t_s_gp_corr stays in place
  Summary summary = spline_wb.optimize(1);
  spline_wb.relinearizeSplineImuTimeOffset();
  std::printf("solve 0: iterations %zu, cost %.12g -> %.12g\n", summary.iterations.size(), summary.initial_cost, summary.final_cost);
  for (int i = 0; i < n_max_outer_iters; i++) {
    summary = spline_wb.optimize(n_max_inner_iters);
    spline_wb.relinearizeSplineImuTimeOffset();
    std::printf("solve %d: iterations %zu, cost %.12g -> %.12g, termination %d\n", i + 1, summary.iterations.size(), summary.initial_cost, summary.final_cost, summary.termination_type);
    if (summary.termination_type == 0) { std::cout << "\n===== Solver converged =====\n\nnumber of outer iterations = " << i + 1 << "\n"; break; }
    if (i == n_max_outer_iters - 1) std::cout << "\n\n===== Solver did not converge =====\n\n";
  }
  spline_wb.relinearizeSplineImuTimeOffset();

  std::cout << "+-+-+-+-+- Final Errors +-+-+-+-+-\n\n";
  e_acc = spline_wb.evaluateErrors(CTBA_ACCEL); e_gyr = spline_wb.evaluateErrors(CTBA_GYRO); e_gp = spline_wb.evaluateErrors(CTBA_GP);
  std::cout << "Number of imu errors: " << e_acc.size() << "\nAccelerometer error mean: " << meanAbove(e_acc, -1.0, &n_err) << "\nGyro error mean: " << meanAbove(e_gyr, -1.0, &n_err) << "\n\n";
  std::cout << "Number of global position errors: " << e_gp.size() << "\nGlobal position error mean: " << meanAbove(e_gp, -1.0, &n_err) << "\n\n";
  std::cout << "Saving optimized values ...\n\n";
  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, true);
  savePredictedImu(true);
  saveBiasSplines();
  save1D(trace_dir, "acc_errors.txt", e_acc); save1D(trace_dir, "gyro_errors.txt", e_gyr); save1D(trace_dir, "global_position_errors.txt", e_gp);
  savePosition(trace_dir, "p_bp.txt", spline_wb.getPbp());
  saveScalar(trace_dir, "globalsensor_imu_time_offset.txt", spline_wb.getSplineImuTimeOffset());
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
