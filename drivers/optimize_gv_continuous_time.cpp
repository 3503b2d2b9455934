// optimize_gv_continuous_time — the reference driver of the same name (visual + global position, BASELINE config 3) re-hosted on the C ABI:
//     /root/reference/src/optimize_gv_continuous_time.cpp:173-565
// Same YAML keys (experiments/*/*/full_batch_optimization_gv/continuous_time/*.yaml), same inputs (EuRoC-format calibration YAML, global
// position measurements, COLMAP cameras.bin / images.bin / points3D.bin / colmap_cam_estimates.txt, the fitted spline_knots.txt, T_wg_init.txt /
// scale_init.txt / p_bp.txt of the alignment step), the same call order on the spline object, the same outer loop (1 iteration, then up to
// n_max_outer_iters x n_max_inner_iters with the time offset re-linearised in between) and the same result files.  Everything the
// reference does inside optimize() runs on the GPU (ctba_solve); computeValidReprojections stays a host pass (ctba_compute_valid_reprojections).
// Difference on the host side: points3D.txt rows are in points3D.bin
// order (the reference writes its unordered_map's iteration order, save.h:200-208).
//   usage: optimize_gv_continuous_time <config.yaml> [cuda device] [order 4|6]
#include <chrono>
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
  const std::string gp_fn = config.as<std::string>("gp_fn"), colmap_dir = config.as<std::string>("colmap_dir");
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

  // wall-clock phases of the tool ("[timing]" lines): decides whether ingest / selection are worth moving to the device
  auto t_phase = std::chrono::steady_clock::now();
  auto lap = [&](const char* what) {
    const auto now = std::chrono::steady_clock::now();
    std::printf("[timing] %s: %.3f s\n", what, std::chrono::duration<double>(now - t_phase).count());
    t_phase = now;
  };
  SimpleCamImuCalib calib;
  require(loadCalibForDataset(config, dataset_fn, cam_id, calib), "Camera-Imu calibration could not be loaded");   // UZH_FPV: Kalibr camchain + imu.yaml; EuRoC / Custom: EuRoC format
  std::vector<GlobalPosMeasurement> gp_meas;
  require(loadGpMeasurements(gp_fn, gp_meas), "Could not load global measurements.");
  std::cout << "Loaded " << gp_meas.size() << " global measurements.\n\n";

  std::vector<ColmapCamera> cameras;
  readCamerasBinary(colmap_dir + "cameras.bin", cameras);
  require(!cameras.empty(), "no cameras");
  std::cout << "Loaded: " << cameras.size() << " cameras.\n\n";
  const Pose T_bc = poseInverse(calib.T_cam_imu_);                                  // :267-270
  ColmapScene sc;
  const int grid_cell_size = config.as<int>("grid_cell_size");
  readImagesAndPoints3DBinary(colmap_dir + "images.bin", colmap_dir + "points3D.bin", cameras, sc);
  std::cout << "Loaded: " << sc.frames.size() << " images.\n\nLoaded: " << sc.lm_id.size() << " 3D points.\n\n";
  lap("calibration + global positions + COLMAP binary ingest (cameras.bin, images.bin, points3D.bin)");
  for (int32_t l : sc.lm_index) require(l >= 0, "images.bin refers to a 3D point that points3D.bin does not hold");

  std::vector<Quat> knots_q_gb; std::vector<Vec3> knots_r_gb; std::vector<double> knots_ts;
  require(loadSplineFromFile(init_spline_fn, knots_q_gb, knots_r_gb, knots_ts) && !knots_ts.empty(), "Could not load spline.");
  std::cout << "Loaded " << knots_q_gb.size() << " initial spline knots.\n\n";
  Pose T_wg; double scale = 0;
  require(loadPose(alignment_dir + "/T_wg_init.txt", T_wg), "Could not T_wg");
  require(loadScalar<double>(alignment_dir + "/scale_init.txt", scale), "Could not scale");
  std::cout << "Loaded scale: " << scale << "\n\n";
  // knots and 3-D points from the COLMAP frame g to the world frame w (:320-337)
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
  const ColmapCamera& c0 = cameras[0];                                              // setCameras: the spline uses cameras_[0] (:1899-1903)
  spline_wb.setCamera(c0.model_id, {{c0.fx, c0.fy, c0.cx, c0.cy}}, {{c0.dist_coeffs[0], c0.dist_coeffs[1], c0.dist_coeffs[2], c0.dist_coeffs[3]}}, T_bc,
                      calib.timeshift_cam_imu_);
  spline_wb.initializeFromNumTraj(knots_q_wb, knots_r_wb);
  spline_wb.setInitSplineGlobSensTimeOffset(0.0);
  Vec3 p_bp;
  require(loadVector3d(alignment_dir + "/p_bp.txt", p_bp), "Could not load p_bp");
  spline_wb.setPbp(p_bp);
  spline_wb.setOptimizePbp(config.as<bool>("optimize_pBP"));
  spline_wb.setOptimizeTimeOffsetSplineGlobalSensor(config.as<bool>("optimize_spline_globsens_time_offset"));
  spline_wb.setScale(scale);
  spline_wb.setOptimizeScale(config.as<bool>("optimize_scale"));
  spline_wb.addRotationLocalParametrization();
  spline_wb.set3DPoints(points_w);

  lap("spline / alignment files, frame transforms, handle set-up");
  // computeValidReprojections (:392-406) on every kept 2-D point, then the valid ones enter the problem
  const int F = (int)sc.frames.size();
  std::vector<int64_t> frame_t_ns(F);
  for (int f = 0; f < F; ++f) frame_t_ns[f] = static_cast<int64_t>(frames_ts[f] * 1e9);
  std::vector<uint8_t> valid(std::max<size_t>(sc.lm_index.size(), 1)); std::vector<int32_t> n_valid(std::max(F, 1));
  int rc = ctba_compute_valid_reprojections(spline_wb.handle(), F, frame_t_ns.data(), sc.pt_begin.data(), sc.uv.data(), sc.lm_index.data(), (int)sc.frames[0].w,
                                            (int)sc.frames[0].h, grid_cell_size, valid.data(), n_valid.data());
  if (rc < 0) throw Error(rc, ctba_last_error(spline_wb.handle()));
  lap("computeValidReprojections (host, ctba_compute_valid_reprojections)");

  const double gp_meas_std = gpStdForDataset(config, dataset_fn);
  std::cout << "Sigma global measurements: " << gp_meas_std << " [m]\n\n";
  {
    std::vector<int64_t> t(gp_meas.size()); std::vector<double> z(3 * gp_meas.size());
    for (size_t i = 0; i < gp_meas.size(); i++) { t[i] = static_cast<int64_t>(gp_meas[i].t * 1e9); for (int c = 0; c < 3; ++c) z[3 * i + c] = gp_meas[i].pos[c]; }
    rc = ctba_add_gp(spline_wb.handle(), (int)t.size(), t.data(), z.data(), gp_meas_std);
    if (rc < 0) throw Error(rc, ctba_last_error(spline_wb.handle()));
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

  lap("measurements into the problem (ctba_add_gp, ctba_add_frames: host work items)");
  std::cout << "+-+-+-+-+- Initial Errors +-+-+-+-+-\n\n";
  int n_err = 0;
  std::vector<double> init_gp_errors = spline_wb.evaluateErrors(CTBA_GP), init_reproj_errors = spline_wb.evaluateErrors(CTBA_REPROJ);
  std::cout << "Number of global position errors: " << init_gp_errors.size() << "\nInitial global position error mean: " << meanAbove(init_gp_errors, -1.0, &n_err) << "\n\n";
  const double m0 = meanAbove(init_reproj_errors, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nInitial reprojection error mean: " << m0 << "\n\n";
  std::cout << "Saving initial values ...\n\n";
  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, false);
  save1D(trace_dir, "initial_global_position_errors.txt", init_gp_errors); save1D(trace_dir, "initial_reprojection_errors.txt", init_reproj_errors);
  savePoints(trace_dir, "initial_points3D.txt", points_w);
  saveScalar(trace_dir, "initial_spline_globalsensor_time_offset.txt", spline_wb.getSplineGlobSensTimeOffset());
  saveScalar(trace_dir, "initial_scale.txt", spline_wb.getScale());
  savePosition(trace_dir, "initial_p_bp.txt", spline_wb.getPbp());

  const int n_max_outer_iters = config.as<int>("n_max_outer_iters"), n_max_inner_iters = config.as<int>("n_max_inner_iters");
  spline_wb.loadOptimizationParams(n_max_outer_iters, n_max_inner_iters, config.as<int>("n_threads"));
  lap("initial errors (first upload + evaluation) and initial result files");
  std::cout << "\n\nOptimizing ...\n\n";
  Summary summary = spline_wb.optimize(1);                                          // :485-486
  spline_wb.relinearizeSplineGlobSensTimeOffset();
  std::printf("solve 0: iterations %zu, cost %.12g -> %.12g\n", summary.iterations.size(), summary.initial_cost, summary.final_cost);
  for (int i = 0; i < n_max_outer_iters; i++) {
    summary = spline_wb.optimize(n_max_inner_iters);
    spline_wb.relinearizeSplineGlobSensTimeOffset();
    std::printf("solve %d: iterations %zu, cost %.12g -> %.12g, termination %d\n", i + 1, summary.iterations.size(), summary.initial_cost, summary.final_cost, summary.termination_type);
    if (summary.termination_type == 0) { std::cout << "\n===== Solver converged =====\n\nnumber of outer iterations = " << i + 1 << "\n"; break; }
    if (i == n_max_outer_iters - 1) std::cout << "\n\n===== Solver did not converge =====\n\n";
  }
  spline_wb.relinearizeSplineGlobSensTimeOffset();
  lap("all solves (structure analysis + LM iterations on the device)");

  std::cout << "+-+-+-+-+- Final Errors +-+-+-+-+-\n\n";
  std::vector<double> opti_gp_errors = spline_wb.evaluateErrors(CTBA_GP), opti_reproj_errors = spline_wb.evaluateErrors(CTBA_REPROJ);
  std::cout << "Number of global position errors: " << opti_gp_errors.size() << "\nGlobal position error mean: " << meanAbove(opti_gp_errors, -1.0, &n_err) << "\n\n";
  const double m1 = meanAbove(opti_reproj_errors, 1e-6, &n_err);
  std::cout << "Number of reprojection errors: " << n_err << "\nReprojection error mean: " << m1 << "\n\n";
  std::cout << "Optimized spline - global sensor time offset: " << spline_wb.getSplineGlobSensTimeOffset() << " [s]\n";
  std::cout << "Optimized colmap - global reference scale: " << spline_wb.getScale() << "\nScale correction was: " << spline_wb.getScaleCorr() << "\n\n";

  saveSplineAt(trace_dir, spline_wb, sampling_spline_rate_hz, true);
  // the reference passes `false` here too (:551): the final errors overwrite the initial_* files; reproduced
  save1D(trace_dir, "initial_global_position_errors.txt", opti_gp_errors); save1D(trace_dir, "initial_reprojection_errors.txt", opti_reproj_errors);
  savePoints(trace_dir, "points3D.txt", spline_wb.get3DPoints());
  saveScalar(trace_dir, "spline_globalsensor_time_offset.txt", spline_wb.getSplineGlobSensTimeOffset());
  saveScalar(trace_dir, "scale.txt", spline_wb.getScale());
  savePosition(trace_dir, "p_bp.txt", spline_wb.getPbp());
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
