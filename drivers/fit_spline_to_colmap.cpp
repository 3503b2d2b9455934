// fit_spline_to_colmap — the reference tool of the same name re-hosted on the C ABI:
//     /root/reference/src/fit_spline_to_colmap.cpp:62-260
// Same YAML keys (experiments/*/*/colmap_fitted_spline/*.yaml: colmap_fn, colmap_spline_dir, start_t, end_t, spline_control_nodes_dt_s,
// num_errors, n_max_iters, n_threads), same initialisation (knots linearly interpolated on the COLMAP body poses), same measurements
// (num_errors position + orientation samples, noise 1) and the same result files (spline_knots.txt, spline.txt under
// <colmap_spline_dir>/colmap_fitted_spline/order_<N>/dt_<ms>_ms/).  Solve() is ctba_solve on the GPU (band solver with pose priors).
//   usage: fit_spline_to_colmap <config.yaml> [cuda device] [order 4|6]
#include <cstdio>
#include <cstdlib>
#include <sys/stat.h>
#include "ctba_io.hpp"

using namespace ctba;
using namespace ctba::io;

static void require(bool ok, const std::string& msg) {   // GVI_FUSION_ASSERT_STREAM (util/assert.h:65-70)
  if (!ok) { std::cerr << msg << "\n"; std::exit(2); }
}
static double meanOf(const std::vector<double>& v, size_t b, size_t e) { double s = 0; for (size_t i = b; i < e; ++i) s += v[i]; return e > b ? s / (double)(e - b) : 0.0; }

template <int order>
static int run(const Yaml& config, int device) {
  const std::string colmap_fn = config.as<std::string>("colmap_fn");
  std::string colmap_spline_dir = config.as<std::string>("colmap_spline_dir");
  mkdir(colmap_spline_dir.c_str(), 0777);
  colmap_spline_dir += "/colmap_fitted_spline"; mkdir(colmap_spline_dir.c_str(), 0777);
  const double colmap_start_t = config.as<double>("start_t"), colmap_end_t = config.as<double>("end_t");
  const double dt_s = config.as<double>("spline_control_nodes_dt_s");
  const int dt_ms = dt_s * 1e3;
  std::string trace_dir = colmap_spline_dir + "/order_" + std::to_string(order); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/dt_" + std::to_string(dt_ms) + "_ms"; mkdir(trace_dir.c_str(), 0777);
  std::cout << "+-+-+-+-+- Files +-+-+-+-+-\ncolmap_fn: " << colmap_fn << "\noutput dir: " << trace_dir << "\n";

  std::vector<Quat> colmap_r_gb; std::vector<Vec3> colmap_p_gb; std::vector<double> colmap_ts;
  require(loadTrajectoryFromFile(colmap_fn, colmap_r_gb, colmap_p_gb, colmap_ts), "Could not load colmap trajectory.");
  std::cout << "Loaded " << colmap_p_gb.size() << " colmap estimates.\n\n";
  std::vector<Quat> sampled_r; std::vector<Vec3> sampled_p; std::vector<double> sampled_ts;
  for (size_t i = 0; i < colmap_ts.size(); i++)
    if (colmap_ts[i] >= colmap_start_t && colmap_ts[i] <= colmap_end_t) { sampled_p.push_back(colmap_p_gb[i]); sampled_r.push_back(colmap_r_gb[i]); sampled_ts.push_back(colmap_ts[i]); }
  require(sampled_ts.size() >= 2, "no colmap estimates between start_t and end_t");
  const double start_knot_t_s = sampled_ts.front(), end_knot_t_s = sampled_ts.back();

  const int64_t dt_ns = static_cast<int64_t>(dt_s * 1e9);
  const int64_t start_t_ns = start_knot_t_s * 1e9;
  BSplineSO3R3<order> spline_gb(dt_ns, start_t_ns, device);
  std::vector<double> knots_ts;
  for (double knot_t = start_knot_t_s; knot_t <= end_knot_t_s; knot_t += dt_s) knots_ts.push_back(knot_t);   // :140-146
  std::vector<Quat> init_knots_r; std::vector<Vec3> init_knots_p;
  linearInterpolation(sampled_p, sampled_r, sampled_ts, knots_ts, init_knots_p, init_knots_r);
  spline_gb.initializeFromNumTraj(init_knots_r, init_knots_p);
  spline_gb.addRotationLocalParametrization();

  const double n_errs = config.as<double>("num_errors");
  const double dt_meas = (sampled_ts.back() - sampled_ts.front()) / n_errs;
  std::vector<double> meas_ts;
  for (double meas_t = sampled_ts.front(); meas_t <= sampled_ts.back(); meas_t += dt_meas) meas_ts.push_back(meas_t);   // :157-164
  std::vector<Quat> meas_r; std::vector<Vec3> meas_p;
  linearInterpolation(sampled_p, sampled_r, sampled_ts, meas_ts, meas_p, meas_r);
  // measurements outside the spline are skipped by the add* members, as in the reference (:1055-1062); the error vectors count what was kept
  size_t n_pos = 0, n_ori = 0;
  for (size_t i = 0; i < meas_ts.size(); i++) n_pos += spline_gb.addPositionMeasurement(meas_p[i], static_cast<int64_t>(meas_ts[i] * 1e9), 1.0);
  for (size_t i = 0; i < meas_ts.size(); i++) n_ori += spline_gb.addOrientationMeasurement(meas_r[i], static_cast<int64_t>(meas_ts[i] * 1e9), 1.0);

  // evaluatePositionErrors / evaluateOrientationErrors (:1581-1680): the priors in the order they were added, positions first
  std::vector<double> e0 = spline_gb.evaluateErrors(CTBA_PRIOR);
  require(e0.size() == n_pos + n_ori, "prior count mismatch");
  std::cout << "Number of position errors: " << n_pos << "\nInitial position error mean: " << meanOf(e0, 0, n_pos) << " [m]\n\n";
  std::cout << "Number of orientation errors: " << n_ori << "\nInitial orientation error mean: " << meanOf(e0, n_pos, e0.size()) << " [rad]\n\n";

  const int n_max_iters = config.as<int>("n_max_iters");
  spline_gb.loadOptimizationParams(n_max_iters, 0, config.as<int>("n_threads"));
  std::cout << "\n\nOptimizing ...\n\n";
  Summary summary = spline_gb.optimize(n_max_iters);
  std::printf("iterations %zu, cost %.9g -> %.9g, termination %d, %.3f ms\n", summary.iterations.size(), summary.initial_cost, summary.final_cost,
              summary.termination_type, 1e3 * summary.total_time_in_seconds);

  std::vector<double> e1 = spline_gb.evaluateErrors(CTBA_PRIOR);
  std::cout << "Number of position errors: " << n_pos << "\nPosition error mean: " << meanOf(e1, 0, n_pos) << " [m]\n\n";
  std::cout << "Number of orientation errors: " << n_ori << "\nOrientation error mean: " << meanOf(e1, n_pos, e1.size()) << " [rad]\n\n";

  std::cout << "Saving spline ...\n\n";
  saveSplineKnots(trace_dir, spline_gb, knots_ts);
  saveSpline(trace_dir, spline_gb);
  return 0;
}

int main(int argc, char** argv) {
  require(argc > 1, "Please, pass the .yaml filename");
  try {
    const Yaml config = Yaml::LoadFile(argv[1]);
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    const int order = argc > 3 ? std::atoi(argv[3]) : 6;   // the reference hard-codes `const int order = 6` (:76)
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
