// fit_spline_to_gp_measurements — the reference tool of the same name re-hosted on the C ABI:
//     /root/reference/src/fit_spline_to_gp_measurements.cpp:67-208
// Same YAML keys (gp_fn, spline_dir, start_t, end_t, spline_control_nodes_dt_s, num_errors, n_max_iters, n_threads): a spline is fitted to
// the global position measurements alone (position priors of noise 1, identity orientations), result files spline_knots.txt and spline.txt
// under <spline_dir>/fit_spline_on_gp_meas/order_<N>/dt_<ms>_ms/.  The rotation knots are not part of the problem (no residual touches
// them), exactly as in the reference where only the position knots are added to the ceres::Problem.
//   usage: fit_spline_to_gp_measurements <config.yaml> [cuda device] [order 4|6]
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
  const std::string gp_fn = config.as<std::string>("gp_fn");
  std::string spline_dir = config.as<std::string>("spline_dir");
  mkdir(spline_dir.c_str(), 0777);
  spline_dir += "/fit_spline_on_gp_meas"; mkdir(spline_dir.c_str(), 0777);
  const double start_t = config.as<double>("start_t"), end_t = config.as<double>("end_t");
  const double dt_s = config.as<double>("spline_control_nodes_dt_s");
  const int dt_ms = dt_s * 1e3;
  std::string trace_dir = spline_dir + "/order_" + std::to_string(order); mkdir(trace_dir.c_str(), 0777);
  trace_dir += "/dt_" + std::to_string(dt_ms) + "_ms"; mkdir(trace_dir.c_str(), 0777);
  std::cout << "+-+-+-+-+- Files +-+-+-+-+-\ngp_fn: " << gp_fn << "\noutput dir: " << trace_dir << "\n";

  std::vector<GlobalPosMeasurement> gp_meas;
  require(loadGpMeasurements(gp_fn, gp_meas), "Could not load global measurements.");
  std::cout << "Loaded " << gp_meas.size() << " global measurements.\n\n";
  std::vector<Vec3> sampled_p; std::vector<Quat> identity_oris; std::vector<double> sampled_ts;
  for (const GlobalPosMeasurement& m : gp_meas)
    if (m.t >= start_t && m.t <= end_t) { sampled_ts.push_back(m.t); sampled_p.push_back(m.pos); identity_oris.push_back(Quat{{0, 0, 0, 1}}); }
  require(sampled_ts.size() >= 2, "no global measurements between start_t and end_t");
  // the knots span ALL measurements, not only the sampled ones (:127-128); if start_t / end_t cut any off, the interpolation below
  // stops short and the size check fails, as in the reference
  const double start_knot_t_s = gp_meas.front().t, end_knot_t_s = gp_meas.back().t;

  const int64_t dt_ns = static_cast<int64_t>(dt_s * 1e9);
  const int64_t start_t_ns = start_knot_t_s * 1e9;
  BSplineSO3R3<order> spline_gp(dt_ns, start_t_ns, device);
  std::vector<double> knots_ts;
  for (double knot_t = start_knot_t_s; knot_t <= end_knot_t_s; knot_t += dt_s) knots_ts.push_back(knot_t);
  std::vector<Quat> init_knots_r; std::vector<Vec3> init_knots_p;
  linearInterpolation(sampled_p, identity_oris, sampled_ts, knots_ts, init_knots_p, init_knots_r);
  spline_gp.initializeFromNumTraj(init_knots_r, init_knots_p);
  spline_gp.addRotationLocalParametrization();

  const double n_errs = config.as<double>("num_errors");
  const double dt_meas = (sampled_ts.back() - sampled_ts.front()) / n_errs;
  std::vector<double> meas_ts;
  for (double meas_t = sampled_ts.front(); meas_t <= sampled_ts.back(); meas_t += dt_meas) meas_ts.push_back(meas_t);
  std::vector<Quat> meas_r; std::vector<Vec3> meas_p;
  linearInterpolation(sampled_p, identity_oris, sampled_ts, meas_ts, meas_p, meas_r);
  size_t n_pos = 0;
  for (size_t i = 0; i < meas_ts.size(); i++) n_pos += spline_gp.addPositionMeasurement(meas_p[i], static_cast<int64_t>(meas_ts[i] * 1e9), 1.0);

  std::vector<double> e0 = spline_gp.evaluateErrors(CTBA_PRIOR);   // evaluatePositionErrors (:1581-1630)
  require(e0.size() == n_pos, "prior count mismatch");
  std::cout << "Number of position errors: " << n_pos << "\nInitial position error mean: " << meanOf(e0, 0, n_pos) << " [m]\n\n";

  const int n_max_iters = config.as<int>("n_max_iters");
  spline_gp.loadOptimizationParams(n_max_iters, 0, config.as<int>("n_threads"));
  std::cout << "\n\nOptimizing ...\n\n";
  Summary summary = spline_gp.optimize(n_max_iters);
  std::printf("iterations %zu, cost %.9g -> %.9g, termination %d, %.3f ms\n", summary.iterations.size(), summary.initial_cost, summary.final_cost,
              summary.termination_type, 1e3 * summary.total_time_in_seconds);
  std::vector<double> e1 = spline_gp.evaluateErrors(CTBA_PRIOR);
  std::cout << "Number of position errors: " << n_pos << "\nPosition error mean: " << meanOf(e1, 0, n_pos) << " [m]\n\n";

  std::cout << "Saving spline ...\n\n";
  saveSplineKnots(trace_dir, spline_gp, knots_ts);
  saveSpline(trace_dir, spline_gp);
  return 0;
}

int main(int argc, char** argv) {
  require(argc > 1, "Please, pass the .yaml filename");
  try {
    const Yaml config = Yaml::LoadFile(argv[1]);
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    const int order = argc > 3 ? std::atoi(argv[3]) : 6;   // the reference hard-codes `const int order = 6` (:74)
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
