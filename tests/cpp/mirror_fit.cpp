// fit_spline_to_colmap-shaped run through the C++ mirror on the GPU (src/fit_spline_to_colmap.cpp:139-230): a spline initialised off
// a circular trajectory is fitted to position and orientation samples of it, then sampled like traceSpline does.
#include <cmath>
#include <cstdio>
#include "ctba_bspline_so3r3.hpp"

#define REQUIRE(c) do { if (!(c)) { std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static ctba::Vec3 traj_p(double t) { return {2.0 * std::cos(0.5 * t), 2.0 * std::sin(0.5 * t), 0.3 * std::sin(0.9 * t)}; }
static ctba::Quat traj_q(double t) { const double a = 0.5 * t + 0.2 * std::sin(0.7 * t); return {0.0, 0.0, std::sin(a / 2), std::cos(a / 2)}; }   // yaw

template <int N>
int run() {
  const int64_t dt_ns = 100000000, start_ns = 1000000000;
  const int K = 60;
  ctba::BSplineSO3R3<N> spline(dt_ns, start_ns);
  std::vector<ctba::Quat> q(K);
  std::vector<ctba::Vec3> p(K);
  for (int i = 0; i < K; ++i) {
    // knots placed ON the trajectory (a B-spline does not interpolate its knots) plus an offset: a deliberately poor initial guess
    const double t = (start_ns + (i - (N - 2) / 2.0) * dt_ns) * 1e-9;
    p[i] = traj_p(t); p[i][0] += 0.05; p[i][2] -= 0.03;
    q[i] = traj_q(t + 0.05);
  }
  spline.initializeFromNumTraj(q, p);
  spline.addRotationLocalParametrization();
  int n_meas = 0;
  std::vector<int64_t> ts;
  for (int64_t t_ns = start_ns; t_ns < spline.getEndTns(); t_ns += 12345678) {
    const double t = t_ns * 1e-9;
    REQUIRE(spline.addPositionMeasurement(traj_p(t), t_ns, 1.0));
    ts.push_back(t_ns);
    ++n_meas;
  }
  for (int64_t t_ns : ts) REQUIRE(spline.addOrientationMeasurement(traj_q(t_ns * 1e-9), t_ns, 1.0));
  auto e0 = spline.evaluateErrors(CTBA_PRIOR);
  REQUIRE((int)e0.size() == 2 * n_meas);
  spline.loadOptimizationParams(20, 0, 8);
  ctba::Summary s = spline.optimize(20);
  REQUIRE(s.IsSolutionUsable());
  REQUIRE(s.iterations.size() >= 2 && s.num_successful_steps >= 1);
  REQUIRE(s.final_cost < 1e-3 * s.initial_cost);
  auto e1 = spline.evaluateErrors(CTBA_PRIOR);
  double m0 = 0, m1 = 0;
  for (int i = 0; i < n_meas; ++i) { m0 += e0[i] / n_meas; m1 += e1[i] / n_meas; }
  REQUIRE(m1 < 0.02 * m0);                                   // mean position error, evaluatePositionErrors-style
  const int64_t tq = start_ns + 2345678901LL;
  ctba::Pose T = spline.getPose(tq);
  ctba::Vec3 pt = traj_p(tq * 1e-9);
  REQUIRE(std::fabs(T.p[0] - pt[0]) < 2e-3 && std::fabs(T.p[1] - pt[1]) < 2e-3 && std::fabs(T.p[2] - pt[2]) < 2e-3);
  ctba::Quat qt = traj_q(tq * 1e-9);
  const double dot = std::fabs(T.q[0] * qt[0] + T.q[1] * qt[1] + T.q[2] * qt[2] + T.q[3] * qt[3]);
  REQUIRE(dot > 1.0 - 1e-5);
  ctba::Vec3 w = spline.getGyro(tq);                         // body rate of a pure yaw motion: about z
  REQUIRE(std::fabs(w[0]) < 1e-2 && std::fabs(w[1]) < 1e-2 && std::fabs(w[2] - (0.5 + 0.14 * std::cos(0.7 * tq * 1e-9))) < 2e-2);
  auto rows = spline.traceSpline(2000000);                   // 500 Hz
  REQUIRE(rows.size() == (size_t)((spline.getEndTns() - start_ns + 1999999) / 2000000));
  REQUIRE(std::fabs(rows[0][0] - start_ns * 1e-9) < 1e-12);
  bool threw = false;
  try { spline.getPose(start_ns - 1); } catch (const ctba::Error& e) { threw = e.code == CTBA_E_ARG; }
  REQUIRE(threw);
  std::printf("order %d: cost %.3e -> %.3e in %zu iterations, mean position error %.2e -> %.2e m\n", N, s.initial_cost, s.final_cost,
              s.iterations.size() - 1, m0, m1);
  return 0;
}

int main() {
  if (run<4>()) return 1;
  if (run<6>()) return 1;
  std::printf("OK\n");
  return 0;
}
