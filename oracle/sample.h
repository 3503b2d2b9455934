// ORACLE (test infrastructure only — never linked or called by the product): CPU restatement of the spline sampling getters
//   BSplineSO3R3::getAccel       /root/reference/include/spline/bspline_so3r3.h:122-174
//   BSplineSO3R3::getGyro        :176-213
//   BSplineSO3R3::getPose        :221-264
//   BSplineSO3R3::getAccelBias   :266-295        (NOTE: evaluates the bias knots with the ORDER-N_ helper and s + N_ knots, not the cubic one)
//   BSplineSO3R3::getGyroBias    :297-327
// which traceSpline / traceAccBiasSpline / traceGyroBiasSpline (include/util/save.h:99-153) and the drivers' predicted-IMU dumps
// (src/optimize_continuous_time.cpp:88-130) call in a loop over timestamps.
// Parity unpinned against the reference's own outputs (no test or golden vector for these getters in the reference).
#pragma once
#include <cstdint>
#include <vector>
#include "problem.h"

namespace orc_getters {
using namespace oracle;

enum { POSE = 0, ACCEL_BODY = 1, ACCEL_WORLD = 2, GYRO_BODY = 3, GYRO_WORLD = 4, ACCEL_BIAS = 5, GYRO_BIAS = 6 };

template <int N>
int sample_one(const SplineProblem& sp, int what, int64_t time_ns, double* out) {
  const int64_t st_ns = time_ns - sp.start_t_ns;
  if (st_ns < 0) return -1;                                   // GVI_FUSION_ASSERT_STREAM(st_ns >= 0)
  if (what == ACCEL_BIAS || what == GYRO_BIAS) {
    const std::vector<double>& b = what == ACCEL_BIAS ? sp.acc_bias : sp.gyr_bias;
    const int64_t s = st_ns / sp.dt_ns_bias;
    const double u = double(st_ns % sp.dt_ns_bias) / double(sp.dt_ns_bias);
    if (size_t(s + N) > b.size() / 3) return -1;
    std::vector<const double*> vec;
    for (int i = 0; i < N; ++i) vec.push_back(&b[3 * (s + i)]);
    V3<double> bias;
    SplineHelper<N>::template evaluate<double, 0, double>(vec.data(), u, sp.inv_dt_bias, &bias);
    out[0] = bias.x; out[1] = bias.y; out[2] = bias.z;
    return 0;
  }
  const int64_t s = st_ns / sp.dt_ns;
  const double u = double(st_ns % sp.dt_ns) / double(sp.dt_ns);
  if (size_t(s + N) > sp.numKnots()) return -1;
  std::vector<const double*> vq, vp;
  for (int i = 0; i < N; ++i) { vq.push_back(&sp.so3[4 * (s + i)]); vp.push_back(&sp.r3[3 * (s + i)]); }
  SO3<double> rot;
  if (what == POSE) {
    V3<double> trans;
    SplineHelper<N>::template evaluate_lie<double, double>(vq.data(), u, sp.inv_dt, &rot, nullptr, nullptr, nullptr);
    SplineHelper<N>::template evaluate<double, 0, double>(vp.data(), u, sp.inv_dt, &trans);
    out[0] = trans.x; out[1] = trans.y; out[2] = trans.z; out[3] = rot.x; out[4] = rot.y; out[5] = rot.z; out[6] = rot.w;   // save.h:112-117
    return 0;
  }
  if (what == ACCEL_BODY || what == ACCEL_WORLD) {
    V3<double> a;
    SplineHelper<N>::template evaluate_lie<double, double>(vq.data(), u, sp.inv_dt, &rot, nullptr, nullptr, nullptr);
    SplineHelper<N>::template evaluate<double, 2, double>(vp.data(), u, sp.inv_dt, &a);
    V3<double> ag{a.x + sp.g[0], a.y + sp.g[1], a.z + sp.g[2]};
    if (what == ACCEL_BODY) ag = rot.inverse().act(ag);
    out[0] = ag.x; out[1] = ag.y; out[2] = ag.z;
    return 0;
  }
  V3<double> gyro;
  SplineHelper<N>::template evaluate_lie<double, double>(vq.data(), u, sp.inv_dt, &rot, &gyro, nullptr, nullptr);
  if (what == GYRO_WORLD) gyro = rot.inverse().act(gyro);     // sic: the reference applies the INVERSE rotation here (:203-207)
  out[0] = gyro.x; out[1] = gyro.y; out[2] = gyro.z;
  return 0;
}

inline int sample(const SplineProblem& sp, int what, int64_t n, const int64_t* t_ns, double* out) {
  const int w = what == POSE ? 7 : 3;
  for (int64_t i = 0; i < n; ++i) {
    int rc;
    if (sp.N == 4) rc = sample_one<4>(sp, what, t_ns[i], out + w * i);
    else if (sp.N == 5) rc = sample_one<5>(sp, what, t_ns[i], out + w * i);
    else rc = sample_one<6>(sp, what, t_ns[i], out + w * i);
    if (rc) return rc;
  }
  return 0;
}
}  // namespace orc_getters
