// ctba_bspline_so3r3.hpp — C++ host mirror of the reference's spline class over the C ABI (ctba.h).
//
//   gvi_fusion::BSplineSO3R3<N>        /root/reference/include/spline/bspline_so3r3.h:60-1990
//
// Same member names, argument meaning and call order as the reference class, so that a driver written against the reference
// (src/optimize_{,vi_,gv_,gi_}continuous_time.cpp, src/fit_spline_to_colmap.cpp) compiles against this header by replacing the Eigen /
// Sophus argument types with the plain value types below (Vec3 = 3 doubles, Quat = x y z w, Pose = translation + quaternion).  The
// object owns a ctba_problem handle instead of a ceres::Problem; add*Measurement appends to the handle, optimize() is ctba_solve.
// Error behaviour: where the reference aborts through GVI_FUSION_ASSERT_STREAM (util/assert.h:65-70) this class throws ctba::Error with
// the library's message; the skip rules of the add* members (measurement outside the spline) are the library's (= the reference's).
// Header-only, C++14, no dependencies besides ctba.h.  CUDA-only behind the ABI: without a usable device the constructor throws.
#pragma once
#include <array>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "ctba.h"

namespace ctba {

using Vec3 = std::array<double, 3>;
using Quat = std::array<double, 4>;                       // x y z w  (Sophus::SO3d::data(), so3.hpp:185-187)
struct Pose { Vec3 p; Quat q; };                          // Sophus::SE3d: translation(), unit_quaternion()

struct Error : std::runtime_error { int code; Error(int c, const std::string& m) : std::runtime_error(m), code(c) {} };

// ceres::Solver::Summary as far as the drivers read it (termination_type == ceres::CONVERGENCE, costs, step counts)
struct Summary {
  int termination_type = 1;                               // 0 CONVERGENCE, 1 NO_CONVERGENCE, 2 FAILURE
  double initial_cost = 0, final_cost = 0;
  int num_successful_steps = 0, num_unsuccessful_steps = 0;
  std::vector<ctba_iteration> iterations;
  double total_time_in_seconds = 0;
  bool IsSolutionUsable() const { return termination_type != 2; }
};

// ImuCalibration as far as the spline reads it (sensor/imu.h; noise densities come from the experiment YAML)
struct ImuNoise { double acc_noise_density = 1, gyro_noise_density = 1, acc_random_walk = 1, gyro_random_walk = 1; };

// one frame of Frame::points2D restricted to the valid reprojections (sensor/camera.h:114-184)
struct FrameObservations {
  int64_t t_ns = 0;
  std::vector<double> uv;          // 2 per point
  std::vector<int32_t> point3d;    // dense landmark index per point
};

template <int N>
class BSplineSO3R3 {
  static_assert(N == 4 || N == 6, "the reference instantiates orders 4 and 6");

 public:
  static constexpr int N_ = N;
  static constexpr double ns_to_s_ = 1e-9, s_to_ns_ = 1e9;

  // BSplineSO3R3(int64_t time_interval_ns, int64_t start_time_ns = 0)                 :75-117
  explicit BSplineSO3R3(int64_t time_interval_ns, int64_t start_time_ns = 0, int device = 0) : dt_ns_(time_interval_ns), start_t_ns_(start_time_ns) {
    const int rc = ctba_create(N, time_interval_ns, start_time_ns, device, &h_);
    if (rc != CTBA_OK) throw Error(rc, "ctba_create failed (no CUDA device or bad arguments); there is no CPU fallback");
  }
  ~BSplineSO3R3() { if (h_) ctba_destroy(h_); }
  BSplineSO3R3(const BSplineSO3R3&) = delete;
  BSplineSO3R3& operator=(const BSplineSO3R3&) = delete;
  ctba_problem* handle() { return h_; }

  // ---- state -------------------------------------------------------------------------------------------------------------
  // initializeFromNumTraj(rotations, positions)                                        :336-353
  void initializeFromNumTraj(const std::vector<Quat>& so3, const std::vector<Vec3>& r3) {
    if (so3.size() != r3.size()) throw Error(CTBA_E_ARG, "initializeFromNumTraj: sizes differ");
    chk(ctba_set_knots(h_, (int)so3.size(), so3.empty() ? nullptr : so3[0].data(), r3.empty() ? nullptr : r3[0].data()));
  }
  // initializeBiases(dt_s, init_acc_bias, init_gyro_bias)                              :355-369
  void initializeBiases(double dt_s, const Vec3& acc0, const Vec3& gyr0) { chk(ctba_initialize_biases(h_, dt_s, acc0.data(), gyr0.data())); }
  void addRotationLocalParametrization() {}                                            // :378-399: implied (every SO(3) knot is updated on the manifold)
  size_t numKnots() const { return (size_t)ctba_num_knots(h_); }                       // :1791
  size_t numBiasKnots() const { return (size_t)ctba_num_bias_knots(h_); }
  int64_t getStartTns() const { return start_t_ns_; }                                  // :1859
  int64_t getDeltaTns() const { return dt_ns_; }                                       // :1877
  int64_t minTimeNs() const { return start_t_ns_; }                                    // :1789
  int64_t maxTimeNs() const { return start_t_ns_ + ((int64_t)numKnots() - N + 1) * dt_ns_ - 1; }        // :1784-1787
  int64_t getEndTns() const { return (int64_t)((double)(((int64_t)numKnots() - N + 1) * dt_ns_ + start_t_ns_) - 1e6); }   // through double, like the reference   // :1862-1867
  Pose getKnot(int i) {                                                                // :215-219
    fetchKnots();
    if (i < 0 || i >= (int)kq_.size()) throw Error(CTBA_E_ARG, "getKnot: index out of range");
    return Pose{kp_[i], kq_[i]};
  }

  // set3DPoints / get3DPoints: landmark identity is the dense index                    :1907-1908
  void set3DPoints(const std::vector<Vec3>& p_wl) { n_lm_ = p_wl.size(); chk(ctba_set_landmarks(h_, (int)p_wl.size(), p_wl.empty() ? nullptr : p_wl[0].data())); }
  std::vector<Vec3> get3DPoints() {
    std::vector<Vec3> o(n_lm_);
    if (n_lm_) chk(ctba_get_landmarks(h_, o[0].data()));
    return o;
  }
  // setCameras (first camera): COLMAP model id, intrinsics fx fy cx cy, 4 distortion coefficients, T_b_c, cam-imu time offset   :1899-1903
  void setCamera(int model, const std::array<double, 4>& intr, const std::array<double, 4>& dist, const Pose& T_b_c, double t_offset_cam_imu_s) {
    chk(ctba_set_camera(h_, model, intr.data(), dist.data(), T_b_c.q.data(), T_b_c.p.data(), t_offset_cam_imu_s));
  }
  void setImu(const ImuNoise& n) { imu_ = n; }                                         // setImu(imu): :1905
  void setScale(double s) { scale_ = s; pushGlobals(); }                               // :1880
  void setPbp(const Vec3& p) { p_bp_ = p; pushGlobals(); }                             // :1841
  void setGravity(const Vec3& g) { g_ = g; pushGlobals(); }                            // :1792
  void setInitSplineGlobSensTimeOffset(double t_s) { t_s_gp_s_ = t_s; pushGlobals(); } // :1845
  void setOptimizeScale(bool b) { opt_[0] = b; pushFlags(); }                          // :1794-1897
  void setOptimizePbp(bool b) { opt_[1] = b; pushFlags(); }
  void setOptimizeGravity(bool b) { opt_[2] = b; pushFlags(); }
  void setOptimizeCamImuExtrinsics(bool b) { opt_[3] = b; pushFlags(); }
  void setOptimizeTimeOffsetSplineGlobalSensor(bool b) { opt_[4] = b; pushFlags(); }

  // ---- measurements (each call appends to the problem; skip rules as in the reference) --------------------------------------
  // addAccelMeasurement(meas, time_ns) + addGyroMeasurement(meas, time_ns)             :523-702
  // (the reference adds the two residual blocks of one IMU sample back to back, optimize_continuous_time.cpp:567-583)
  bool addImuMeasurement(const Vec3& lin_acc, const Vec3& ang_vel, int64_t time_ns) {
    return chk(ctba_add_imu(h_, 1, &time_ns, lin_acc.data(), ang_vel.data(), imu_.acc_noise_density, imu_.gyro_noise_density)) == 1;
  }
  int addImuMeasurements(const std::vector<int64_t>& t_ns, const std::vector<Vec3>& lin_acc, const std::vector<Vec3>& ang_vel) {
    if (t_ns.size() != lin_acc.size() || t_ns.size() != ang_vel.size()) throw Error(CTBA_E_ARG, "addImuMeasurements: sizes differ");
    if (t_ns.empty()) return 0;
    return chk(ctba_add_imu(h_, (int)t_ns.size(), t_ns.data(), lin_acc[0].data(), ang_vel[0].data(), imu_.acc_noise_density, imu_.gyro_noise_density));
  }
  // addGlobalPosMeasurement(meas, time_ns, noise)                                      :880-958
  bool addGlobalPosMeasurement(const Vec3& meas, int64_t time_ns, double noise) { return chk(ctba_add_gp(h_, 1, &time_ns, meas.data(), noise)) == 1; }
  // addImagePointsMeasurement(frame, noise): the frame's valid reprojections            :960-1046
  bool addImagePointsMeasurement(const FrameObservations& f, double noise) {
    const int32_t begin[2] = {0, (int32_t)f.point3d.size()};
    if (f.uv.size() != 2 * f.point3d.size()) throw Error(CTBA_E_ARG, "addImagePointsMeasurement: uv / point3d sizes differ");
    return chk(ctba_add_frames(h_, 1, &f.t_ns, begin, f.uv.data(), f.point3d.data(), noise)) == 1;
  }
  // addImuBiasCost(bias_err_freq_hz, bias_err_dt_s)                                    :706-792
  int addImuBiasCost(double bias_err_freq_hz, double bias_err_dt_s) {
    return chk(ctba_add_bias_cost(h_, bias_err_freq_hz, bias_err_dt_s, imu_.acc_random_walk, imu_.gyro_random_walk));
  }
  void addGravityNormError() { chk(ctba_add_gravity_norm(h_)); }                       // :1213-1227
  // addPositionMeasurement / addOrientationMeasurement (spline fitting)                 :1048-1123
  bool addPositionMeasurement(const Vec3& meas, int64_t time_ns, double noise) { return chk(ctba_add_position_measurements(h_, 1, &time_ns, meas.data(), noise)) == 1; }
  bool addOrientationMeasurement(const Quat& meas, int64_t time_ns, double noise) { return chk(ctba_add_orientation_measurements(h_, 1, &time_ns, meas.data(), noise)) == 1; }
  // ---- alignment of a frozen spline to the world frame (src/align_spline_to_global_frame.cpp:153-226) --------------------------
  void setWorldColmapTransformation(const Pose& T) {                                   // :1825
    const double t[7] = {T.q[0], T.q[1], T.q[2], T.q[3], T.p[0], T.p[1], T.p[2]};
    chk(ctba_set_world_alignment(h_, t, nullptr));
  }
  // addWorldColmapAlignmentMeasurement(meas, time_ns, noise)                           :1121-1211
  bool addWorldColmapAlignmentMeasurement(const Vec3& meas, int64_t time_ns, double noise) {
    return chk(ctba_add_alignment_measurements(h_, 1, &time_ns, meas.data(), noise)) == 1;
  }
  Pose getWorldColmapTransformation() {                                                // :1827-1834
    double t[7]; chk(ctba_get_world_alignment(h_, t, nullptr));
    return Pose{{t[4], t[5], t[6]}, {t[0], t[1], t[2], t[3]}};
  }
  Pose getWorldColmapTransformationCorr() {                                            // :1836-1840
    double c[7]; chk(ctba_get_world_alignment(h_, nullptr, c));
    return Pose{{c[4], c[5], c[6]}, {c[0], c[1], c[2], c[3]}};
  }
  std::vector<double> evaluateWorldColmapAlignmentErrors() { return evaluateErrors(CTBA_ALIGN); }   // :1682-1790
  double getScaleCorr() { return globals()[0]; }                                        // :1882
  double getSplineGlobSensTimeOffsetCorr() { return globals()[9]; }                     // :1852

  // ---- solve ----------------------------------------------------------------------------------------------------------------
  // loadOptimizationParams(n_max_outer_iters, n_max_inner_iters, n_threads)            :1760-1766 (threads: the GPU decides)
  void loadOptimizationParams(int n_max_outer_iters, int n_max_inner_iters, int /*n_threads*/) { n_outer_ = n_max_outer_iters; n_inner_ = n_max_inner_iters; }
  // ceres::Solver::Summary optimize(int max_num_iterations)                            :1768-1782
  Summary optimize(int max_num_iterations) {
    Summary s;
    s.iterations.resize((size_t)max_num_iterations + 2);
    ctba_summary cs{};
    cs.iterations = s.iterations.data(); cs.iterations_capacity = (int)s.iterations.size();
    chk(ctba_solve(h_, max_num_iterations, &cs));
    s.iterations.resize((size_t)cs.num_iterations);
    s.termination_type = cs.termination_type; s.initial_cost = cs.initial_cost; s.final_cost = cs.final_cost;
    s.num_successful_steps = cs.num_successful_steps; s.num_unsuccessful_steps = cs.num_unsuccessful_steps;
    s.total_time_in_seconds = cs.time_total_s;
    knots_fresh_ = false;
    return s;
  }
  void relinearizeSplineImuTimeOffset() { relinearize(1); }                             // :1807-1811
  void relinearizeSplineGlobSensTimeOffset() { relinearize(2); }                        // :1853-1857

  // ---- results --------------------------------------------------------------------------------------------------------------
  Pose getPose(int64_t time_ns) {                                                      // :221-264
    double o[7];
    chk(ctba_sample(h_, CTBA_SAMPLE_POSE, 1, &time_ns, o));
    return Pose{{o[0], o[1], o[2]}, {o[3], o[4], o[5], o[6]}};
  }
  Vec3 getAccel(int64_t time_ns, bool in_world_frame = false) { return sample3(in_world_frame ? CTBA_SAMPLE_ACCEL_WORLD : CTBA_SAMPLE_ACCEL_BODY, time_ns); }   // :122-174
  Vec3 getGyro(int64_t time_ns, bool in_world_frame = false) { return sample3(in_world_frame ? CTBA_SAMPLE_GYRO_WORLD : CTBA_SAMPLE_GYRO_BODY, time_ns); }      // :176-213
  Vec3 getAccelBias(int64_t time_ns) { return sample3(CTBA_SAMPLE_ACCEL_BIAS, time_ns); }   // :266-295
  Vec3 getGyroBias(int64_t time_ns) { return sample3(CTBA_SAMPLE_GYRO_BIAS, time_ns); }     // :297-327
  // traceSpline's loop (include/util/save.h:99-121) as one device call: rows t_s tx ty tz qx qy qz qw
  std::vector<std::array<double, 8>> traceSpline(int64_t dt_ns) {
    std::vector<int64_t> t;
    for (int64_t t_ns = getStartTns(); t_ns < getEndTns(); t_ns += dt_ns) t.push_back(t_ns);
    std::vector<double> o(7 * t.size());
    if (!t.empty()) chk(ctba_sample(h_, CTBA_SAMPLE_POSE, (int64_t)t.size(), t.data(), o.data()));
    std::vector<std::array<double, 8>> rows(t.size());
    for (size_t i = 0; i < t.size(); ++i) { rows[i][0] = t[i] * ns_to_s_; for (int c = 0; c < 7; ++c) rows[i][1 + c] = o[7 * i + c]; }
    return rows;
  }
  // evaluate{Reprojection,Accel,Gyro,GlobalPos,Position/Orientation}Errors               :1229-1680
  std::vector<double> evaluateErrors(int kind) {
    const int n = chk(ctba_evaluate_errors(h_, kind, nullptr));
    std::vector<double> e((size_t)n);
    if (n) chk(ctba_evaluate_errors(h_, kind, e.data()));
    return e;
  }
  double getScale() { return globals()[18] + globals()[0]; }                            // :1881
  Vec3 getPbp() { auto g = globals(); return {g[10], g[11], g[12]}; }                    // :1843
  Vec3 getGravity() { auto g = globals(); return {g[13], g[14], g[15]}; }                // :1793
  double getSplineImuTimeOffset() { return globals()[16]; }                             // :1805
  double getSplineGlobSensTimeOffset() { return globals()[17]; }                        // :1851
  std::array<int64_t, 8> counts() { std::array<int64_t, 8> c{}; chk(ctba_counts(h_, c.data())); return c; }

 private:
  int chk(int rc) const {
    if (rc < 0) throw Error(rc, std::string(ctba_last_error(h_)));
    return rc;
  }
  void relinearize(int mask) {
    chk(ctba_set_option(h_, "relinearize_mask", mask));
    chk(ctba_relinearize_time_offsets(h_));
    chk(ctba_set_option(h_, "relinearize_mask", 3));
  }
  void pushGlobals() { chk(ctba_set_globals(h_, scale_, p_bp_.data(), g_.data(), t_s_gp_s_)); }
  void pushFlags() { chk(ctba_set_flags(h_, opt_[0], opt_[1], opt_[2], opt_[3], opt_[4])); }
  std::array<double, 19> globals() { std::array<double, 19> o{}; chk(ctba_get_globals(h_, o.data())); return o; }
  Vec3 sample3(int what, int64_t t_ns) { Vec3 o{}; chk(ctba_sample(h_, what, 1, &t_ns, o.data())); return o; }
  void fetchKnots() {
    if (knots_fresh_) return;
    const size_t n = numKnots();
    kq_.resize(n); kp_.resize(n);
    if (n) chk(ctba_get_knots(h_, kq_[0].data(), kp_[0].data()));
    knots_fresh_ = true;
  }

  ctba_problem* h_ = nullptr;
  int64_t dt_ns_, start_t_ns_;
  ImuNoise imu_;
  double scale_ = 0.0, t_s_gp_s_ = 0.0;
  Vec3 p_bp_{{0, 0, 0}}, g_{{0, 0, 0}};
  bool opt_[5] = {false, true, false, true, true};        // ctor defaults :94-98: scale, p_bp, gravity, cam-imu extrinsics, spline/global-sensor offset
  int n_outer_ = 0, n_inner_ = 0;
  size_t n_lm_ = 0;
  std::vector<Quat> kq_; std::vector<Vec3> kp_;
  bool knots_fresh_ = false;
};

}  // namespace ctba
