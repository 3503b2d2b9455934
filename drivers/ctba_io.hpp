// ctba_io.hpp — the on-disk formats either side of the solve, dependency-free C++14.
//
// Loaders restate the reference's parsing idioms exactly, including "read one record too many, then pop_back" (so a file must end with a
// newline to give the same count) and "a '#' is only recognised as the first character of a record":
//   loadGpMeasurements   /root/reference/include/sensor/global_position_sensor.h:43-74     t x y z
//   loadSplineFromFile   /root/reference/include/spline/spline_common.h:125-163            t tx ty tz qx qy qz qw
//   loadPose             /root/reference/include/util/load.h:18-49                         3x4 row-major R|t, R re-orthonormalised
//   loadVector3d / loadScalar                          load.h:52-101
//   fixRotationMatrix    /root/reference/include/util/utils.h:34-50                        U V^T of the SVD (here: polar iteration, same matrix)
// Writers: createTrace / tracePose / tracePosition / traceScalar / traceSpline   /root/reference/include/util/save.h:25-39, 99-117, 156-178
// YAML: the experiment files are flat `key: value` maps (experiments/**/**.yaml); Yaml reads exactly that subset.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <map>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "ctba_bspline_so3r3.hpp"

namespace ctba {
namespace io {

struct GlobalPosMeasurement { double t; Vec3 pos; };

inline void skipComment(std::ifstream& fs) {
  if (fs.peek() == '#') fs.ignore(std::numeric_limits<std::streamsize>::max(), '\n');
}

inline bool loadGpMeasurements(const std::string& filename, std::vector<GlobalPosMeasurement>& measurements) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open file: " << filename << "\n"; return false; }
  while (fs.good() && !fs.eof()) {
    skipComment(fs);
    double stamp = 0, x = 0, y = 0, z = 0;
    fs >> stamp >> x >> y >> z;
    measurements.push_back(GlobalPosMeasurement{stamp, Vec3{{x, y, z}}});
  }
  if (!measurements.empty()) measurements.pop_back();   // the loop reads one record past the last line (:69-71)
  return true;
}

// loadImuMeasurementsFromFile (sensor/imu.h:355-388): id t wx wy wz ax ay az
struct ImuMeasurement { double t; Vec3 lin_acc, ang_vel; };
inline bool loadImuMeasurementsFromFile(const std::string& filename, std::vector<ImuMeasurement>& measurements) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open imu file: " << filename << "\n"; return false; }
  while (fs.good() && !fs.eof()) {
    skipComment(fs);
    size_t imu_id = 0; double stamp = 0, wx = 0, wy = 0, wz = 0, ax = 0, ay = 0, az = 0;
    fs >> imu_id >> stamp >> wx >> wy >> wz >> ax >> ay >> az;
    measurements.push_back(ImuMeasurement{stamp, Vec3{{ax, ay, az}}, Vec3{{wx, wy, wz}}});
  }
  if (!measurements.empty()) measurements.pop_back();
  return true;
}

inline Quat normalized(const Quat& q) {   // Sophus::SO3d(Eigen::Quaterniond) normalises (so3.hpp:481-487)
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  return Quat{{q[0] / n, q[1] / n, q[2] / n, q[3] / n}};
}

inline bool loadSplineFromFile(const std::string& filename, std::vector<Quat>& so3_knots, std::vector<Vec3>& r3_knots, std::vector<double>& ts) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open file: " << filename << "\n"; return false; }
  while (fs.good() && !fs.eof()) {
    skipComment(fs);
    double stamp = 0, tx = 0, ty = 0, tz = 0, qw = 1, qx = 0, qy = 0, qz = 0;
    fs >> stamp >> tx >> ty >> tz >> qx >> qy >> qz >> qw;
    ts.push_back(stamp);
    so3_knots.push_back(normalized(Quat{{qx, qy, qz, qw}}));
    r3_knots.push_back(Vec3{{tx, ty, tz}});
  }
  if (!ts.empty()) { ts.pop_back(); so3_knots.pop_back(); r3_knots.pop_back(); }   // :154-157
  return true;
}

// Nearest rotation to a 3x3 matrix: R (R^T R)^(-1/2) = U V^T of the SVD, by Newton's polar iteration X <- (X + X^-T) / 2
// (quadratically convergent for the nearly orthogonal matrices of a 9-digit pose file).  det < 0 => the matrix is negated (utils.h:44-47).
inline void fixRotationMatrix(const double R[9], double R_new[9]) {
  double X[9];
  for (int i = 0; i < 9; ++i) X[i] = R[i];
  for (int it = 0; it < 50; ++it) {
    const double c00 = X[4] * X[8] - X[5] * X[7], c01 = X[5] * X[6] - X[3] * X[8], c02 = X[3] * X[7] - X[4] * X[6];
    const double det = X[0] * c00 + X[1] * c01 + X[2] * c02;
    const double C[9] = {c00, c01, c02,                                                    // cofactor matrix = det * X^-T
                         X[2] * X[7] - X[1] * X[8], X[0] * X[8] - X[2] * X[6], X[1] * X[6] - X[0] * X[7],
                         X[1] * X[5] - X[2] * X[4], X[2] * X[3] - X[0] * X[5], X[0] * X[4] - X[1] * X[3]};
    double diff = 0;
    for (int i = 0; i < 9; ++i) { const double y = 0.5 * (X[i] + C[i] / det); diff = std::fmax(diff, std::fabs(y - X[i])); X[i] = y; }
    if (diff < 1e-16) break;
  }
  const double det = X[0] * (X[4] * X[8] - X[5] * X[7]) + X[1] * (X[5] * X[6] - X[3] * X[8]) + X[2] * (X[3] * X[7] - X[4] * X[6]);
  for (int i = 0; i < 9; ++i) R_new[i] = det < 0 ? -X[i] : X[i];
}

// Eigen::Quaterniond(Matrix3d) (Shepperd's branches, as Eigen's QuaternionBase::operator=(MatrixBase)); x y z w
inline Quat quatFromMatrix(const double R[9]) {
  double q[4];   // w x y z
  const double t = R[0] + R[4] + R[8];
  if (t > 0.0) {
    double s = std::sqrt(t + 1.0);
    q[0] = 0.5 * s; s = 0.5 / s;
    q[1] = (R[7] - R[5]) * s; q[2] = (R[2] - R[6]) * s; q[3] = (R[3] - R[1]) * s;
  } else {
    int i = 0;
    if (R[4] > R[0]) i = 1;
    if (R[8] > R[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double s = std::sqrt(R[4 * i] - R[4 * j] - R[4 * k] + 1.0);
    q[1 + i] = 0.5 * s; s = 0.5 / s;
    q[0] = (R[3 * k + j] - R[3 * j + k]) * s;
    q[1 + j] = (R[3 * j + i] + R[3 * i + j]) * s;
    q[1 + k] = (R[3 * k + i] + R[3 * i + k]) * s;
  }
  return normalized(Quat{{q[1], q[2], q[3], q[0]}});
}

inline void matrixFromQuat(const Quat& q, double R[9]) {   // Eigen::Quaterniond::toRotationMatrix
  const double tx = 2 * q[0], ty = 2 * q[1], tz = 2 * q[2];
  const double twx = tx * q[3], twy = ty * q[3], twz = tz * q[3], txx = tx * q[0], txy = ty * q[0], txz = tz * q[0], tyy = ty * q[1], tyz = tz * q[1], tzz = tz * q[2];
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz; R[2] = txz + twy;
  R[3] = txy + twz; R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy; R[7] = tyz + twx; R[8] = 1 - (txx + tyy);
}

// ---- Sophus::SO3d group operations on x y z w quaternions, as the tools use them on the host (so3.hpp:229-231, 247-290, 325-339, 583-619) ----
inline Quat quatMul(const Quat& a, const Quat& b) {   // SO3 product (normalising, so3.hpp:325-339)
  return normalized(Quat{{a[3] * b[0] + a[0] * b[3] + a[1] * b[2] - a[2] * b[1], a[3] * b[1] + a[1] * b[3] + a[2] * b[0] - a[0] * b[2],
                          a[3] * b[2] + a[2] * b[3] + a[0] * b[1] - a[1] * b[0], a[3] * b[3] - a[0] * b[0] - a[1] * b[1] - a[2] * b[2]}});
}
inline Quat quatInverse(const Quat& q) { return Quat{{-q[0], -q[1], -q[2], q[3]}}; }
inline Quat so3Exp(const Vec3& om) {
  const double t2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2];
  double imag, real;
  if (t2 < 1e-20) { const double t4 = t2 * t2; imag = 0.5 - t2 / 48.0 + t4 / 3840.0; real = 1.0 - t2 / 8.0 + t4 / 384.0; }
  else { const double t = std::sqrt(t2), h = 0.5 * t; imag = std::sin(h) / t; real = std::cos(h); }
  return Quat{{imag * om[0], imag * om[1], imag * om[2], real}};
}
inline Vec3 so3Log(const Quat& q) {
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2], w = q[3];
  double f;
  if (n2 < 1e-20) f = 2.0 / w - (2.0 / 3.0) * n2 / (w * w * w);
  else { const double n = std::sqrt(n2); f = std::fabs(w) < 1e-10 ? (w > 0 ? M_PI : -M_PI) / n : 2.0 * std::atan(n / w) / n; }
  return Vec3{{f * q[0], f * q[1], f * q[2]}};
}
// Sophus::interpolate (interpolate.hpp:27-34): a * exp(p * log(a^-1 b))
inline Quat so3Interpolate(const Quat& a, const Quat& b, double p) {
  const Vec3 l = so3Log(quatMul(quatInverse(a), b));
  return quatMul(a, so3Exp(Vec3{{p * l[0], p * l[1], p * l[2]}}));
}

// loadTrajectoryFromFile (util/trajectory_utils.h:17-55): the same 8 columns and idiom as loadSplineFromFile
inline bool loadTrajectoryFromFile(const std::string& filename, std::vector<Quat>& ori, std::vector<Vec3>& pos, std::vector<double>& ts) {
  return loadSplineFromFile(filename, ori, pos, ts);
}

// linearInterpolation (util/trajectory_utils.h:96-143), including its stop rule: samples at or after ts.back() are never produced and the
// size check then fails (the reference aborts; here: std::runtime_error)
inline void linearInterpolation(const std::vector<Vec3>& pos, const std::vector<Quat>& ori, const std::vector<double>& ts,
                                const std::vector<double>& interpolated_ts, std::vector<Vec3>& interpolated_pos, std::vector<Quat>& interpolated_ori) {
  size_t it_i = 0, t_i = 0;
  const size_t n = interpolated_ts.size(), m = ts.size() - 1;
  while (n > 0 && ts.size() >= 2) {
    if (interpolated_ts[it_i] >= ts[t_i] && interpolated_ts[it_i] < ts[t_i + 1]) {
      const double dt = (interpolated_ts[it_i] - ts[t_i]) / (ts[t_i + 1] - ts[t_i]);
      const Vec3 &p1 = pos[t_i], &p2 = pos[t_i + 1];
      interpolated_pos.push_back(Vec3{{p1[0] + dt * (p2[0] - p1[0]), p1[1] + dt * (p2[1] - p1[1]), p1[2] + dt * (p2[2] - p1[2])}});
      interpolated_ori.push_back(so3Interpolate(ori[t_i], ori[t_i + 1], dt));
      if (++it_i == n) break;
    } else {
      if (++t_i == m) break;
    }
  }
  if (interpolated_pos.size() != n)
    throw std::runtime_error("interpolated_pos.size() (= " + std::to_string(interpolated_pos.size()) + ") != interpolated_ts.size() (= " + std::to_string(n) + ")");
}

// ---- rigid transforms on Pose (Sophus::SE3d as far as the drivers use it) ----
inline Vec3 rotate(const Quat& q, const Vec3& v) { double R[9]; matrixFromQuat(q, R); return Vec3{{R[0] * v[0] + R[1] * v[1] + R[2] * v[2], R[3] * v[0] + R[4] * v[1] + R[5] * v[2], R[6] * v[0] + R[7] * v[1] + R[8] * v[2]}}; }
inline Pose poseMul(const Pose& A, const Pose& B) { const Vec3 t = rotate(A.q, B.p); return Pose{{{t[0] + A.p[0], t[1] + A.p[1], t[2] + A.p[2]}}, quatMul(A.q, B.q)}; }
inline Pose poseInverse(const Pose& T) { const Quat qi = quatInverse(T.q); const Vec3 t = rotate(qi, T.p); return Pose{{-t[0], -t[1], -t[2]}, qi}; }

// loadFrameTimes (util/colmap_utils.h:14-42): first column of colmap_cam_estimates.txt
inline bool loadFrameTimes(const std::string& filename, std::vector<double>& timestamps) {
  std::vector<Quat> q; std::vector<Vec3> p;
  return loadSplineFromFile(filename, q, p, timestamps);
}

// SimpleCamImuCalib + loadEuRoCSimpleCamImuCalib (util/load_calibration.h:20-60, 166-201): the EuRoC-format calibration YAML
// (datasets/EuRoC/calib/*/calib.yaml: cameras: [- camera: {...}, T_B_C: {...}], imu_params: {...}), read with a reader for that layout.
// The UZH-FPV (Kalibr camchain) variant, loadUZHFPVSimpleCamImuCalib, is not built.
struct SimpleCamImuCalib {
  int idx_ = 0;
  std::string camera_model_, distortion_model_;
  double resolution_[2] = {0, 0}, distortion_coeffs_[4] = {0, 0, 0, 0}, intrinsics_[4] = {0, 0, 0, 0};
  Pose T_cam_imu_{{{0, 0, 0}}, {{0, 0, 0, 1}}};
  double timeshift_cam_imu_ = 0, imu_rate_ = 0, acc_noise_density_ = 0, acc_random_walk_ = 0, gyro_noise_density_ = 0, gyro_random_walk_ = 0;
};
namespace detail {
inline std::string valueAfter(const std::string& seg, const std::string& key, size_t from = 0) {   // "key: value" -> value (to end of line / comment)
  const size_t k = seg.find(key, from);
  if (k == std::string::npos) throw std::runtime_error("calibration yaml: missing '" + key + "'");
  size_t e = seg.find('\n', k);
  std::string v = seg.substr(k + key.size(), (e == std::string::npos ? seg.size() : e) - k - key.size());
  const size_t h = v.find('#'); if (h != std::string::npos) v = v.substr(0, h);
  const size_t a = v.find_first_not_of(" \t\r\"'"), b = v.find_last_not_of(" \t\r\"'");
  return a == std::string::npos ? std::string() : v.substr(a, b - a + 1);
}
inline std::vector<double> dataAfter(const std::string& seg, const std::string& section) {       // section: ... data: [a, b, ...] (may span lines)
  const size_t k = seg.find(section);
  if (k == std::string::npos) throw std::runtime_error("calibration yaml: missing '" + section + "'");
  const size_t d = seg.find("data:", k), lb = seg.find('[', d), rb = seg.find(']', lb);
  if (d == std::string::npos || lb == std::string::npos || rb == std::string::npos) throw std::runtime_error("calibration yaml: no data list after '" + section + "'");
  std::string body = seg.substr(lb + 1, rb - lb - 1);
  for (char& c : body) if (c == ',') c = ' ';
  std::istringstream ss(body); std::vector<double> out; double v;
  while (ss >> v) out.push_back(v);
  return out;
}
}  // namespace detail
inline bool loadEuRoCSimpleCamImuCalib(const std::string& calib_fn, int cam_id, SimpleCamImuCalib& calib) {
  std::ifstream fs(calib_fn.c_str());
  if (!fs.is_open()) throw std::runtime_error("Could not open yaml file: " + calib_fn);   // YAML::LoadFile throws BadFile
  std::vector<std::string> lines; std::string line;
  while (std::getline(fs, line)) lines.push_back(line);
  auto top_level = [](const std::string& l) { return !l.empty() && l[0] != ' ' && l[0] != '\t' && l[0] != '#' && l[0] != '-'; };
  std::vector<size_t> cams;
  size_t imu_line = lines.size();
  for (size_t i = 0; i < lines.size(); ++i) {
    const size_t a = lines[i].find_first_not_of(" \t");
    if (a != std::string::npos && lines[i].compare(a, 9, "- camera:") == 0) cams.push_back(i);
    if (lines[i].compare(0, 11, "imu_params:") == 0) imu_line = i;
  }
  if (cam_id < 0 || (size_t)cam_id >= cams.size() || imu_line == lines.size()) throw std::runtime_error("calibration yaml: camera " + std::to_string(cam_id) + " or imu_params not found");
  auto segment = [&](size_t b, bool stop_at_camera) {
    std::string seg;
    for (size_t i = b; i < lines.size(); ++i) {
      if (i > b && (top_level(lines[i]) || (stop_at_camera && std::find(cams.begin(), cams.end(), i) != cams.end()))) break;
      seg += lines[i]; seg += '\n';
    }
    return seg;
  };
  const std::string cam = segment(cams[(size_t)cam_id], true), imu = segment(imu_line, false);
  calib.idx_ = cam_id;
  calib.camera_model_ = detail::valueAfter(cam, " type:");
  calib.resolution_[0] = std::stod(detail::valueAfter(cam, "image_width:")); calib.resolution_[1] = std::stod(detail::valueAfter(cam, "image_height:"));
  calib.distortion_model_ = detail::valueAfter(cam, " type:", cam.find("distortion:"));
  const std::vector<double> dc = detail::dataAfter(cam, "distortion:"), in = detail::dataAfter(cam, "intrinsics:"), T = detail::dataAfter(cam, "T_B_C:");
  if (dc.size() < 4 || in.size() < 4 || T.size() < 16) throw std::runtime_error("calibration yaml: short data list");
  for (int i = 0; i < 4; ++i) { calib.distortion_coeffs_[i] = dc[i]; calib.intrinsics_[i] = in[i]; }
  const double R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};           // the list is row-major (Matrix4d(data).transpose(), :189-190)
  Pose T_imu_cam{{{T[3], T[7], T[11]}}, quatFromMatrix(R)};
  calib.T_cam_imu_ = poseInverse(T_imu_cam);
  calib.timeshift_cam_imu_ = -1.0 * std::stod(detail::valueAfter(imu, "delay_imu_cam:"));
  calib.imu_rate_ = std::stod(detail::valueAfter(imu, "imu_rate:"));
  calib.acc_noise_density_ = std::stod(detail::valueAfter(imu, "sigma_acc_c:")); calib.acc_random_walk_ = std::stod(detail::valueAfter(imu, "sigma_acc_bias_c:"));
  calib.gyro_noise_density_ = std::stod(detail::valueAfter(imu, "sigma_omega_c:")); calib.gyro_random_walk_ = std::stod(detail::valueAfter(imu, "sigma_omega_bias_c:"));
  return true;
}

inline bool loadPose(const std::string& filename, Pose& pose) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open file: " << filename << "\n"; return false; }
  skipComment(fs);
  double R[9], t[3];
  fs >> R[0] >> R[1] >> R[2] >> t[0] >> R[3] >> R[4] >> R[5] >> t[1] >> R[6] >> R[7] >> R[8] >> t[2];
  double Rn[9];
  fixRotationMatrix(R, Rn);
  pose.q = quatFromMatrix(Rn);
  pose.p = Vec3{{t[0], t[1], t[2]}};
  return true;
}

inline bool loadVector3d(const std::string& filename, Vec3& v) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open file: " << filename << "\n"; return false; }
  skipComment(fs);
  fs >> v[0] >> v[1] >> v[2];
  return true;
}

template <typename T>
bool loadScalar(const std::string& filename, T& scalar) {
  std::ifstream fs(filename.c_str());
  if (!fs.is_open()) { std::cerr << "Could not open file: " << filename << "\n"; return false; }
  skipComment(fs);
  T s{};
  fs >> s;
  scalar = s;
  return true;
}

// ---- writers (save.h) ----
inline void createTrace(const std::string& trace_dir, const std::string& trace_fn, std::ofstream* fs) {
  const std::string full_fn = trace_dir + "/" + trace_fn;
  fs->open(full_fn.c_str());
  if (fs->fail()) throw std::runtime_error("Fail to create trace: " + full_fn);   // the reference aborts (save.h:30-34)
  std::cout << "Created trace: " << full_fn << "\n";
}
inline void traceScalar(double v, std::ofstream& s) { s << v << std::endl; }
inline void tracePose(const Pose& pose, std::ofstream& s) {
  double R[9];
  matrixFromQuat(pose.q, R);
  s << R[0] << " " << R[1] << " " << R[2] << " " << pose.p[0] << "\n"
    << R[3] << " " << R[4] << " " << R[5] << " " << pose.p[1] << "\n"
    << R[6] << " " << R[7] << " " << R[8] << " " << pose.p[2] << "\n"
    << 0.0 << " " << 0.0 << " " << 0.0 << " " << 1.0 << std::endl;
}
inline void tracePosition(const Vec3& pos, std::ofstream& s) {
  s << "# x y z" << std::endl;
  s << pos[0] << " " << pos[1] << " " << pos[2] << std::endl;
}
// traceSpline (save.h:99-117): one device call for all samples instead of a getPose per line
template <class Spline>
void traceSpline(Spline& spline, int64_t dt_ns, std::ofstream& s) {
  s << "# timestamp tx ty tz qx qy qz qw" << std::endl;
  for (const auto& r : spline.traceSpline(dt_ns))
    s << std::setprecision(15) << r[0] << " " << r[1] << " " << r[2] << " " << r[3] << " " << r[4] << " " << r[5] << " " << r[6] << " " << r[7] << std::endl;
}
// the tools' saveSpline / savePose / savePosition / saveScalar (src/align_spline_to_global_frame.cpp:40-85, src/fit_spline_to_colmap.cpp)
template <class Spline>
void saveSpline(const std::string& trace_dir, Spline& spline) {
  std::ofstream f; createTrace(trace_dir, "spline.txt", &f); f.precision(9);
  traceSpline(spline, static_cast<int64_t>((1.0 / 100.0) * 1e9), f);
}
inline void savePose(const std::string& trace_dir, const std::string& fn, const Pose& pose) { std::ofstream f; createTrace(trace_dir, fn, &f); f.precision(9); tracePose(pose, f); }
inline void savePosition(const std::string& trace_dir, const std::string& fn, const Vec3& p) { std::ofstream f; createTrace(trace_dir, fn, &f); f.precision(9); tracePosition(p, f); }
inline void saveScalar(const std::string& trace_dir, const std::string& fn, double v) { std::ofstream f; createTrace(trace_dir, fn, &f); f.precision(16); traceScalar(v, f); }
// traceSplineKnots (save.h:81-96) / saveSplineKnots of fit_spline_to_colmap.cpp:35-44: knot i with the caller's time stamp ts[i], 15 digits
template <class Spline>
void saveSplineKnots(const std::string& trace_dir, Spline& spline, const std::vector<double>& ts) {
  std::ofstream f; createTrace(trace_dir, "spline_knots.txt", &f); f.precision(15);
  f << "# timestamp tx ty tz qx qy qz qw" << std::endl;
  for (size_t i = 0; i < spline.numKnots(); ++i) {
    const Pose k = spline.getKnot((int)i);
    f << std::setprecision(15) << ts.at(i) << " " << k.p[0] << " " << k.p[1] << " " << k.p[2] << " " << k.q[0] << " " << k.q[1] << " " << k.q[2] << " " << k.q[3] << std::endl;
  }
}
// spline_knots.txt as loadSplineFromFile reads it (the knot time stamps are t0 + i dt)
template <class Spline>
void saveSplineKnots(const std::string& trace_dir, const std::string& fn, Spline& spline) {
  std::ofstream f; createTrace(trace_dir, fn, &f);
  f << "# timestamp tx ty tz qx qy qz qw" << std::endl;
  for (size_t i = 0; i < spline.numKnots(); ++i) {
    const Pose k = spline.getKnot((int)i);
    f << std::setprecision(17) << (spline.getStartTns() + (int64_t)i * spline.getDeltaTns()) * 1e-9 << " " << k.p[0] << " " << k.p[1] << " " << k.p[2] << " "
      << k.q[0] << " " << k.q[1] << " " << k.q[2] << " " << k.q[3] << std::endl;
  }
}

// ---- flat YAML (`key: value`, '#' comments, optional quotes) — what YAML::LoadFile + config["k"].as<T>() is used for in the drivers ----
class Yaml {
 public:
  static Yaml LoadFile(const std::string& fn) {
    std::ifstream fs(fn.c_str());
    if (!fs.is_open()) throw std::runtime_error("Could not open yaml file: " + fn);
    Yaml y;
    std::string line, open_key;
    while (std::getline(fs, line)) {
      bool in_q = false; char qc = 0; size_t cut = line.size();
      for (size_t i = 0; i < line.size(); ++i) {
        const char c = line[i];
        if (in_q) { if (c == qc) in_q = false; }
        else if (c == '"' || c == '\'') { in_q = true; qc = c; }
        else if (c == '#' && (i == 0 || line[i - 1] == ' ' || line[i - 1] == '\t')) { cut = i; break; }
      }
      line = line.substr(0, cut);
      if (!open_key.empty()) {                       // continuation of a flow sequence `key: [a, b,` ... `c]`
        y.kv_[open_key] += " " + trim(line);
        if (line.find(']') != std::string::npos) open_key.clear();
        continue;
      }
      const size_t colon = line.find(':');
      if (colon == std::string::npos) continue;
      std::string key = trim(line.substr(0, colon)), val = trim(line.substr(colon + 1));
      if (val.size() >= 2 && (val.front() == '"' || val.front() == '\'') && val.back() == val.front()) val = val.substr(1, val.size() - 2);
      if (!key.empty()) y.kv_[key] = val;
      if (!key.empty() && val.find('[') != std::string::npos && val.find(']') == std::string::npos) open_key = key;
    }
    return y;
  }
  bool has(const std::string& k) const { return kv_.count(k) != 0; }
  // config["k"].as<std::vector<double>>() for a one-line flow sequence `k: [a, b, c]`
  std::vector<double> asList(const std::string& k) const {
    std::string v = as<std::string>(k);
    for (char& c : v) if (c == '[' || c == ']' || c == ',') c = ' ';
    std::istringstream ss(v); std::vector<double> out; double d;
    while (ss >> d) out.push_back(d);
    return out;
  }
  template <class T> T as(const std::string& k) const {
    auto it = kv_.find(k);
    if (it == kv_.end()) throw std::runtime_error("yaml: missing key '" + k + "'");   // yaml-cpp throws InvalidNode / BadConversion
    return conv(it->second, k, static_cast<T*>(nullptr));
  }
 private:
  static std::string trim(const std::string& s) {
    const size_t a = s.find_first_not_of(" \t\r\n"), b = s.find_last_not_of(" \t\r\n");
    return a == std::string::npos ? std::string() : s.substr(a, b - a + 1);
  }
  static std::string conv(const std::string& v, const std::string&, std::string*) { return v; }
  static double conv(const std::string& v, const std::string& k, double*) {
    std::istringstream ss(v); double d; if (!(ss >> d)) throw std::runtime_error("yaml: '" + k + "' is not a number"); return d;
  }
  static int conv(const std::string& v, const std::string& k, int*) {
    std::istringstream ss(v); int d; if (!(ss >> d)) throw std::runtime_error("yaml: '" + k + "' is not an integer"); return d;
  }
  static bool conv(const std::string& v, const std::string& k, bool*) {   // yaml-cpp: true/false, yes/no, on/off, y/n in any of three capitalisations
    std::string l; for (char c : v) l += (char)std::tolower((unsigned char)c);
    if (l == "true" || l == "yes" || l == "on" || l == "y") return true;
    if (l == "false" || l == "no" || l == "off" || l == "n") return false;
    throw std::runtime_error("yaml: '" + k + "' is not a boolean");
  }
  std::map<std::string, std::string> kv_;
};

// loadUZHFPVSimpleCamImuCalib (util/load_calibration.h:83-160): Kalibr camchain (cam0 / cam1 maps with flow sequences that may start on
// the line after the key) + Kalibr imu.yaml
namespace detail {
inline std::vector<double> listAfterKey(const std::string& seg, const std::string& key) {
  const size_t k = seg.find(key);
  if (k == std::string::npos) throw std::runtime_error("calibration yaml: missing '" + key + "'");
  const size_t lb = seg.find('[', k), rb = seg.find(']', lb);
  if (lb == std::string::npos || rb == std::string::npos) throw std::runtime_error("calibration yaml: no list after '" + key + "'");
  std::string body = seg.substr(lb + 1, rb - lb - 1);
  for (char& c : body) if (c == ',') c = ' ';
  std::istringstream ss(body); std::vector<double> out; double v;
  while (ss >> v) out.push_back(v);
  return out;
}
}  // namespace detail
inline bool loadUZHFPVSimpleCamImuCalib(const std::string& camimu_calib_fn, const std::string& imu_calib_fn, int cam_id, SimpleCamImuCalib& calib) {
  if (cam_id != 0 && cam_id != 1) return false;
  std::ifstream fs(camimu_calib_fn.c_str());
  if (!fs.is_open()) throw std::runtime_error("Could not open yaml file: " + camimu_calib_fn);
  const std::string want = cam_id == 0 ? "cam0:" : "cam1:";
  std::string line, cam; bool in = false;
  while (std::getline(fs, line)) {
    const bool top = !line.empty() && line[0] != ' ' && line[0] != '\t' && line[0] != '#' && line[0] != '[';
    if (top) in = line.compare(0, want.size(), want) == 0;
    if (in) { cam += line; cam += '\n'; }
  }
  if (cam.empty()) throw std::runtime_error("calibration yaml: '" + want + "' not found");
  calib.idx_ = cam_id;
  calib.camera_model_ = detail::valueAfter(cam, "camera_model:");
  calib.distortion_model_ = detail::valueAfter(cam, "distortion_model:");
  const std::vector<double> dc = detail::listAfterKey(cam, "distortion_coeffs:"), res = detail::listAfterKey(cam, "resolution:"),
                            in4 = detail::listAfterKey(cam, "intrinsics:"), T = detail::listAfterKey(cam, "T_cam_imu:");
  if (dc.size() < 4 || res.size() < 2 || in4.size() < 4 || T.size() < 16) throw std::runtime_error("calibration yaml: short list");
  for (int i = 0; i < 4; ++i) { calib.distortion_coeffs_[i] = dc[i]; calib.intrinsics_[i] = in4[i]; }
  calib.resolution_[0] = res[0]; calib.resolution_[1] = res[1];
  calib.timeshift_cam_imu_ = std::stod(detail::valueAfter(cam, "timeshift_cam_imu:"));
  const double R[9] = {T[0], T[1], T[2], T[4], T[5], T[6], T[8], T[9], T[10]};           // row-major list (Matrix4d(data).transpose(), :115-117)
  calib.T_cam_imu_ = Pose{{{T[3], T[7], T[11]}}, quatFromMatrix(R)};
  const Yaml imu = Yaml::LoadFile(imu_calib_fn);
  calib.imu_rate_ = imu.as<double>("update_rate");
  calib.acc_noise_density_ = imu.as<double>("accelerometer_noise_density"); calib.acc_random_walk_ = imu.as<double>("accelerometer_random_walk");
  calib.gyro_noise_density_ = imu.as<double>("gyroscope_noise_density"); calib.gyro_random_walk_ = imu.as<double>("gyroscope_random_walk");
  return true;
}
// the drivers' dataset switch (src/optimize_continuous_time.cpp:288-311, 551-565): calibration format and the key of the GP noise
inline bool loadCalibForDataset(const Yaml& config, const std::string& dataset, int cam_id, SimpleCamImuCalib& calib) {
  if (dataset == "UZH_FPV") return loadUZHFPVSimpleCamImuCalib(config.as<std::string>("camimu_calib_fn"), config.as<std::string>("imu_calib_fn"), cam_id, calib);
  if (dataset == "EuRoC" || dataset == "Custom") return loadEuRoCSimpleCamImuCalib(config.as<std::string>("camimu_calib_fn"), cam_id, calib);
  throw std::runtime_error("Unknown dataset");
}
inline double gpStdForDataset(const Yaml& config, const std::string& dataset) {
  if (dataset == "UZH_FPV") return config.as<double>("leica_std");
  if (dataset == "EuRoC" || dataset == "Custom") return config.as<double>("gp_std");
  throw std::runtime_error("Unknown dataset");
}

}  // namespace io
}  // namespace ctba
