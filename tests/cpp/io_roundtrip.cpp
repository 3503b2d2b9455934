// Reads the files tests/test_drivers.py wrote with drivers/ctba_io.hpp and prints what it parsed (one "key v v v" line each);
// then writes a pose / position / scalar back with the reference's precisions.
#include <cstdio>
#include "../../drivers/ctba_io.hpp"
using namespace ctba;
using namespace ctba::io;
int main(int argc, char** argv) {
  if (argc < 2) return 1;
  const std::string d = argv[1];
  std::vector<GlobalPosMeasurement> gp;
  if (!loadGpMeasurements(d + "/gp.txt", gp)) return 2;
  std::printf("gp_count %zu\n", gp.size());
  if (!gp.empty()) std::printf("gp_last %.17g %.17g %.17g %.17g\n", gp.back().t, gp.back().pos[0], gp.back().pos[1], gp.back().pos[2]);
  std::vector<Quat> kq; std::vector<Vec3> kp; std::vector<double> ts;
  if (!loadSplineFromFile(d + "/spline_knots.txt", kq, kp, ts)) return 3;
  std::printf("knot_count %zu\n", ts.size());
  std::printf("knot1 %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", ts[1], kp[1][0], kp[1][1], kp[1][2], kq[1][0], kq[1][1], kq[1][2], kq[1][3]);
  Pose T;
  if (!loadPose(d + "/T.txt", T)) return 4;
  std::printf("pose %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", T.q[0], T.q[1], T.q[2], T.q[3], T.p[0], T.p[1], T.p[2]);
  double sc = 0; Vec3 v{};
  if (!loadScalar<double>(d + "/scalar.txt", sc) || !loadVector3d(d + "/vec.txt", v)) return 5;
  std::printf("scalar %.17g\nvec %.17g %.17g %.17g\n", sc, v[0], v[1], v[2]);
  if (loadScalar<double>(d + "/does_not_exist.txt", sc)) return 6;
  Yaml y = Yaml::LoadFile(d + "/cfg.yaml");
  std::printf("yaml %s|%d|%.17g|%d|%d|%d\n", y.as<std::string>("gp_fn").c_str(), y.as<int>("spline_order"), y.as<double>("spline_control_nodes_dt_s"),
              (int)y.as<bool>("optimize_pBP"), (int)y.as<bool>("optimize_spline_globsens_time_offset"), (int)y.has("nope"));
  bool threw = false;
  try { y.as<int>("nope"); } catch (const std::exception&) { threw = true; }
  if (!threw) return 7;
  savePose(d, "T_out.txt", T); savePosition(d, "p_out.txt", v); saveScalar(d, "s_out.txt", sc);
  // linearInterpolation (trajectory_utils.h:96-143) of a trajectory file at the times of interp_ts.txt (one per line)
  std::vector<Quat> tq; std::vector<Vec3> tp; std::vector<double> tt;
  if (loadTrajectoryFromFile(d + "/traj.txt", tq, tp, tt)) {
    std::vector<double> its; std::ifstream f((d + "/interp_ts.txt").c_str()); double t;
    while (f >> t) its.push_back(t);
    std::vector<Quat> oq; std::vector<Vec3> op;
    linearInterpolation(tp, tq, tt, its, op, oq);
    for (size_t i = 0; i < its.size(); ++i)
      std::printf("interp%zu %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", i, op[i][0], op[i][1], op[i][2], oq[i][0], oq[i][1], oq[i][2], oq[i][3]);
    its.push_back(tt.back());            // a sample AT the last stamp is never produced: the reference's size assert fires
    bool thr = false;
    std::vector<Quat> oq2; std::vector<Vec3> op2;
    try { linearInterpolation(tp, tq, tt, its, op2, oq2); } catch (const std::runtime_error&) { thr = true; }
    if (!thr) return 8;
  }
  // calibration files: Kalibr camchain + imu.yaml (UZH-FPV) and the EuRoC layout
  {
    std::ifstream probe((d + "/camchain.yaml").c_str());
    if (probe.is_open()) {
      for (int cam = 0; cam < 2; ++cam) {
        SimpleCamImuCalib c;
        if (!loadUZHFPVSimpleCamImuCalib(d + "/camchain.yaml", d + "/imu.yaml", cam, c)) return 9;
        const Pose Tbc = poseInverse(c.T_cam_imu_);
        std::printf("kalibr%d %s|%s %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", cam,
                    c.camera_model_.c_str(), c.distortion_model_.c_str(), c.intrinsics_[0], c.intrinsics_[1], c.intrinsics_[2], c.intrinsics_[3], c.distortion_coeffs_[0],
                    c.distortion_coeffs_[1], c.distortion_coeffs_[2], c.distortion_coeffs_[3], c.resolution_[0], c.resolution_[1], c.timeshift_cam_imu_, c.imu_rate_,
                    c.acc_noise_density_, c.acc_random_walk_, c.gyro_noise_density_, c.gyro_random_walk_, Tbc.q[0], Tbc.q[1], Tbc.q[2], Tbc.q[3], Tbc.p[0], Tbc.p[1], Tbc.p[2]);
      }
      SimpleCamImuCalib bad;
      if (loadUZHFPVSimpleCamImuCalib(d + "/camchain.yaml", d + "/imu.yaml", 2, bad)) return 10;
      SimpleCamImuCalib e;
      if (!loadEuRoCSimpleCamImuCalib(d + "/euroc.yaml", 1, e)) return 11;
      const Pose Tbc = poseInverse(e.T_cam_imu_);
      std::printf("euroc1 %s|%s %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", e.camera_model_.c_str(), e.distortion_model_.c_str(),
                  e.intrinsics_[0], e.intrinsics_[3], e.distortion_coeffs_[0], e.distortion_coeffs_[3], e.timeshift_cam_imu_, e.imu_rate_, Tbc.q[0], Tbc.q[1], Tbc.q[2], Tbc.q[3],
                  Tbc.p[0], Tbc.p[1], Tbc.p[2]);
    }
  }
  std::printf("OK\n");
  return 0;
}
