// align_colmap_to_imu — the reference tool of the same name, a host-only step (no solver, nothing for the GPU to do):
//     /root/reference/src/align_colmap_to_imu.cpp:51-194
// The first IMU samples after the first COLMAP image give the gravity direction in the body frame; T_IB rotates the body z axis onto it
// (initializePoseFromImu, :51-82), and T_IG = T_IB (T_GC_0 T_cam_imu)^-1 takes the COLMAP frame to that gravity-aligned frame.
// Same YAML keys (cam_id, camimu_calib_fn, [imu_calib_fn], imu_fn, colmap_dir, result_dir, dataset), outputs T_IB.txt and T_IG.txt.
//   usage: align_colmap_to_imu <config.yaml>
#include <cstdio>
#include <cstdlib>
#include <sys/stat.h>
#include "ctba_io.hpp"

using namespace ctba;
using namespace ctba::io;

static void require(bool ok, const std::string& msg) {   // GVI_FUSION_ASSERT_STREAM (util/assert.h:65-70)
  if (!ok) { std::cerr << msg << "\n"; std::exit(2); }
}

// initializePoseFromImu (:51-82): T_WB = exp(-angle * (e_z x e_acc) / |e_z x e_acc|), zero translation
static Pose initializePoseFromImu(const std::vector<ImuMeasurement>& m) {
  Vec3 acc{{0, 0, 0}};
  for (const ImuMeasurement& s : m) for (int c = 0; c < 3; ++c) acc[c] += s.lin_acc[c];
  for (int c = 0; c < 3; ++c) acc[c] /= static_cast<double>(m.size());
  const double n = std::sqrt(acc[0] * acc[0] + acc[1] * acc[1] + acc[2] * acc[2]);
  const Vec3 e{{acc[0] / n, acc[1] / n, acc[2] / n}};
  Vec3 ax{{-e[1], e[0], 0.0}};                                                      // e_z x e_acc
  const double an = std::sqrt(ax[0] * ax[0] + ax[1] * ax[1]);                       // (the reference does not guard a vanishing cross product either)
  const double angle = std::acos(e[2]);
  for (int c = 0; c < 3; ++c) ax[c] = -ax[c] / an * angle;
  return Pose{{{0, 0, 0}}, so3Exp(ax)};
}

// Imu::getMeasurementsContainingEdges(t_k, t_k_p_1, out, false) (sensor/imu.h:86-135)
static bool measurementsContainingEdges(const std::vector<ImuMeasurement>& all, double t_k, double t_k_p_1, std::vector<ImuMeasurement>& out) {
  if (all.empty() || all.front().t > t_k || all.back().t < t_k_p_1) return false;
  size_t i0 = 0;
  for (; i0 < all.size(); ++i0) if (all[i0].t > t_k) { --i0; break; }
  size_t i1 = i0;
  for (; i1 < all.size(); ++i1) if (all[i1].t > t_k_p_1) break;
  if (i1 >= all.size()) return false;                                               // the reference dereferences end() here
  out.insert(out.begin(), all.begin() + (long)i0, all.begin() + (long)i1 + 1);
  return true;
}

int main(int argc, char** argv) {
  require(argc > 1, "Please, pass the .yaml filename");
  try {
    const Yaml config = Yaml::LoadFile(argv[1]);
    const int cam_id = config.as<int>("cam_id");
    const std::string imu_fn = config.as<std::string>("imu_fn"), colmap_dir = config.as<std::string>("colmap_dir"), trace_dir = config.as<std::string>("result_dir");
    mkdir(trace_dir.c_str(), 0777);
    const std::string dataset_fn = config.as<std::string>("dataset");
    std::cout << "Dataset: " << dataset_fn << "\nimu fn: " << imu_fn << "\ncolmap dir: " << colmap_dir << "\nresult dir: " << trace_dir << "\n\n";
    SimpleCamImuCalib calib;
    require(loadCalibForDataset(config, dataset_fn, cam_id, calib), "Camera-Imu calibration could not be loaded");
    std::vector<ImuMeasurement> full, imu;
    require(loadImuMeasurementsFromFile(imu_fn, full), "Could not load imu");
    std::cout << "Loaded " << full.size() << " imu measurements.\n\n";
    std::vector<Quat> q_gc; std::vector<Vec3> p_gc; std::vector<double> cam_ts;
    require(loadSplineFromFile(colmap_dir + "colmap_cam_estimates.txt", q_gc, p_gc, cam_ts) && !cam_ts.empty(), "Could not load COLMAP camera poses.");   // loadSE3TrajectoryFromFile
    const double ts_first_imu = cam_ts[0];
    for (const ImuMeasurement& m : full) if (m.t >= ts_first_imu) imu.push_back(m);
    require(imu.size() > 11, "fewer than 12 imu measurements after the first image");
    std::vector<ImuMeasurement> init_meas;
    require(measurementsContainingEdges(imu, imu[0].t, imu[10].t, init_meas), "no imu measurements around the first image");
    const Pose T_IB_0 = initializePoseFromImu(init_meas);
    const Pose T_GB_0 = poseMul(Pose{p_gc[0], q_gc[0]}, calib.T_cam_imu_);
    const Pose T_IG = poseMul(T_IB_0, poseInverse(T_GB_0));
    savePose(trace_dir, "T_IB.txt", T_IB_0);
    savePose(trace_dir, "T_IG.txt", T_IG);
  } catch (const std::exception& e) {
    std::cerr << e.what() << "\n";
    return 2;
  }
  return 0;
}
