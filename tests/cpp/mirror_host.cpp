// Host-side checks of the C++ mirror (include/ctba_bspline_so3r3.hpp) on a planning-only handle (device -1): problem assembly, the
// reference's skip rules (bspline_so3r3.h:525-563, 882-903, 962-973, 1052-1059) and the error behaviour; no GPU needed.
#include <cmath>
#include <cstdio>
#include "ctba_bspline_so3r3.hpp"

#define REQUIRE(c) do { if (!(c)) { std::printf("FAILED %s:%d  %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

int main() {
  using Spline = ctba::BSplineSO3R3<6>;
  const int64_t dt_ns = 100000000, start_ns = 5000000000;   // 0.1 s knots, spline starts at t = 5 s
  Spline spline(dt_ns, start_ns, /*device=*/-1);
  std::vector<ctba::Quat> q(30, ctba::Quat{{0, 0, 0, 1}});
  std::vector<ctba::Vec3> p(30);
  for (int i = 0; i < 30; ++i) p[i] = {0.1 * i, 0.0, 1.0};
  spline.initializeFromNumTraj(q, p);
  spline.addRotationLocalParametrization();
  REQUIRE(spline.numKnots() == 30);
  REQUIRE(spline.getStartTns() == start_ns && spline.getDeltaTns() == dt_ns);
  REQUIRE(spline.maxTimeNs() == start_ns + (30 - 6 + 1) * dt_ns - 1);
  REQUIRE(spline.getEndTns() == start_ns + (30 - 6 + 1) * dt_ns - 1000000);
  ctba::Pose k3 = spline.getKnot(3);
  REQUIRE(std::fabs(k3.p[0] - 0.3) < 1e-15 && k3.q[3] == 1.0);
  // global positions: before the start and beyond the last window are skipped, like the reference
  REQUIRE(!spline.addGlobalPosMeasurement({0, 0, 0}, start_ns - 1, 0.01));
  REQUIRE(spline.addGlobalPosMeasurement({0, 0, 0}, start_ns, 0.01));
  REQUIRE(spline.addGlobalPosMeasurement({0, 0, 0}, spline.maxTimeNs(), 0.01));
  REQUIRE(!spline.addGlobalPosMeasurement({0, 0, 0}, spline.maxTimeNs() + 1, 0.01));
  // IMU needs the bias spline
  spline.setImu({0.1, 0.01, 0.002, 4e-5});
  spline.initializeBiases(0.5, {0, 0, 0}, {0, 0, 0});
  REQUIRE(spline.numBiasKnots() == 6);                       // int(0.1 * 30 / 0.5)  (:355-369)
  REQUIRE(spline.addImuMeasurement({0, 0, 9.81}, {0, 0, 0}, start_ns + 1234567));
  REQUIRE(!spline.addImuMeasurement({0, 0, 9.81}, {0, 0, 0}, start_ns - 5));
  // pose priors
  REQUIRE(spline.addPositionMeasurement({1, 2, 3}, start_ns + 7 * dt_ns, 1.0));
  REQUIRE(spline.addOrientationMeasurement({0, 0, 0, 1}, start_ns + 7 * dt_ns, 1.0));
  REQUIRE(!spline.addPositionMeasurement({1, 2, 3}, start_ns + 26 * dt_ns, 1.0));
  // a frame with two observations
  spline.set3DPoints({{1, 0, 0}, {0, 1, 0}});
  spline.setCamera(4, {450, 450, 376, 240}, {0, 0, 0, 0}, ctba::Pose{{0, 0, 0}, {0, 0, 0, 1}}, 0.0);
  ctba::FrameObservations f; f.t_ns = start_ns + 3 * dt_ns + 17; f.uv = {100, 120, 300, 200}; f.point3d = {0, 1};
  REQUIRE(spline.addImagePointsMeasurement(f, 1.0));
  f.t_ns = start_ns - 1;
  REQUIRE(!spline.addImagePointsMeasurement(f, 1.0));
  auto c = spline.counts();   // frames, observations, accel, gyro, gp, bias blocks, gravity
  REQUIRE(c[0] == 1 && c[1] == 2 && c[2] == 1 && c[3] == 1 && c[4] == 2);
  REQUIRE(ctba_num_priors(spline.handle()) == 2);
  // world alignment (align_spline_to_global_frame): transformation composition and the skip rules, on a second handle
  {
    Spline al(dt_ns, start_ns, /*device=*/-1);
    al.initializeFromNumTraj(q, p);
    al.setWorldColmapTransformation(ctba::Pose{{1, 2, 3}, {0, 0, 0.6, 0.8}});
    al.setScale(2.0);
    al.setInitSplineGlobSensTimeOffset(0.05);
    REQUIRE(!al.addWorldColmapAlignmentMeasurement({0, 0, 0}, start_ns + 49999999, 1.0));   // before start + offset
    REQUIRE(al.addWorldColmapAlignmentMeasurement({0, 0, 0}, start_ns + 50000000, 1.0));
    REQUIRE(!al.addWorldColmapAlignmentMeasurement({0, 0, 0}, al.maxTimeNs() + 50000001, 1.0));
    REQUIRE(ctba_num_alignment_measurements(al.handle()) == 1);
    ctba::Pose T = al.getWorldColmapTransformation(), Tc = al.getWorldColmapTransformationCorr();
    REQUIRE(std::fabs(T.q[2] - 0.6) < 1e-15 && std::fabs(T.q[3] - 0.8) < 1e-15 && T.p[1] == 2.0);
    REQUIRE(Tc.q[3] == 1.0 && Tc.p[0] == 0.0);
    REQUIRE(al.getScaleCorr() == 0.0 && al.getSplineGlobSensTimeOffsetCorr() == 0.0);
    bool t2 = false;
    try { al.optimize(3); } catch (const ctba::Error& e) { t2 = e.code == CTBA_E_CUDA; }
    REQUIRE(t2);
  }
  // relinearizeSplineImuTimeOffset / relinearizeSplineGlobSensTimeOffset fold their own offset only (:1807-1811, :1853-1857)
  {
    Spline rl(dt_ns, start_ns, /*device=*/-1);
    rl.initializeFromNumTraj(q, p);
    rl.setInitSplineGlobSensTimeOffset(0.5);
    const double corr[10] = {0, 0, 0, 0, 1, 0, 0, 0, 0.01, 0.02};   // t_offset_cam_imu_corr = 0.01, t_s_gp_corr = 0.02
    REQUIRE(ctba_set_corrections(rl.handle(), corr) == CTBA_OK);
    rl.relinearizeSplineImuTimeOffset();
    REQUIRE(std::fabs(rl.getSplineImuTimeOffset() - 0.01) < 1e-15 && rl.getSplineGlobSensTimeOffset() == 0.5 && rl.getSplineGlobSensTimeOffsetCorr() == 0.02);
    rl.relinearizeSplineGlobSensTimeOffset();
    REQUIRE(std::fabs(rl.getSplineGlobSensTimeOffset() - 0.52) < 1e-15 && rl.getSplineGlobSensTimeOffsetCorr() == 0.0);
    REQUIRE(ctba_set_corrections(rl.handle(), corr) == CTBA_OK);
    REQUIRE(ctba_relinearize_time_offsets(rl.handle()) == CTBA_OK);   // the plain C call still folds both
    REQUIRE(std::fabs(rl.getSplineImuTimeOffset() - 0.02) < 1e-15 && std::fabs(rl.getSplineGlobSensTimeOffset() - 0.54) < 1e-15);
  }
  // no device: the compute entry points fail loudly (there is no CPU fallback)
  bool threw = false;
  try { spline.optimize(3); } catch (const ctba::Error& e) { threw = e.code == CTBA_E_CUDA; }
  REQUIRE(threw);
  threw = false;
  try { spline.getPose(start_ns + 5); } catch (const ctba::Error& e) { threw = e.code == CTBA_E_CUDA; }
  REQUIRE(threw);
  threw = false;
  try { ctba::BSplineSO3R3<4> bad(0, 0, -1); } catch (const ctba::Error& e) { threw = e.code == CTBA_E_ARG; }
  REQUIRE(threw);
  std::printf("OK\n");
  return 0;
}
