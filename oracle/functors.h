// ORACLE — TEST INFRASTRUCTURE ONLY (never shipped, never on the product path).
//
// The six hot cost functors of the reference, restated line for line as templates on the
// scalar type so that one source yields values (double) and autodiff Jacobians (Jet<K>),
// exactly like the reference's ceres::DynamicAutoDiffCostFunction use:
//   ReprojectionCostFunctor<N>          include/optimization/continuous_time/reprojection_cost_functor.h:47-109
//   AccelerometerCostFunctor<N>         include/optimization/continuous_time/imu_cost_functor.h:48-77
//   GyroCostFunctor<N,SO3,false>        include/optimization/continuous_time/imu_cost_functor.h:106-139
//   BiasCostFunctor<4>                  include/optimization/continuous_time/imu_cost_functor.h:351-369
//   GlobalPositionSensorCostFunctor<N>  include/optimization/continuous_time/global_position_cost_functor.h:45-73
//   GravityNormCostFunctor              include/optimization/gravity_norm_cost_functor.h:26-39
//   GlobalPositionSensorWithAlignmentCorrectionCostFunctor<N>   global_position_cost_functor.h:82-150
// Parameter-block order inside each functor is the order wired by bspline_so3r3.h
// (:596-610, :685-694, :928-939, :1007-1028, :771-788, :1222-1224).
#pragma once
#include <vector>
#include "ref_math.h"

namespace oracle {

template <int N>
struct ReprojectionCostFunctor {
  const CameraModel* camera;
  double scale, u, inv_dt, inv_std;
  int n_reproj;
  const double* uv;  // 2*n_reproj measured pixels (the valid points of the frame, in frame order)

  template <class T>
  bool operator()(T const* const* sKnots, T* sResiduals) const {
    SO3<T> R_w_b;
    SplineHelper<N>::template evaluate_lie<T, double>(sKnots, u, inv_dt, &R_w_b);
    V3<T> p_w_b;
    SplineHelper<N>::template evaluate<T, 0, double>(sKnots + N, u, inv_dt, &p_w_b);

    T scale_corr = sKnots[2 * N][0];
    T s = scale + scale_corr;

    SO3<T> R_b_c_corr = SO3<T>::fromRaw(sKnots[2 * N + 1]);
    V3<T> p_b_c_corr{sKnots[2 * N + 2][0], sKnots[2 * N + 2][1], sKnots[2 * N + 2][2]};

    // camera->T_b_c.rotationMatrix() : double rotation of the calibration
    SO3<double> q_bc0{camera->q_bc[0], camera->q_bc[1], camera->q_bc[2], camera->q_bc[3]};
    M3<double> Rbc0d = q_bc0.matrix();
    M3<T> Rbc0;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Rbc0.m[i][j] = T(Rbc0d.m[i][j]);
    V3<T> pbc0{T(camera->p_bc[0]), T(camera->p_bc[1]), T(camera->p_bc[2])};

    M3<T> R_b_c = mul(Rbc0, R_b_c_corr.matrix());
    V3<T> p_b_c = mul(Rbc0, p_b_c_corr) + pbc0;

    M3<T> Rwb = R_w_b.matrix();
    M3<T> R_w_c = mul(Rwb, R_b_c);
    // s * R_w_b.matrix() * p_b_c : (s*R)*p
    M3<T> sR;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) sR.m[i][j] = s * Rwb.m[i][j];
    V3<T> p_w_c = mul(sR, p_b_c) + p_w_b;

    M3<T> R_c_w = transpose(R_w_c);
    // -1.0 * R_c_w * p_w_c
    M3<T> mR;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) mR.m[i][j] = -1.0 * R_c_w.m[i][j];
    V3<T> p_c_w = mul(mR, p_w_c);

    for (int n = 0; n < n_reproj; ++n) {
      V3<T> p_w_li{sKnots[2 * N + 3 + n][0], sKnots[2 * N + 3 + n][1], sKnots[2 * N + 3 + n][2]};
      V3<T> p_c_li = mul(R_c_w, p_w_li) + p_c_w;
      T reproj_uv[2];
      project(*camera, p_c_li, reproj_uv);
      T thr = T(100.0);
      T e0 = reproj_uv[0] - uv[2 * n + 0];
      T e1 = reproj_uv[1] - uv[2 * n + 1];
      T sq = e0 * e0 + e1 * e1;
      if (sq > thr) {
        sResiduals[2 * n + 0] = T(0.0);
        sResiduals[2 * n + 1] = T(0.0);
      } else {
        sResiduals[2 * n + 0] = inv_std * e0;
        sResiduals[2 * n + 1] = inv_std * e1;
      }
    }
    return true;
  }
};

template <int N>
struct AccelerometerCostFunctor {
  static constexpr int Nb = 4;
  double meas[3];
  double st_s, k, dt_s, inv_dt, u_bias, inv_dt_bias, inv_std;

  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    T t_corr_s = sKnots[3 * N + 1][0];
    T st_corr_s = st_s - t_corr_s;
    T u = (st_corr_s - k * dt_s) / dt_s;
    SO3<T> R_w_b;
    SplineHelper<N>::template evaluate_lie<T, T>(sKnots, u, inv_dt, &R_w_b);
    V3<T> accel_w;
    SplineHelper<N>::template evaluate<T, 2, T>(sKnots + N, u, inv_dt, &accel_w);
    V3<T> bias;
    SplineHelper<Nb>::template evaluate<T, 0, double>(sKnots + 2 * N, u_bias, inv_dt_bias, &bias);
    V3<T> g{sKnots[3 * N][0], sKnots[3 * N][1], sKnots[3 * N][2]};
    V3<T> a_b = R_w_b.inverse().act(accel_w + g);
    r[0] = inv_std * (a_b.x - meas[0] + bias.x);
    r[1] = inv_std * (a_b.y - meas[1] + bias.y);
    r[2] = inv_std * (a_b.z - meas[2] + bias.z);
    return true;
  }
};

template <int N>
struct GyroCostFunctor {
  static constexpr int Nb = 4;
  double meas[3];
  double st_s, k, dt_s, inv_dt, u_bias, inv_dt_bias, inv_std;

  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    T t_corr_s = sKnots[2 * N][0];
    T st_corr_s = st_s - t_corr_s;
    T u = (st_corr_s - k * dt_s) / dt_s;
    V3<T> rot_vel;
    SplineHelper<N>::template evaluate_lie<T, T>(sKnots, u, inv_dt, (SO3<T>*)nullptr, &rot_vel);
    V3<T> bias;
    SplineHelper<Nb>::template evaluate<T, 0, double>(sKnots + N, u_bias, inv_dt_bias, &bias);
    r[0] = inv_std * (rot_vel.x - meas[0] + bias.x);
    r[1] = inv_std * (rot_vel.y - meas[1] + bias.y);
    r[2] = inv_std * (rot_vel.z - meas[2] + bias.z);
    return true;
  }
};

struct BiasCostFunctor {
  static constexpr int N = 4;
  double u_0, u_1, inv_dt, inv_std;
  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    V3<T> b0, b1;
    SplineHelper<N>::template evaluate<T, 0, double>(sKnots, u_0, inv_dt, &b0);
    SplineHelper<N>::template evaluate<T, 0, double>(sKnots + N, u_1, inv_dt, &b1);
    r[0] = inv_std * (b0.x - b1.x);
    r[1] = inv_std * (b0.y - b1.y);
    r[2] = inv_std * (b0.z - b1.z);
    return true;
  }
};

template <int N>
struct GlobalPositionSensorCostFunctor {
  double meas[3];
  double scale, st_s, k, dt_s, inv_dt, inv_std;
  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    T t_corr_s = sKnots[2 * N + 1][0];
    T st_corr_s = st_s - t_corr_s;
    T u = (st_corr_s - k * dt_s) / dt_s;
    SO3<T> R_w_b;
    SplineHelper<N>::template evaluate_lie<T, T>(sKnots, u, inv_dt, &R_w_b);
    V3<T> p_w_b;
    SplineHelper<N>::template evaluate<T, 0, T>(sKnots + N, u, inv_dt, &p_w_b);
    V3<T> p_b_p{sKnots[2 * N][0], sKnots[2 * N][1], sKnots[2 * N][2]};
    T scale_corr = sKnots[2 * N + 2][0];
    T s = scale + scale_corr;
    M3<T> Rwb = R_w_b.matrix();
    M3<T> sR;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) sR.m[i][j] = s * Rwb.m[i][j];
    V3<T> p_w_p = mul(sR, p_b_p) + p_w_b;
    r[0] = inv_std * (p_w_p.x - meas[0]);
    r[1] = inv_std * (p_w_p.y - meas[1]);
    r[2] = inv_std * (p_w_p.z - meas[2]);
    return true;
  }
};

// global_position_cost_functor.h:82-150 — parameter blocks: N rotation knots, N position knots, t_corr, scale_corr, R_w_g_corr, p_w_g_corr, p_b_p
// (bspline_so3r3.h:1151-1188).  Eigen evaluates `s * R_w_g * p_g_b` and `s * R_w_b * p_b_p` left to right: the matrix is scaled first.
template <int N>
struct GlobalPositionSensorWithAlignmentCorrectionCostFunctor {
  double meas[3];
  double q_wg[4], p_wg0[3];   // T_w_g
  double scale, st_s, k, dt_s, inv_dt, inv_std;
  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    T t_corr_s = sKnots[2 * N][0];
    T st_corr_s = st_s - t_corr_s;
    T u = (st_corr_s - k * dt_s) / dt_s;
    SO3<T> R_g_b;
    SplineHelper<N>::template evaluate_lie<T, T>(sKnots, u, inv_dt, &R_g_b);
    V3<T> p_g_b;
    SplineHelper<N>::template evaluate<T, 0, T>(sKnots + N, u, inv_dt, &p_g_b);
    T scale_corr = sKnots[2 * N + 1][0];
    T s = scale + scale_corr;
    SO3<T> R_w_g_corr = SO3<T>::fromRaw(sKnots[2 * N + 2]);
    V3<T> p_w_g_corr{sKnots[2 * N + 3][0], sKnots[2 * N + 3][1], sKnots[2 * N + 3][2]};
    V3<T> p_b_p{sKnots[2 * N + 4][0], sKnots[2 * N + 4][1], sKnots[2 * N + 4][2]};
    SO3<double> q0{q_wg[0], q_wg[1], q_wg[2], q_wg[3]};
    M3<double> R0d = q0.matrix();
    M3<T> R0;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) R0.m[i][j] = T(R0d.m[i][j]);
    M3<T> R_w_g = mul(R0, R_w_g_corr.matrix());
    V3<T> p_w_g{T(p_wg0[0]) + p_w_g_corr.x, T(p_wg0[1]) + p_w_g_corr.y, T(p_wg0[2]) + p_w_g_corr.z};
    M3<T> R_w_b = mul(R_w_g, R_g_b.matrix());
    M3<T> sRwg, sRwb;
    for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { sRwg.m[i][j] = s * R_w_g.m[i][j]; sRwb.m[i][j] = s * R_w_b.m[i][j]; }
    V3<T> p_w_b = mul(sRwg, p_g_b) + p_w_g;
    V3<T> p_w_p = mul(sRwb, p_b_p) + p_w_b;
    r[0] = inv_std * (p_w_p.x - meas[0]);
    r[1] = inv_std * (p_w_p.y - meas[1]);
    r[2] = inv_std * (p_w_p.z - meas[2]);
    return true;
  }
};

// position_cost_functor.h:24-57
template <int N>
struct PositionCostFunctor {
  double meas[3];
  double u, inv_dt, inv_std;
  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    V3<T> p;
    SplineHelper<N>::template evaluate<T, 0, double>(sKnots, u, inv_dt, &p);
    r[0] = inv_std * (p.x - meas[0]);
    r[1] = inv_std * (p.y - meas[1]);
    r[2] = inv_std * (p.z - meas[2]);
    return true;
  }
};

// orientation_cost_functor.h:24-61:  dr = r * measurement.inverse();  residuals = inv_std * 2 dr.unit_quaternion().vec()
template <int N>
struct OrientationCostFunctor {
  double meas[4];   // x y z w
  double u, inv_dt, inv_std;
  template <class T>
  bool operator()(T const* const* sKnots, T* r) const {
    SO3<T> rot;
    SplineHelper<N>::template evaluate_lie<T, double>(sKnots, u, inv_dt, &rot);
    SO3<T> mi = SO3<T>{T(meas[0]), T(meas[1]), T(meas[2]), T(meas[3])}.inverse();
    SO3<T> dr = rot * mi;
    r[0] = inv_std * (2.0 * dr.x);
    r[1] = inv_std * (2.0 * dr.y);
    r[2] = inv_std * (2.0 * dr.z);
    return true;
  }
};

struct GravityNormCostFunctor {
  double g_norm = 9.80665;
  double weight = 1e6;
  template <class T>
  bool operator()(T const* const* parameters, T* r) const {
    const T* g = parameters[0];
    T n2 = g[0] * g[0] + g[1] * g[1] + g[2] * g[2];
    r[0] = weight * (n2 - (g_norm * g_norm));
    return true;
  }
};

// ---------------------------------------------------------------------------------
// ceres::DynamicAutoDiffCostFunction<Functor, 4>::Evaluate restated [Ceres, recalled]:
// the functor is re-run ceil(#parameters / Stride) times with Jet<double,Stride>, each pass
// seeding Stride consecutive global parameter scalars.  `jac` (may be null) receives the
// GLOBAL-size Jacobian, row-major num_residuals x total_params.
template <class Functor, int Stride = 4>
inline void dynamic_autodiff(const Functor& f, const std::vector<const double*>& blocks,
                             const std::vector<int>& sizes, int num_residuals, double* residuals,
                             double* jac) {
  const int nb = (int)blocks.size();
  int total = 0;
  std::vector<int> start(nb);
  for (int b = 0; b < nb; ++b) { start[b] = total; total += sizes[b]; }
  if (!jac) {
    std::vector<const double*> ptrs(blocks);
    f(ptrs.data(), residuals);
    return;
  }
  using J = Jet<Stride>;
  std::vector<J> storage(total);
  std::vector<const J*> jp(nb);
  for (int b = 0; b < nb; ++b) {
    for (int i = 0; i < sizes[b]; ++i) storage[start[b] + i] = J(blocks[b][i]);
    jp[b] = storage.data() + start[b];
  }
  std::vector<J> out(num_residuals);
  for (int p0 = 0; p0 < total; p0 += Stride) {
    const int np = std::min(Stride, total - p0);
    for (int k = 0; k < np; ++k) storage[p0 + k].v[k] = 1.0;
    f(jp.data(), out.data());
    for (int k = 0; k < np; ++k) storage[p0 + k].v[k] = 0.0;
    for (int r = 0; r < num_residuals; ++r) {
      for (int k = 0; k < np; ++k) jac[(size_t)r * total + p0 + k] = out[r].v[k];
      if (p0 == 0) residuals[r] = out[r].a;
    }
  }
}

}  // namespace oracle
