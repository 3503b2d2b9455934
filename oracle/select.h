// ORACLE (test infrastructure only — never linked or called by the product): CPU restatement of
//   BSplineSO3R3::computeValidReprojections            /root/reference/include/spline/bspline_so3r3.h:406-521
//   Frame::setGrid / Frame::addPoint2D                 /root/reference/include/sensor/camera.h:155-184
// kept in the reference's own shape (Frame / Point2D objects, grids of Point2D pointers, std::find over points3D_ids) so that it
// is an independent statement of the same procedure as the flat-array product code (csrc/ctba_select.cu).
// Parity unpinned against the reference's own outputs: the reference has no test or golden vector for this function.
#pragma once
#include <algorithm>
#include <cstdint>
#include <vector>
#include "problem.h"
#include "dt_problem.h"

namespace orc_select {
using namespace oracle;

struct Point2D {                      // camera.h:96-111
  double u, v;
  uint64_t point3d_id;
  bool is_reproj_valid = false;
  double vel_u = 0.0, vel_v = 0.0;    // Point2D::velocity
};

struct Frame {                        // camera.h:114-184
  std::vector<Point2D> points2D;
  std::vector<uint64_t> points3D_ids;
  int64_t time_ns = 0;
  int cell_size_px = 0, n_grid_columns = 0, n_grid_rows = 0;
  std::vector<std::vector<size_t>> grid;
  int n_valid_reprojections = 0;
  void setGrid(int w, int h, int cell) {   // :172-184
    cell_size_px = cell;
    n_grid_columns = w / cell_size_px;
    if (w % cell_size_px > 0) n_grid_columns += 1;
    n_grid_rows = h / cell_size_px;
    if (h % cell_size_px > 0) n_grid_rows += 1;
    grid.resize(n_grid_rows * n_grid_columns);
  }
  void addPoint2D(double x, double y, uint64_t id3d) {   // :155-168
    Point2D pt; pt.u = x; pt.v = y; pt.point3d_id = id3d;
    points2D.push_back(pt);
    int idx_clmn = x / cell_size_px;
    int idx_row = y / cell_size_px;
    int idx = idx_clmn + idx_row * n_grid_columns;
    grid[idx].push_back(points2D.size() - 1);
    points3D_ids.push_back(id3d);
  }
};

inline void computeValidReprojections(SplineProblem& sp, std::vector<Frame>& frames) {
  if (frames.empty()) return;
  const int cell_size_px_ = frames[0].cell_size_px, n_grid_columns_ = frames[0].n_grid_columns, n_grid_rows_ = frames[0].n_grid_rows;   // :371-375
  const size_t n_cells = (size_t)n_grid_columns_ * n_grid_rows_;
  std::vector<std::vector<Point2D*>> grid_k_min_1(n_cells);
  for (auto frame = frames.begin(); frame != frames.end(); ++frame) {
    const int64_t st_ns = frame->time_ns - sp.start_t_ns;
    if (st_ns < 0) continue;
    const int64_t s = st_ns / sp.dt_ns;
    if (size_t(s + sp.N) > sp.numKnots()) continue;
    std::vector<std::vector<Point2D*>> grid_k(n_cells);
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      for (size_t point_id = 0; point_id < grid_k_min_1[cell_id].size(); point_id++) {
        const uint64_t point3D_id = grid_k_min_1[cell_id][point_id]->point3d_id;
        auto it = std::find(frame->points3D_ids.begin(), frame->points3D_ids.end(), point3D_id);
        if (it != frame->points3D_ids.end()) {
          const size_t point3D_idx = it - frame->points3D_ids.begin();
          Point2D* p = &frame->points2D[point3D_idx];
          p->is_reproj_valid = true;
          int idx_clmn = p->u / cell_size_px_;
          int idx_row = p->v / cell_size_px_;
          grid_k[idx_clmn + idx_row * n_grid_columns_].push_back(p);
        }
      }
    }
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      if (grid_k[cell_id].size() > 0) continue;
      if (frame->grid[cell_id].size() == 0) continue;
      Point2D* p = &frame->points2D[frame->grid[cell_id][0]];
      p->is_reproj_valid = true;
      grid_k[cell_id].push_back(p);
    }
    // getPose(time_ns) :221-270, then :477-483
    const double u = double(st_ns % sp.dt_ns) / double(sp.dt_ns);
    std::vector<const double*> kn;
    for (int j = 0; j < sp.N; ++j) kn.push_back(&sp.so3[4 * (s + j)]);
    for (int j = 0; j < sp.N; ++j) kn.push_back(&sp.r3[3 * (s + j)]);
    SO3<double> R_wb; V3<double> p_wb;
    if (sp.N == 4) {
      SplineHelper<4>::evaluate_lie<double, double>(kn.data(), u, sp.inv_dt, &R_wb, nullptr, nullptr, nullptr);
      SplineHelper<4>::evaluate<double, 0, double>(kn.data() + 4, u, sp.inv_dt, &p_wb);
    } else {
      SplineHelper<6>::evaluate_lie<double, double>(kn.data(), u, sp.inv_dt, &R_wb, nullptr, nullptr, nullptr);
      SplineHelper<6>::evaluate<double, 0, double>(kn.data() + 6, u, sp.inv_dt, &p_wb);
    }
    const SO3<double> R_bc{sp.cam.q_bc[0], sp.cam.q_bc[1], sp.cam.q_bc[2], sp.cam.q_bc[3]};
    const SO3<double> R_wc = R_wb * R_bc;
    const double scale = sp.scale + sp.scale_corr;
    const V3<double> t_bc{sp.cam.p_bc[0], sp.cam.p_bc[1], sp.cam.p_bc[2]};
    const V3<double> rb = R_wb.act(t_bc);
    const V3<double> p_wc{scale * rb.x + p_wb.x, scale * rb.y + p_wb.y, scale * rb.z + p_wb.z};
    const SO3<double> R_cw = R_wc.inverse();
    size_t n_reprojs = 0;
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      for (size_t point_id = 0; point_id < grid_k[cell_id].size(); point_id++) {
        const uint64_t l = grid_k[cell_id][point_id]->point3d_id;
        const V3<double> d{sp.lm[3 * l] - p_wc.x, sp.lm[3 * l + 1] - p_wc.y, sp.lm[3 * l + 2] - p_wc.z};
        const V3<double> p_CLi = R_cw.act(d);
        const double nrm = std::sqrt(p_CLi.x * p_CLi.x + p_CLi.y * p_CLi.y + p_CLi.z * p_CLi.z);
        if ((nrm < 0.3) || (p_CLi.z < 0.0)) grid_k[cell_id][point_id]->is_reproj_valid = false;
        else n_reprojs += 1;
      }
    }
    frame->n_valid_reprojections = (int)n_reprojs;
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) grid_k_min_1[cell_id].clear();
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++)
      for (size_t point_id = 0; point_id < grid_k[cell_id].size(); point_id++) grid_k_min_1[cell_id].push_back(grid_k[cell_id][point_id]);
  }
}

// DiscreteTimeEstimator::computeValidReprojections   /root/reference/include/discrete_time_estimator/estimator.h:179-291
// frame k is judged at state k's pose; no frame is skipped (the reference asserts that frame and state times agree)
inline void computeValidReprojectionsDt(DtProblem& dp, std::vector<Frame>& frames) {
  if (frames.empty()) return;
  const int cell_size_px_ = frames[0].cell_size_px, n_grid_columns_ = frames[0].n_grid_columns, n_grid_rows_ = frames[0].n_grid_rows;   // initializeGrid :167-172
  const size_t n_cells = (size_t)n_grid_columns_ * n_grid_rows_;
  std::vector<std::vector<Point2D*>> grid_k_min_1(n_cells);
  int k = 0;
  for (auto frame = frames.begin(); frame != frames.end(); ++frame) {
    std::vector<std::vector<Point2D*>> grid_k(n_cells);
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      for (size_t point_id = 0; point_id < grid_k_min_1[cell_id].size(); point_id++) {
        const uint64_t point3D_id = grid_k_min_1[cell_id][point_id]->point3d_id;
        auto it = std::find(frame->points3D_ids.begin(), frame->points3D_ids.end(), point3D_id);
        if (it != frame->points3D_ids.end()) {
          const size_t point3D_idx = it - frame->points3D_ids.begin();
          Point2D* p = &frame->points2D[point3D_idx];
          p->is_reproj_valid = true;
          int idx_clmn = p->u / cell_size_px_;
          int idx_row = p->v / cell_size_px_;
          grid_k[idx_clmn + idx_row * n_grid_columns_].push_back(p);
        }
      }
    }
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      if (grid_k[cell_id].size() > 0) continue;
      if (frame->grid[cell_id].size() == 0) continue;
      Point2D* p = &frame->points2D[frame->grid[cell_id][0]];
      p->is_reproj_valid = true;
      grid_k[cell_id].push_back(p);
    }
    // :243-250
    const double* T = &dp.pose[7 * (size_t)k];
    const SO3<double> R_wb{T[0], T[1], T[2], T[3]};
    const SO3<double> R_bc{dp.camera.cam.q_bc[0], dp.camera.cam.q_bc[1], dp.camera.cam.q_bc[2], dp.camera.cam.q_bc[3]};
    const SO3<double> R_wc = R_wb * R_bc;
    const double scale = dp.scale + dp.scale_corr;
    const V3<double> rb = R_wb.act(V3<double>{dp.camera.cam.p_bc[0], dp.camera.cam.p_bc[1], dp.camera.cam.p_bc[2]});
    const V3<double> p_wc{scale * rb.x + T[4], scale * rb.y + T[5], scale * rb.z + T[6]};
    const SO3<double> R_cw = R_wc.inverse();
    size_t n_reprojs = 0;
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) {
      for (size_t point_id = 0; point_id < grid_k[cell_id].size(); point_id++) {
        const uint64_t l = grid_k[cell_id][point_id]->point3d_id;
        const V3<double> d{dp.lm[3 * l] - p_wc.x, dp.lm[3 * l + 1] - p_wc.y, dp.lm[3 * l + 2] - p_wc.z};
        const V3<double> p_CLi = R_cw.act(d);
        const double nrm = std::sqrt(p_CLi.x * p_CLi.x + p_CLi.y * p_CLi.y + p_CLi.z * p_CLi.z);
        if ((nrm < 0.3) || (p_CLi.z < 0.0)) grid_k[cell_id][point_id]->is_reproj_valid = false;
        else n_reprojs += 1;
      }
    }
    frame->n_valid_reprojections = (int)n_reprojs;
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++) grid_k_min_1[cell_id].clear();
    for (size_t cell_id = 0; cell_id < n_cells; cell_id++)
      for (size_t point_id = 0; point_id < grid_k[cell_id].size(); point_id++) grid_k_min_1[cell_id].push_back(grid_k[cell_id][point_id]);
    k++;
  }
}
// DiscreteTimeEstimator::computeFeatureVelocity   estimator.h:293-325 (frame times in seconds)
inline void computeFeatureVelocity(std::vector<Frame>& frames, const std::vector<double>& ts) {
  for (size_t f = 0; f + 1 < frames.size(); ++f) {
    Frame& frame = frames[f]; Frame& next = frames[f + 1];
    const double dt = ts[f + 1] - ts[f];
    for (auto& point2D : frame.points2D) {
      if (!point2D.is_reproj_valid) continue;
      auto it = std::find(next.points3D_ids.begin(), next.points3D_ids.end(), point2D.point3d_id);
      if (it != next.points3D_ids.end()) {
        const Point2D& n2 = next.points2D[it - next.points3D_ids.begin()];
        point2D.vel_u = (n2.u - point2D.u) / dt; point2D.vel_v = (n2.v - point2D.v) / dt;
      }
    }
  }
}

}  // namespace orc_select
