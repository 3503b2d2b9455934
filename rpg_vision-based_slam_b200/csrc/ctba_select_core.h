// ctba_select_core.h — the grid-based choice of the reprojections that enter a problem, shared by the continuous-time and the
// discrete-time entry points (host only):
//   BSplineSO3R3::computeValidReprojections            include/spline/bspline_so3r3.h:406-521
//   DiscreteTimeEstimator::computeValidReprojections   include/discrete_time_estimator/estimator.h:179-291
// (+ the per-frame feature grid of Frame::addPoint2D / setGrid, include/sensor/camera.h:155-184).  The two differ only in where the
// frame's pose comes from (the spline at the frame time / the frame's state) and in the frames they skip; `pose(f, &R_wb, &p_wb)` returns
// false for a skipped frame.  Sequential from frame to frame (the grid of frame k-1 seeds frame k); flat arrays, one stamp array for the
// landmark -> point lookup instead of std::find.
#pragma once
#include <cmath>
#include <string>
#include <vector>
#include "../../include/ctba.h"
#include "ctba_math.h"

namespace ctba {

template <class PoseFn>
int select_valid_reprojections(int F, const int32_t* pt_begin, const double* uv, const int32_t* lm_index, int64_t L, const double* lm, int width, int height,
                               int cell_size_px, const m3& R_bc, const v3& p_bc, double scale, PoseFn&& pose, uint8_t* valid, int32_t* n_valid, std::string* err) {
  // Frame::setGrid  camera.h:172-184
  const int ncol = width / cell_size_px + (width % cell_size_px > 0 ? 1 : 0);
  const int nrow = height / cell_size_px + (height % cell_size_px > 0 ? 1 : 0);
  const int ncell = ncol * nrow;
  const int M = pt_begin[F];
  for (int i = 0; i < M; ++i) valid[i] = 0;             // Point2D ctor: is_reproj_valid(false)
  for (int f = 0; f < F; ++f) n_valid[f] = 0;
  auto cell_of = [&](int q) -> int {
    const int cx = (int)(uv[2 * (size_t)q] / cell_size_px), cy = (int)(uv[2 * (size_t)q + 1] / cell_size_px);   // :448-450, camera.h:163-165
    if (cx < 0 || cx >= ncol || cy < 0 || cy >= nrow || uv[2 * (size_t)q] < 0 || uv[2 * (size_t)q + 1] < 0) return -1;
    return cx + cy * ncol;
  };
  std::vector<int32_t> stamp((size_t)(L > 1 ? L : 1), -1), where((size_t)(L > 1 ? L : 1), 0);
  std::vector<int32_t> prev;                               // grid_{k-1}, cell-major
  std::vector<std::vector<int32_t>> grid(ncell);
  std::vector<int32_t> first(ncell);
  for (int f = 0; f < F; ++f) {
    m3 R_wb; v3 p_wb;
    if (!pose(f, &R_wb, &p_wb)) continue;
    for (auto& g : grid) g.clear();
    for (int c = 0; c < ncell; ++c) first[c] = -1;
    for (int q = pt_begin[f]; q < pt_begin[f + 1]; ++q) {
      const int l = lm_index[q];
      if (l < 0 || l >= L) { *err = "compute_valid_reprojections: Point 3D not found (bspline_so3r3.h:466)"; return CTBA_E_ARG; }
      const int c = cell_of(q);
      if (c < 0) { *err = "compute_valid_reprojections: feature outside the image grid"; return CTBA_E_ARG; }
      if (stamp[l] != f) { stamp[l] = f; where[l] = q; }   // std::find: first occurrence
      if (first[c] < 0) first[c] = q;                      // Frame::grid[c][0]
    }
    // track the 2-D points of the previous grid (:427-456)
    for (int32_t q0 : prev) {
      const int l = lm_index[q0];
      if (stamp[l] != f) continue;
      const int q = where[l];
      valid[q] = 1;
      grid[cell_of(q)].push_back(q);
    }
    // one new point per free cell (:459-473)
    for (int c = 0; c < ncell; ++c) {
      if (!grid[c].empty() || first[c] < 0) continue;
      valid[first[c]] = 1;
      grid[c].push_back(first[c]);
    }
    // drop points closer than 30 cm to the camera or behind it, at the CURRENT pose (:477-502)
    const m3 R_wc = mul(R_wb, R_bc);
    const v3 p_wc = scale * mul(R_wb, p_bc) + p_wb;
    const m3 R_cw = transpose(R_wc);
    int n = 0;
    for (int c = 0; c < ncell; ++c)
      for (int32_t q : grid[c]) {
        const int l = lm_index[q];
        const v3 pl{lm[3 * (size_t)l], lm[3 * (size_t)l + 1], lm[3 * (size_t)l + 2]};
        const v3 pc = mul(R_cw, pl - p_wc);
        const double nn = std::sqrt(pc.x * pc.x + pc.y * pc.y + pc.z * pc.z);
        if (nn < 0.3 || pc.z < 0.0) valid[q] = 0; else ++n;
      }
    n_valid[f] = n;
    prev.clear();
    for (int c = 0; c < ncell; ++c) for (int32_t q : grid[c]) prev.push_back(q);   // :505-517 (rejected points stay in the grid)
  }
  return CTBA_OK;
}

}  // namespace ctba
