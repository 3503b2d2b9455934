// ctba_colmap.hpp — COLMAP binary ingest of the drivers, restated from
//     readCamerasBinary                /root/reference/include/sensor/camera.h:245-270
//     readImagesAndPoints3DBinary      /root/reference/include/sensor/camera.h:273-441
//     numberParameters                 /root/reference/include/sensor/camera_calib.h:17-47
//     readBinaryLittleEndian           /root/reference/include/util/endian.h:51-69
// including the reference's rules: image id - 1, quaternion stored qw qx qy qz and normalised, 2-D points without a 3-D point skipped
// (kInvalidPoint3DId), exact-duplicate 2-D points skipped (same x AND y as one earlier point, :351-360), a 3-D point kept once per image
// (:363-369), and the camera record quirk (fx fy cx cy, then numberParameters(model) MORE doubles: for model 4 that is four past the
// record — only the first four are used).  Output is flat (CSR) so that it feeds ctba_compute_valid_reprojections and
// ctba_add_frames directly; landmark identity becomes the index of the point in points3D.bin order.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <fstream>
#include <limits>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <vector>

namespace ctba {
namespace io {

inline size_t numberParameters(int model_id) {
  static const int n[11] = {3, 4, 4, 5, 8, 4, 12, 5, 4, 5, 12};
  if (model_id < 0 || model_id > 10) throw std::runtime_error("Camera model id " + std::to_string(model_id) + " not supported!");
  return (size_t)n[model_id];
}

template <typename T>
T readBinaryLittleEndian(std::istream* s) {   // little-endian hosts only (x86-64 / aarch64), like every machine the reference ran on
  T v{};
  s->read(reinterpret_cast<char*>(&v), sizeof(T));
  return v;
}

struct ColmapCamera {
  uint32_t cam_id = 0; int model_id = 0; uint64_t w = 0, h = 0;
  double fx = 0, fy = 0, cx = 0, cy = 0;
  std::vector<double> dist_coeffs;   // numberParameters(model) entries; [0..3] are the distortion coefficients the models use
};

inline void readCamerasBinary(const std::string& path, std::vector<ColmapCamera>& cameras) {
  std::ifstream file(path, std::ios::binary);
  if (!file.is_open()) throw std::runtime_error("cannot open " + path);
  const size_t num_cameras = readBinaryLittleEndian<uint64_t>(&file);
  for (size_t i = 0; i < num_cameras; i++) {
    ColmapCamera c;
    c.cam_id = readBinaryLittleEndian<uint32_t>(&file);
    c.model_id = readBinaryLittleEndian<int>(&file);
    c.dist_coeffs.assign(numberParameters(c.model_id), 0.0);
    if (c.model_id != 4 && c.model_id != 5 && c.model_id != 7) throw std::runtime_error("Camera model id " + std::to_string(c.model_id) + " not supported!");
    c.w = readBinaryLittleEndian<uint64_t>(&file); c.h = readBinaryLittleEndian<uint64_t>(&file);
    c.fx = readBinaryLittleEndian<double>(&file); c.fy = readBinaryLittleEndian<double>(&file);
    c.cx = readBinaryLittleEndian<double>(&file); c.cy = readBinaryLittleEndian<double>(&file);
    for (double& d : c.dist_coeffs) d = readBinaryLittleEndian<double>(&file);   // past the end of the file: zeros
    cameras.push_back(c);
  }
}

struct ColmapFrame {
  uint32_t image_id = 0, cam_id = 0;          // image_id = COLMAP id - 1
  double q_cw[4] = {0, 0, 0, 1};              // x y z w, normalised
  double p_cw[3] = {0, 0, 0};
  uint64_t w = 0, h = 0;
  std::string name;
};

struct ColmapScene {
  std::vector<ColmapFrame> frames;
  // every kept 2-D point, frame-major (Frame::points2D order): CSR pt_begin[F+1]
  std::vector<int32_t> pt_begin{0};
  std::vector<double> uv;                     // 2 per point
  std::vector<uint32_t> point2D_id;           // index of the point in the image's COLMAP list
  std::vector<uint64_t> point3D_id;           // COLMAP id
  std::vector<int32_t> lm_index;              // dense landmark index (-1: the 3-D point is not in points3D.bin)
  // points3D.bin in file order
  std::vector<uint64_t> lm_id;
  std::vector<double> lm_xyz, lm_error;       // 3 per point / 1 per point
  std::vector<int32_t> track_begin{0};        // CSR over (image_id - 1, point2D_idx)
  std::vector<uint32_t> track_image, track_point2D;
  std::unordered_map<uint64_t, int32_t> lm_of_id;
};

inline void readImagesAndPoints3DBinary(const std::string& path_images, const std::string& path_points3d, const std::vector<ColmapCamera>& cameras,
                                        ColmapScene& sc) {
  std::ifstream fi(path_images, std::ios::binary);
  if (!fi.is_open()) throw std::runtime_error("cannot open " + path_images);
  const uint64_t kInvalidPoint3DId = std::numeric_limits<uint64_t>::max();
  const size_t num_reg_images = readBinaryLittleEndian<uint64_t>(&fi);
  for (size_t i = 0; i < num_reg_images; ++i) {
    ColmapFrame fr;
    fr.image_id = readBinaryLittleEndian<uint32_t>(&fi) - 1;
    const double qw = readBinaryLittleEndian<double>(&fi), qx = readBinaryLittleEndian<double>(&fi), qy = readBinaryLittleEndian<double>(&fi),
                 qz = readBinaryLittleEndian<double>(&fi);
    const double n = std::sqrt(qw * qw + qx * qx + qy * qy + qz * qz);
    fr.q_cw[0] = qx / n; fr.q_cw[1] = qy / n; fr.q_cw[2] = qz / n; fr.q_cw[3] = qw / n;
    for (int c = 0; c < 3; ++c) fr.p_cw[c] = readBinaryLittleEndian<double>(&fi);
    fr.cam_id = readBinaryLittleEndian<uint32_t>(&fi);
    const size_t cam_idx = static_cast<size_t>(fr.cam_id) - 1;
    if (cam_idx >= cameras.size()) throw std::runtime_error("images.bin: camera id out of range");
    fr.w = cameras[cam_idx].w; fr.h = cameras[cam_idx].h;
    char ch;
    do { fi.read(&ch, 1); if (ch != '\0') fr.name += ch; } while (ch != '\0' && fi.good());
    const size_t num_potential_points2D = readBinaryLittleEndian<uint64_t>(&fi);
    std::unordered_map<double, size_t> x_coords, y_coords;
    std::vector<uint64_t> seen3d;
    for (size_t j = 0; j < num_potential_points2D; ++j) {
      const double x = readBinaryLittleEndian<double>(&fi), y = readBinaryLittleEndian<double>(&fi);
      const uint64_t id3 = readBinaryLittleEndian<uint64_t>(&fi);
      if (id3 == kInvalidPoint3DId) continue;
      auto sx = x_coords.find(x); auto sy = y_coords.find(y);
      if (sx != x_coords.end() && sy != y_coords.end() && sx->second == sy->second) continue;      // note 1 (:351-360)
      if (std::find(seen3d.begin(), seen3d.end(), id3) != seen3d.end()) continue;                  // note 2 (:363-369)
      sc.uv.push_back(x); sc.uv.push_back(y);
      sc.point2D_id.push_back(static_cast<uint32_t>(j));
      sc.point3D_id.push_back(id3);
      seen3d.push_back(id3);
      x_coords[x] = j; y_coords[y] = j;                                                            // insert_or_assign
    }
    sc.pt_begin.push_back((int32_t)sc.point3D_id.size());
    sc.frames.push_back(fr);
  }
  fi.close();

  std::ifstream fp(path_points3d, std::ios::binary);
  if (!fp.is_open()) throw std::runtime_error("cannot open " + path_points3d);
  const size_t num_points3D = readBinaryLittleEndian<uint64_t>(&fp);
  for (size_t i = 0; i < num_points3D; ++i) {
    const uint64_t id = readBinaryLittleEndian<uint64_t>(&fp);
    double xyz[3];
    for (int c = 0; c < 3; ++c) xyz[c] = readBinaryLittleEndian<double>(&fp);
    for (int c = 0; c < 3; ++c) readBinaryLittleEndian<uint8_t>(&fp);   // r g b
    const double err = readBinaryLittleEndian<double>(&fp);
    const size_t track_length = readBinaryLittleEndian<uint64_t>(&fp);
    std::vector<uint32_t> t_img(track_length), t_pt(track_length);
    for (size_t j = 0; j < track_length; ++j) { t_img[j] = readBinaryLittleEndian<uint32_t>(&fp) - 1; t_pt[j] = readBinaryLittleEndian<uint32_t>(&fp); }
    if (sc.lm_of_id.count(id)) continue;   // unordered_map::insert keeps the first entry of a repeated id (:437)
    sc.track_image.insert(sc.track_image.end(), t_img.begin(), t_img.end());
    sc.track_point2D.insert(sc.track_point2D.end(), t_pt.begin(), t_pt.end());
    sc.track_begin.push_back((int32_t)sc.track_image.size());
    sc.lm_of_id[id] = (int32_t)sc.lm_id.size();
    sc.lm_id.push_back(id); sc.lm_error.push_back(err);
    for (int c = 0; c < 3; ++c) sc.lm_xyz.push_back(xyz[c]);
  }
  sc.lm_index.resize(sc.point3D_id.size());
  for (size_t q = 0; q < sc.point3D_id.size(); ++q) {
    auto it = sc.lm_of_id.find(sc.point3D_id[q]);
    sc.lm_index[q] = it == sc.lm_of_id.end() ? -1 : it->second;
  }
}

}  // namespace io
}  // namespace ctba
