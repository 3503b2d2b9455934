// Parses a COLMAP binary export with drivers/ctba_colmap.hpp and prints what it kept, for tests/test_drivers.py.
#include <cstdio>
#include "../../drivers/ctba_colmap.hpp"
using namespace ctba::io;
int main(int argc, char** argv) {
  if (argc < 2) return 1;
  const std::string d = argv[1];
  try {
    std::vector<ColmapCamera> cams;
    readCamerasBinary(d + "/cameras.bin", cams);
    for (const auto& c : cams)
      std::printf("cam %u %d %llu %llu %.17g %.17g %.17g %.17g %zu %.17g %.17g %.17g %.17g\n", c.cam_id, c.model_id, (unsigned long long)c.w, (unsigned long long)c.h,
                  c.fx, c.fy, c.cx, c.cy, c.dist_coeffs.size(), c.dist_coeffs[0], c.dist_coeffs[1], c.dist_coeffs[2], c.dist_coeffs[3]);
    ColmapScene sc;
    readImagesAndPoints3DBinary(d + "/images.bin", d + "/points3D.bin", cams, sc);
    for (size_t f = 0; f < sc.frames.size(); ++f) {
      const ColmapFrame& fr = sc.frames[f];
      std::printf("frame %u %u %s %.17g %.17g %.17g %.17g %.17g %.17g %.17g %d\n", fr.image_id, fr.cam_id, fr.name.c_str(), fr.q_cw[0], fr.q_cw[1], fr.q_cw[2],
                  fr.q_cw[3], fr.p_cw[0], fr.p_cw[1], fr.p_cw[2], sc.pt_begin[f + 1] - sc.pt_begin[f]);
      for (int q = sc.pt_begin[f]; q < sc.pt_begin[f + 1]; ++q)
        std::printf("pt %zu %u %.17g %.17g %llu %d\n", f, sc.point2D_id[q], sc.uv[2 * q], sc.uv[2 * q + 1], (unsigned long long)sc.point3D_id[q], sc.lm_index[q]);
    }
    for (size_t l = 0; l < sc.lm_id.size(); ++l)
      std::printf("lm %zu %llu %.17g %.17g %.17g %.17g %d\n", l, (unsigned long long)sc.lm_id[l], sc.lm_xyz[3 * l], sc.lm_xyz[3 * l + 1], sc.lm_xyz[3 * l + 2],
                  sc.lm_error[l], sc.track_begin[l + 1] - sc.track_begin[l]);
  } catch (const std::exception& e) { std::printf("ERROR %s\n", e.what()); return 2; }
  std::printf("OK\n");
  return 0;
}
