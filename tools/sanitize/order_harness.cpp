#include "internal.hpp"  // -I amrvolumerenderer_b200/csrc
#include <cstdio>
#include <random>
#include <string>
namespace amrvr { int report_error(int status, const std::string& msg) { (void)msg; return status; } }
using namespace amrvr;
int main() {
  std::mt19937 rng(5);
  std::uniform_real_distribution<double> u(0.0, 1.0);
  long checksum = 0;
  for (int trial = 0; trial < 300; ++trial) {
    int n = 1 + (rng() % (trial < 295 ? 60 : 300));
    int P = 1 + rng() % 8;
    std::vector<amrvr_box_record> recs(n);
    for (int i = 0; i < n; ++i) {
      recs[i] = amrvr_box_record{};
      for (int d = 0; d < 3; ++d) {
        double lo = (rng() % 8) / 8.0, len = (1 + rng() % 3) / 8.0;
        recs[i].min_corner[d] = lo; recs[i].max_corner[d] = lo + len; recs[i].dims[d] = 4;
      }
      recs[i].owner = (int)((long)i * P / n);
    }
    amrvr_camera cam{};
    for (int d = 0; d < 3; ++d) { cam.eye[d] = -2 + 5 * u(rng); cam.look_at[d] = 0.5; }
    cam.up[1] = 1; cam.fov_y_degrees = 45; cam.near_plane = 0.1f; cam.far_plane = 10;
    std::vector<int> order(P);
    std::vector<std::string> dots;
    bool ok = order_ranks(P, n, recs.data(), cam, 1.3f, true, *host_order_backend(), order.data(), trial % 7 == 0 ? &dots : nullptr);
    checksum += ok; for (int r : order) checksum += r;
    std::vector<float> hints(n); std::vector<int> lorder(n);
    layer_order(n, recs.data(), cam, hints.data(), lorder.data());
    double bmin[3] = {0,0,0}, bmax[3] = {1.4,1.4,1.4};
    checksum += (long)(1e6 * reference_sample_distance(n, recs.data(), bmin, bmax));
    CameraBasis basis; make_camera_basis(cam, basis);
    // footprints (incl. cameras inside and behind boxes), overlay edges, automatic camera
    const int W = 1 + static_cast<int>(rng() % 2048), H = 1 + static_cast<int>(rng() % 2048);
    std::vector<Rect> rects(n);
    box_footprints(n, recs.data(), basis, cam, W, H, rects.data());
    for (const Rect& r : rects) {
      if (r.w < 0 || r.h < 0 || r.x0 < 0 || r.y0 < 0 || r.y0 + r.h > H) { std::printf("bad rect\n"); return 1; }
      checksum += r.w + r.h;
    }
    OverlayEdges edges;
    make_overlay_edges(bmin, bmax, cam, 1 + static_cast<int>(rng() % 3), W, H, edges);
    checksum += edges.count;
    amrvr_camera autocam{};
    const double up[3] = {0.0, 0.0, 1.0};
    amrvr_auto_camera(bmin, bmax, static_cast<unsigned>(rng()), (trial & 1) ? up : nullptr, &autocam);
    checksum += static_cast<long>(autocam.far_plane);
  }
  float table[1024];
  build_color_table(1.0f, 0.5f, 0.0f, 1.0f, nullptr, 0, table);
  std::printf("checksum %ld\n", checksum);
  return 0;
}
