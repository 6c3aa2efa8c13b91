#include "../../include/amrvr.h"
#include <cstdint>
#include <cstdio>
#include <utility>
#include <random>
#include <string>
#include <vector>
namespace amrvr { int report_error(int status, const std::string& msg) { (void)msg; return status; } }
int main(int argc, char** argv) {
  int opened = 0, refused = 0;
  for (int a = 1; a < argc; ++a) {
    amrvr_plotfile* pf = nullptr;
    if (amrvr_plotfile_open(argv[a], &pf) != 0) { ++refused; continue; }
    amrvr_plotfile_info info{};
    amrvr_plotfile_get_info(pf, &info);
    bool bad = false;
    for (int l = 0; l <= info.finest_level; ++l) {
      amrvr_plotfile_level L{};
      if (amrvr_plotfile_get_level(pf, l, &L) != 0) { bad = true; continue; }
      std::vector<int32_t> boxes(6 * (size_t)(L.nboxes > 0 ? L.nboxes : 1));
      if (amrvr_plotfile_level_boxes(pf, l, boxes.data(), L.nboxes) != 0) { bad = true; continue; }
      int64_t total = 0;
      for (int b = 0; b < L.nboxes; ++b) {
        int64_t c = 1; for (int d = 0; d < 3; ++d) c *= (int64_t)boxes[6*b+3+d] - boxes[6*b+d] + 1;
        total += c;
      }
      if (total > (1 << 26)) { bad = true; continue; }
      std::vector<double> dst((size_t)total + 1);
      for (int comp = 0; comp < info.ncomp && comp < 2; ++comp) {
        int32_t nread = 0;
        if (amrvr_plotfile_read_level(pf, l, comp, 0, 1, 0, 3, dst.data(), total, nullptr, nullptr, 0, &nread) != 0) bad = true;
      }
      for (int b = 0; b < L.nboxes && b < 3; ++b) {
        int64_t c = 1; for (int d = 0; d < 3; ++d) c *= (int64_t)boxes[6*b+3+d] - boxes[6*b+d] + 1;
        std::vector<double> one((size_t)c);
        if (amrvr_plotfile_read_fab(pf, l, b, 0, one.data(), c) != 0) bad = true;
      }
    }
    amrvr_plotfile_close(pf);
    if (bad) ++refused; else ++opened;
  }
  std::printf("opened %d refused %d\n", opened, refused);

  // convexify on random (overlapping, misaligned, partly outside) box sets
  std::mt19937 rng(11);
  long pieces_total = 0;
  for (int trial = 0; trial < 2000; ++trial) {
    const int nc = static_cast<int>(rng() % 12), nf = static_cast<int>(rng() % 12);
    std::vector<int32_t> coarse(6 * static_cast<size_t>(nc) + 6), fine(6 * static_cast<size_t>(nf) + 6);
    auto fill = [&](std::vector<int32_t>& v, int n, int span) {
      for (int b = 0; b < n; ++b)
        for (int d = 0; d < 3; ++d) {
          const int lo = static_cast<int>(rng() % span) - 4;
          v[6 * b + d] = lo;
          v[6 * b + 3 + d] = lo + static_cast<int>(rng() % 9);
        }
    };
    fill(coarse, nc, 24);
    fill(fine, nf, 48);
    const int32_t ratio[3] = {1 + static_cast<int32_t>(rng() % 4), 1 + static_cast<int32_t>(rng() % 4),
                              1 + static_cast<int32_t>(rng() % 4)};
    int32_t npieces = 0;
    amrvr_convexify_boxes(nc, coarse.data(), nf, fine.data(), ratio, nullptr, nullptr, 0, &npieces);
    std::vector<int32_t> out(6 * static_cast<size_t>(npieces) + 6), src(static_cast<size_t>(npieces) + 1);
    if (amrvr_convexify_boxes(nc, coarse.data(), nf, fine.data(), ratio, out.data(), src.data(), npieces,
                              &npieces) != 0) {
      std::printf("convexify failed at trial %d\n", trial);
      return 1;
    }
    pieces_total += npieces;
  }
  std::printf("convexify pieces %ld\n", pieces_total);

  // the same with coordinates at the ends of the int32 range mixed in
  const int32_t ends[] = {INT32_MIN, INT32_MIN + 1, -1, 0, 1, INT32_MAX - 1, INT32_MAX, 1 << 30, -(1 << 30)};
  long extreme_total = 0;
  for (int trial = 0; trial < 20000; ++trial) {
    const int nc = 1 + static_cast<int>(rng() % 3), nf = 1 + static_cast<int>(rng() % 3);
    std::vector<int32_t> coarse(6 * static_cast<size_t>(nc)), fine(6 * static_cast<size_t>(nf));
    auto fill = [&](std::vector<int32_t>& v, int n) {
      for (int b = 0; b < n; ++b)
        for (int d = 0; d < 3; ++d) {
          int32_t a = ends[rng() % 9], z = ends[rng() % 9];
          if (rng() % 2) {
            a = static_cast<int32_t>(rng() % 16) - 4;
            z = a + static_cast<int32_t>(rng() % 8);
          }
          if (a > z) std::swap(a, z);
          v[6 * b + d] = a;
          v[6 * b + 3 + d] = z;
        }
    };
    fill(coarse, nc);
    fill(fine, nf);
    const int32_t ratio[3] = {1 + static_cast<int32_t>(rng() % 4), 1 + static_cast<int32_t>(rng() % 4),
                              1 + static_cast<int32_t>(rng() % 4)};
    int32_t npieces = 0;
    std::vector<int32_t> out(6 * 4096), src(4096);
    amrvr_convexify_boxes(nc, coarse.data(), nf, fine.data(), ratio, out.data(), src.data(), 4096, &npieces);
    extreme_total += npieces;
  }
  std::printf("convexify pieces at the int32 ends %ld\n", extreme_total);
}
