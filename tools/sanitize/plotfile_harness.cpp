#include "../../include/amrvr.h"
#include <cstdio>
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
}
