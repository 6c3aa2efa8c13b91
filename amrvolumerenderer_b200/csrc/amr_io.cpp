// amr_io.cpp -- the data formats on the input side of the path:
// an AMReX plotfile reader and a convexify for already-loaded levels.  Host
// code only; written from the published on-disk format and box calculus, since
// AMReX itself (26.04, CMakeLists.txt:44-52 of the reference) is not vendored.
//
// Reference call sites this file serves:
//   amrex::PlotFileData  -- VolumeRenderer/VolumeRenderer.cpp:599-679
//   amrex::convexify     -- VolumeRenderer/VolumeRenderer.cpp:669-670,
//                           VolumeRenderer/VolumeRendererApi.cpp:97-98
#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cctype>
#include <cerrno>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <system_error>
#include <thread>

#include "internal.hpp"

namespace {

using amrvr::report_error;

struct IBox {
  int32_t lo[3], hi[3];
  int64_t extent(int d) const { return static_cast<int64_t>(hi[d]) - lo[d] + 1; }
  int64_t cells() const { return extent(0) * extent(1) * extent(2); }
  bool ok() const { return hi[0] >= lo[0] && hi[1] >= lo[1] && hi[2] >= lo[2]; }
  // A box that holds data: not empty, and small enough that cells() and byte
  // counts stay far inside int64 (no real FAB is 2^20 cells long on a side).
  bool holds_data() const {
    return ok() && extent(0) <= (1 << 20) && extent(1) <= (1 << 20) && extent(2) <= (1 << 20);
  }
  bool intersects(const IBox& o) const {
    for (int d = 0; d < 3; ++d)
      if (std::max(lo[d], o.lo[d]) > std::min(hi[d], o.hi[d])) return false;
    return true;
  }
  bool contains(const IBox& o) const {
    for (int d = 0; d < 3; ++d)
      if (o.lo[d] < lo[d] || o.hi[d] > hi[d]) return false;
    return true;
  }
};

// ---- tiny text scanner over the plotfile's ASCII headers --------------------
struct Scanner {
  const char* p;
  const char* end;
  std::string what;  // file name, for messages
  void skip_ws() {
    while (p < end && std::isspace(static_cast<unsigned char>(*p))) ++p;
  }
  [[noreturn]] void bad(const std::string& msg) const {
    throw std::runtime_error(what + ": " + msg);
  }
  char peek() {
    skip_ws();
    return (p < end) ? *p : '\0';
  }
  void expect(char c) {
    skip_ws();
    if (p >= end || *p != c) bad(std::string("expected '") + c + "'");
    ++p;
  }
  std::string word() {
    skip_ws();
    const char* b = p;
    while (p < end && !std::isspace(static_cast<unsigned char>(*p))) ++p;
    if (b == p) bad("unexpected end of file");
    return std::string(b, p);
  }
  long long integer() {
    skip_ws();
    char* e = nullptr;
    errno = 0;
    const long long v = std::strtoll(p, &e, 10);
    if (e == p || errno != 0) bad("expected an integer");
    p = e;
    return v;
  }
  double real() {
    skip_ws();
    char* e = nullptr;
    const double v = std::strtod(p, &e);
    if (e == p) bad("expected a number");
    p = e;
    return v;
  }
  // "(i,j,k)" -- fewer than three entries are padded with `pad`
  void intvect(int32_t* v, int32_t pad) {
    expect('(');
    int n = 0;
    for (;;) {
      const long long x = integer();
      if (x < -(1LL << 30) || x > (1LL << 30)) bad("index out of range");
      if (n < 3) v[n] = static_cast<int32_t>(x);
      ++n;
      skip_ws();
      if (p < end && *p == ',') {
        ++p;
        continue;
      }
      break;
    }
    expect(')');
    for (; n < 3; ++n) v[n] = pad;
  }
  // amrex::Box text form "((lo) (hi) (type))"; the index type is optional in old files
  IBox box() {
    IBox b;
    expect('(');
    intvect(b.lo, 0);
    intvect(b.hi, 0);
    if (peek() == '(') {
      int32_t type[3];
      intvect(type, 0);
      if (type[0] != 0 || type[1] != 0 || type[2] != 0) bad("only cell-centred boxes are supported");
    }
    expect(')');
    return b;
  }
};

std::string slurp(const std::string& path) {
  std::ifstream in(path, std::ios::binary);
  if (!in) throw std::runtime_error("cannot open '" + path + "'");
  std::ostringstream ss;
  ss << in.rdbuf();
  return ss.str();
}

struct FabOnDisk {
  int file = -1;  // index into Plotfile::files
  int64_t offset = 0;
};

struct LevelData {
  amrvr_plotfile_level meta{};
  std::vector<IBox> boxes;
  std::vector<FabOnDisk> fod;
  int ncomp = 0;
};

}  // namespace

struct amrvr_plotfile {
  std::string path;
  amrvr_plotfile_info info{};
  std::vector<std::string> names;
  std::vector<LevelData> levels;
  std::vector<std::string> file_names;
  std::vector<int> fds;
  ~amrvr_plotfile() {
    for (int fd : fds)
      if (fd >= 0) ::close(fd);
  }
};

namespace {

// Level_n/Cell_H: amrex::VisMF::Header as operator<< writes it.
void parse_vismf_header(amrvr_plotfile& pf, LevelData& L, const std::string& mf_path) {
  const std::string text = slurp(mf_path + "_H");
  Scanner s{text.data(), text.data() + text.size(), mf_path + "_H"};
  const long long version = s.integer();
  if (version != 1)
    s.bad("VisMF header version " + std::to_string(version) +
          " is not supported (only Version_v1, FABs with their own text header)");
  (void)s.integer();  // how
  L.ncomp = static_cast<int>(s.integer());
  if (s.peek() == '(') {
    int32_t g[3];
    s.intvect(g, 0);
  } else {
    (void)s.integer();  // nGrow; each FAB record states its own (grown) box
  }
  s.expect('(');
  const long long n = s.integer();
  (void)s.integer();  // hash
  if (n < 0 || n > (1 << 24)) s.bad("implausible box count");
  L.boxes.resize(static_cast<size_t>(n));
  for (auto& b : L.boxes) {
    b = s.box();
    if (!b.holds_data()) s.bad("BoxArray holds an empty or implausibly large box");
  }
  s.expect(')');
  const long long nf = s.integer();
  if (nf != n) s.bad("FabOnDisk table and BoxArray differ in length");
  L.fod.resize(static_cast<size_t>(n));
  const std::string dir = mf_path.substr(0, mf_path.find_last_of('/') + 1);
  std::map<std::string, int> index;
  for (size_t i = 0; i < pf.file_names.size(); ++i) index[pf.file_names[i]] = static_cast<int>(i);
  for (auto& f : L.fod) {
    const std::string tag = s.word();
    if (tag != "FabOnDisk:") s.bad("expected 'FabOnDisk:'");
    const std::string full = dir + s.word();
    f.offset = s.integer();
    auto it = index.find(full);
    if (it == index.end()) {
      const int fd = ::open(full.c_str(), O_RDONLY);
      if (fd < 0) throw std::runtime_error("cannot open '" + full + "'");
      pf.file_names.push_back(full);
      pf.fds.push_back(fd);
      it = index.emplace(full, static_cast<int>(pf.fds.size()) - 1).first;
    }
    f.file = it->second;
  }
  L.meta.nboxes = static_cast<int32_t>(n);
}

struct FabRecord {
  IBox box;
  int ncomp = 0;
  int real_bytes = 8;
  bool swap = false;
  int64_t data_offset = 0;
};

// "FAB ((8, (64 11 52 0 1 12 0 1023)),(8, (8 7 6 5 4 3 2 1)))((lo) (hi) (t)) ncomp\n"
FabRecord read_fab_record(const amrvr_plotfile& pf, const FabOnDisk& f) {
  char buf[1024];
  const ssize_t got = ::pread(pf.fds[f.file], buf, sizeof(buf), f.offset);
  if (got <= 0) throw std::runtime_error("cannot read '" + pf.file_names[f.file] + "'");
  const char* nl = static_cast<const char*>(std::memchr(buf, '\n', static_cast<size_t>(got)));
  if (nl == nullptr) throw std::runtime_error(pf.file_names[f.file] + ": FAB header too long");
  Scanner s{buf, nl, pf.file_names[f.file]};
  if (s.word() != "FAB") s.bad("not a FAB record");
  FabRecord r;
  s.expect('(');
  s.expect('(');
  const long long nfmt = s.integer();
  s.expect(',');
  s.expect('(');
  long long bits = 0;
  for (long long i = 0; i < nfmt; ++i) {
    const long long v = s.integer();
    if (i == 0) bits = v;
  }
  s.expect(')');
  s.expect(')');
  s.expect(',');
  s.expect('(');
  const long long nord = s.integer();
  s.expect(',');
  s.expect('(');
  long long first = 0;
  for (long long i = 0; i < nord; ++i) {
    const long long v = s.integer();
    if (i == 0) first = v;
  }
  s.expect(')');
  s.expect(')');
  s.expect(')');
  if (!((bits == 64 && nord == 8) || (bits == 32 && nord == 4)))
    s.bad("unsupported real format (need IEEE 64- or 32-bit)");
  r.real_bytes = static_cast<int>(nord);
  uint16_t probe = 1;
  const bool host_little = *reinterpret_cast<const uint8_t*>(&probe) == 1;
  bool file_little;
  if (first == nord) file_little = true;
  else if (first == 1) file_little = false;
  else s.bad("unsupported byte order");
  r.swap = (file_little != host_little);
  r.box = s.box();
  r.ncomp = static_cast<int>(s.integer());
  r.data_offset = f.offset + (nl - buf) + 1;
  return r;
}

void pread_all(int fd, void* dst, size_t bytes, int64_t offset, const std::string& name) {
  char* out = static_cast<char*>(dst);
  while (bytes > 0) {
    const ssize_t got = ::pread(fd, out, bytes, offset);
    if (got <= 0) throw std::runtime_error("short read from '" + name + "'");
    out += got;
    offset += got;
    bytes -= static_cast<size_t>(got);
  }
}

template <typename T>
T byteswap(T v) {
  unsigned char* b = reinterpret_cast<unsigned char*>(&v);
  std::reverse(b, b + sizeof(T));
  return v;
}

void read_fab_component(const amrvr_plotfile& pf, int level, int box, int component, double* dst) {
  const LevelData& L = pf.levels[static_cast<size_t>(level)];
  const FabOnDisk& f = L.fod[static_cast<size_t>(box)];
  const IBox& valid = L.boxes[static_cast<size_t>(box)];
  const FabRecord r = read_fab_record(pf, f);
  const std::string& name = pf.file_names[f.file];
  if (component >= r.ncomp) throw std::runtime_error(name + ": FAB has too few components");
  if (!r.box.holds_data() || !r.box.contains(valid))
    throw std::runtime_error(name + ": FAB box does not cover its BoxArray entry");
  const int64_t fab_cells = r.box.cells();
  const int64_t at = r.data_offset + static_cast<int64_t>(component) * fab_cells * r.real_bytes;
  const bool same_box = std::memcmp(&r.box, &valid, sizeof(IBox)) == 0;
  if (same_box && r.real_bytes == 8) {
    pread_all(pf.fds[f.file], dst, static_cast<size_t>(fab_cells) * 8, at, name);
    if (r.swap)
      for (int64_t i = 0; i < fab_cells; ++i) dst[i] = byteswap(dst[i]);
    return;
  }
  // ghost cells on disk and/or single precision: stage, then copy the valid region
  std::vector<char> raw(static_cast<size_t>(fab_cells) * r.real_bytes);
  pread_all(pf.fds[f.file], raw.data(), raw.size(), at, name);
  const int64_t fx = r.box.hi[0] - r.box.lo[0] + 1, fy = r.box.hi[1] - r.box.lo[1] + 1;
  const int nx = valid.hi[0] - valid.lo[0] + 1, ny = valid.hi[1] - valid.lo[1] + 1,
            nz = valid.hi[2] - valid.lo[2] + 1;
  for (int k = 0; k < nz; ++k)
    for (int j = 0; j < ny; ++j) {
      const int64_t src = (valid.lo[0] - r.box.lo[0]) + fx * ((j + valid.lo[1] - r.box.lo[1]) +
                          fy * (k + valid.lo[2] - r.box.lo[2]));
      double* out = dst + (static_cast<int64_t>(k) * ny + j) * nx;
      if (r.real_bytes == 8) {
        const double* in = reinterpret_cast<const double*>(raw.data()) + src;
        for (int i = 0; i < nx; ++i) out[i] = r.swap ? byteswap(in[i]) : in[i];
      } else {
        const float* in = reinterpret_cast<const float*>(raw.data()) + src;
        for (int i = 0; i < nx; ++i) out[i] = static_cast<double>(r.swap ? byteswap(in[i]) : in[i]);
      }
    }
}

void parse_plotfile(amrvr_plotfile& pf) {
  const std::string text = slurp(pf.path + "/Header");
  Scanner s{text.data(), text.data() + text.size(), pf.path + "/Header"};
  (void)s.word();  // file version, e.g. HyperCLaw-V1.1
  const long long ncomp = s.integer();
  if (ncomp < 0 || ncomp > 100000) s.bad("implausible component count");
  pf.names.resize(static_cast<size_t>(ncomp));
  for (auto& n : pf.names) n = s.word();
  pf.info.ncomp = static_cast<int32_t>(ncomp);
  pf.info.space_dim = static_cast<int32_t>(s.integer());
  const int dim = pf.info.space_dim;
  if (dim < 1 || dim > 3) s.bad("space dimension must be 1, 2 or 3");
  pf.info.time = s.real();
  pf.info.finest_level = static_cast<int32_t>(s.integer());
  if (pf.info.finest_level < 0 || pf.info.finest_level > 64) s.bad("implausible finest level");
  const int nlev = pf.info.finest_level + 1;
  for (int d = 0; d < dim; ++d) pf.info.prob_lo[d] = s.real();
  for (int d = 0; d < dim; ++d) pf.info.prob_hi[d] = s.real();
  pf.levels.resize(static_cast<size_t>(nlev));
  for (int l = 0; l < pf.info.finest_level; ++l) {
    int32_t* r = pf.levels[static_cast<size_t>(l)].meta.ref_ratio;
    if (s.peek() == '(') {
      s.intvect(r, 1);
    } else {
      r[0] = r[1] = r[2] = static_cast<int32_t>(s.integer());
    }
    for (int d = dim; d < 3; ++d) r[d] = 1;  // VolumeRenderer.cpp:662-665
    if (r[0] < 1 || r[1] < 1 || r[2] < 1) s.bad("refinement ratio must be positive");
  }
  for (int l = 0; l < nlev; ++l) {
    const IBox b = s.box();
    for (int d = 0; d < 3; ++d) {
      pf.levels[static_cast<size_t>(l)].meta.domain_lo[d] = b.lo[d];
      pf.levels[static_cast<size_t>(l)].meta.domain_hi[d] = b.hi[d];
    }
  }
  for (int l = 0; l < nlev; ++l) (void)s.integer();  // level steps
  for (int l = 0; l < nlev; ++l) {
    double* dx = pf.levels[static_cast<size_t>(l)].meta.cell_size;
    dx[0] = dx[1] = dx[2] = 1.0;
    for (int d = 0; d < dim; ++d) dx[d] = s.real();
  }
  (void)s.integer();  // coordinate system
  (void)s.integer();  // boundary width
  for (int l = 0; l < nlev; ++l) {
    (void)s.integer();  // level
    const long long ngrids = s.integer();
    (void)s.real();     // time
    (void)s.integer();  // level steps
    for (long long g = 0; g < ngrids; ++g)
      for (int d = 0; d < dim; ++d) {
        (void)s.real();
        (void)s.real();
      }
    const std::string rel = s.word();
    LevelData& L = pf.levels[static_cast<size_t>(l)];
    if (ncomp > 0) {
      parse_vismf_header(pf, L, pf.path + "/" + rel);
      if (L.meta.nboxes != ngrids) s.bad("grid count differs from " + rel + "_H");
      L.meta.real_bytes = 8;
      if (!L.fod.empty()) L.meta.real_bytes = read_fab_record(pf, L.fod[0]).real_bytes;
    }
  }
}

bool level_ok(const amrvr_plotfile* pf, int level) {
  return pf != nullptr && level >= 0 && level < static_cast<int>(pf->levels.size());
}

// amrex::BoxList's boxDiff: b1 minus b2 as disjoint boxes, cutting the highest
// dimension first and the low side before the high side.
void box_diff(IBox b1, const IBox& b2, std::vector<IBox>& out) {
  if (b2.contains(b1)) return;
  if (!b1.intersects(b2)) {
    out.push_back(b1);
    return;
  }
  for (int d = 2; d >= 0; --d) {
    if (b1.lo[d] < b2.lo[d] && b2.lo[d] <= b1.hi[d]) {
      IBox bn = b1;
      bn.hi[d] = b2.lo[d] - 1;
      out.push_back(bn);
      b1.lo[d] = b2.lo[d];
    }
    if (b1.lo[d] <= b2.hi[d] && b2.hi[d] < b1.hi[d]) {
      IBox bn = b1;
      bn.lo[d] = b2.hi[d] + 1;
      out.push_back(bn);
      b1.hi[d] = b2.hi[d];
    }
  }
}

int32_t floor_div(int32_t a, int32_t b) {  // amrex::coarsen; in 64 bits so that -a cannot overflow
  const int64_t x = a, y = b;
  return static_cast<int32_t>((x >= 0) ? x / y : -((-x + y - 1) / y));
}

void convexify_pair(int ncoarse, const int32_t* coarse, int nfine, const int32_t* fine,
                    const int32_t* ratio, std::vector<IBox>& pieces, std::vector<int32_t>& source) {
  std::vector<IBox> cf(static_cast<size_t>(nfine));
  for (int f = 0; f < nfine; ++f)
    for (int d = 0; d < 3; ++d) {
      cf[static_cast<size_t>(f)].lo[d] = floor_div(fine[6 * f + d], ratio[d]);
      cf[static_cast<size_t>(f)].hi[d] = floor_div(fine[6 * f + 3 + d], ratio[d]);
    }
  std::vector<IBox> cur, next;
  for (int c = 0; c < ncoarse; ++c) {
    IBox b;
    for (int d = 0; d < 3; ++d) {
      b.lo[d] = coarse[6 * c + d];
      b.hi[d] = coarse[6 * c + 3 + d];
    }
    if (!b.ok()) continue;
    cur.assign(1, b);
    for (const IBox& f : cf) {
      if (!f.ok() || !f.intersects(b)) continue;
      next.clear();
      for (const IBox& p : cur) box_diff(p, f, next);
      cur.swap(next);
      if (cur.empty()) break;
    }
    for (const IBox& p : cur) {
      pieces.push_back(p);
      source.push_back(c);
    }
  }
}

}  // namespace

// ------------------------------------------------------------------ C ABI ---

extern "C" int amrvr_plotfile_open(const char* path, amrvr_plotfile** out) {
  if (path == nullptr || out == nullptr) return report_error(AMRVR_ERR_INVALID, "bad argument");
  *out = nullptr;
  if (path[0] == '\0')  // VolumeRenderer.cpp:595-597
    return report_error(AMRVR_ERR_INVALID, "Plotfile path must not be empty.");
  struct stat st;
  if (::stat(path, &st) != 0)  // VolumeRenderer.cpp:1464-1467
    return report_error(AMRVR_ERR_RUNTIME, std::string("plotfile path '") + path + "' does not exist");
  auto pf = std::make_unique<amrvr_plotfile>();
  pf->path = path;
  while (pf->path.size() > 1 && pf->path.back() == '/') pf->path.pop_back();
  try {
    parse_plotfile(*pf);
  } catch (const std::exception& e) {
    return report_error(AMRVR_ERR_RUNTIME, e.what());
  }
  *out = pf.release();
  return AMRVR_OK;
}

extern "C" void amrvr_plotfile_close(amrvr_plotfile* pf) { delete pf; }

extern "C" int amrvr_plotfile_get_info(const amrvr_plotfile* pf, amrvr_plotfile_info* info) {
  if (pf == nullptr || info == nullptr) return report_error(AMRVR_ERR_INVALID, "bad argument");
  *info = pf->info;
  return AMRVR_OK;
}

extern "C" const char* amrvr_plotfile_var_name(const amrvr_plotfile* pf, int component) {
  if (pf == nullptr || component < 0 || component >= static_cast<int>(pf->names.size())) return nullptr;
  return pf->names[static_cast<size_t>(component)].c_str();
}

extern "C" int amrvr_plotfile_var_index(const amrvr_plotfile* pf, const char* name) {
  if (pf == nullptr || name == nullptr) return -1;
  const auto it = std::find(pf->names.begin(), pf->names.end(), std::string(name));
  return (it == pf->names.end()) ? -1 : static_cast<int>(it - pf->names.begin());
}

extern "C" int amrvr_plotfile_get_level(const amrvr_plotfile* pf, int level,
                                        amrvr_plotfile_level* out) {
  if (!level_ok(pf, level) || out == nullptr) return report_error(AMRVR_ERR_INVALID, "bad level");
  *out = pf->levels[static_cast<size_t>(level)].meta;
  return AMRVR_OK;
}

extern "C" int amrvr_plotfile_level_boxes(const amrvr_plotfile* pf, int level, int32_t* lo_hi,
                                          int capacity) {
  if (!level_ok(pf, level) || lo_hi == nullptr) return report_error(AMRVR_ERR_INVALID, "bad level");
  const LevelData& L = pf->levels[static_cast<size_t>(level)];
  if (capacity < static_cast<int>(L.boxes.size()))
    return report_error(AMRVR_ERR_INVALID, "lo_hi is too small");
  for (size_t b = 0; b < L.boxes.size(); ++b) {
    std::memcpy(lo_hi + 6 * b, L.boxes[b].lo, 3 * sizeof(int32_t));
    std::memcpy(lo_hi + 6 * b + 3, L.boxes[b].hi, 3 * sizeof(int32_t));
  }
  return AMRVR_OK;
}

extern "C" int amrvr_plotfile_read_fab(const amrvr_plotfile* pf, int level, int box, int component,
                                       double* dst, int64_t capacity) {
  if (!level_ok(pf, level) || dst == nullptr) return report_error(AMRVR_ERR_INVALID, "bad level");
  const LevelData& L = pf->levels[static_cast<size_t>(level)];
  if (box < 0 || box >= static_cast<int>(L.boxes.size()))
    return report_error(AMRVR_ERR_INVALID, "box index out of range");
  if (component < 0 || component >= L.ncomp)
    return report_error(AMRVR_ERR_INVALID, "component out of range");
  if (capacity < L.boxes[static_cast<size_t>(box)].cells())
    return report_error(AMRVR_ERR_INVALID, "dst is too small");
  try {
    read_fab_component(*pf, level, box, component, dst);
  } catch (const std::exception& e) {
    return report_error(AMRVR_ERR_RUNTIME, e.what());
  }
  return AMRVR_OK;
}

extern "C" int amrvr_plotfile_read_level(const amrvr_plotfile* pf, int level, int component, int rank,
                                         int nranks, int rule, int threads, double* dst,
                                         int64_t capacity, int32_t* box_index_out,
                                         int64_t* offset_out, int box_capacity, int32_t* nread) {
  if (!level_ok(pf, level) || nread == nullptr || nranks < 1 || rank < 0 || rank >= nranks)
    return report_error(AMRVR_ERR_INVALID, "bad argument");
  const LevelData& L = pf->levels[static_cast<size_t>(level)];
  if (component < 0 || component >= L.ncomp)
    return report_error(AMRVR_ERR_INVALID, "component out of range");
  const int n = static_cast<int>(L.boxes.size());
  const int chunk = (n + nranks - 1) / nranks;
  std::vector<int32_t> mine;
  std::vector<int64_t> offs;
  int64_t total = 0;
  for (int b = 0; b < n; ++b) {
    const int owner = (rule == 1) ? std::min(nranks - 1, (chunk > 0) ? b / chunk : 0) : b % nranks;
    if (owner != rank) continue;
    mine.push_back(b);
    offs.push_back(total);
    total += L.boxes[static_cast<size_t>(b)].cells();
  }
  *nread = static_cast<int32_t>(mine.size());
  if (dst == nullptr && capacity == 0) {  // size query
    if (offset_out != nullptr && box_capacity > static_cast<int>(mine.size()))
      offset_out[mine.size()] = total;
    for (size_t k = 0; k < mine.size() && static_cast<int>(k) < box_capacity; ++k) {
      if (box_index_out != nullptr) box_index_out[k] = mine[k];
      if (offset_out != nullptr) offset_out[k] = offs[k];
    }
    return AMRVR_OK;
  }
  if (dst == nullptr || capacity < total) return report_error(AMRVR_ERR_INVALID, "dst is too small");
  if ((box_index_out != nullptr || offset_out != nullptr) && box_capacity < static_cast<int>(mine.size()))
    return report_error(AMRVR_ERR_INVALID, "box table is too small");
  for (size_t k = 0; k < mine.size(); ++k) {
    if (box_index_out != nullptr) box_index_out[k] = mine[k];
    if (offset_out != nullptr) offset_out[k] = offs[k];
  }
  const int nthreads = std::max(1, std::min<int>(threads, static_cast<int>(mine.size())));
  std::atomic<size_t> next{0};
  std::vector<std::string> errors(static_cast<size_t>(nthreads));
  auto work = [&](int t) {
    try {
      for (size_t k = next++; k < mine.size(); k = next++)
        read_fab_component(*pf, level, mine[k], component, dst + offs[k]);
    } catch (const std::exception& e) {
      errors[static_cast<size_t>(t)] = e.what();
    }
  };
  std::vector<std::thread> pool;
  for (int t = 1; t < nthreads; ++t) {
    try {
      pool.emplace_back(work, t);
    } catch (const std::system_error&) {
      break;  // no more threads to be had: the ones running share the work
    }
  }
  work(0);
  for (auto& t : pool) t.join();
  for (const auto& e : errors)
    if (!e.empty()) return report_error(AMRVR_ERR_RUNTIME, e);
  return AMRVR_OK;
}

extern "C" int amrvr_convexify_boxes(int ncoarse, const int32_t* coarse_lo_hi, int nfine,
                                     const int32_t* fine_lo_hi, const int32_t ratio[3],
                                     int32_t* out_lo_hi, int32_t* out_source, int capacity,
                                     int32_t* npieces) {
  if (ncoarse < 0 || nfine < 0 || (ncoarse > 0 && coarse_lo_hi == nullptr) ||
      (nfine > 0 && fine_lo_hi == nullptr) || ratio == nullptr || npieces == nullptr || capacity < 0)
    return report_error(AMRVR_ERR_INVALID, "bad argument");
  if (ratio[0] < 1 || ratio[1] < 1 || ratio[2] < 1)
    return report_error(AMRVR_ERR_INVALID, "refinement ratio must be positive");
  std::vector<IBox> pieces;
  std::vector<int32_t> source;
  convexify_pair(ncoarse, coarse_lo_hi, nfine, fine_lo_hi, ratio, pieces, source);
  *npieces = static_cast<int32_t>(pieces.size());
  if (static_cast<int>(pieces.size()) > capacity || (!pieces.empty() && out_lo_hi == nullptr))
    return report_error(AMRVR_ERR_INVALID, "out_lo_hi is too small");
  for (size_t k = 0; k < pieces.size(); ++k) {
    std::memcpy(out_lo_hi + 6 * k, pieces[k].lo, 3 * sizeof(int32_t));
    std::memcpy(out_lo_hi + 6 * k + 3, pieces[k].hi, 3 * sizeof(int32_t));
    if (out_source != nullptr) out_source[k] = source[k];
  }
  return AMRVR_OK;
}

extern "C" int amrvr_convexify_levels(int nlevels, const amrvr_level* levels,
                                      const int32_t* const* fine_boxes, const int32_t* nfine_boxes,
                                      const int32_t (*ref_ratio)[3], amrvr_fab* fabs_out, int capacity,
                                      int32_t* nfabs_out) {
  if (nlevels < 0 || (nlevels > 0 && (levels == nullptr || nfabs_out == nullptr)) || capacity < 0 ||
      (capacity > 0 && fabs_out == nullptr) || (nlevels > 1 && ref_ratio == nullptr))
    return report_error(AMRVR_ERR_INVALID, "bad argument");
  int written = 0;
  std::vector<int32_t> coarse, fine_local;
  for (int l = 0; l < nlevels; ++l) {
    const amrvr_level& L = levels[l];
    if (L.nfabs < 0 || (L.nfabs > 0 && L.fabs == nullptr))
      return report_error(AMRVR_ERR_INVALID, "bad level");
    nfabs_out[l] = 0;
    if (l == nlevels - 1) {  // the finest level is kept as it is
      if (written + L.nfabs > capacity) return report_error(AMRVR_ERR_INVALID, "fabs_out is too small");
      for (int f = 0; f < L.nfabs; ++f) fabs_out[written++] = L.fabs[f];
      nfabs_out[l] = L.nfabs;
      break;
    }
    coarse.resize(static_cast<size_t>(L.nfabs) * 6);
    for (int f = 0; f < L.nfabs; ++f) {
      std::memcpy(&coarse[6 * static_cast<size_t>(f)], L.fabs[f].lo, 3 * sizeof(int32_t));
      std::memcpy(&coarse[6 * static_cast<size_t>(f) + 3], L.fabs[f].hi, 3 * sizeof(int32_t));
    }
    const int32_t* fb = (fine_boxes != nullptr) ? fine_boxes[l + 1] : nullptr;
    int nfb = (fine_boxes != nullptr && nfine_boxes != nullptr) ? nfine_boxes[l + 1] : 0;
    if (fb == nullptr) {
      const amrvr_level& F = levels[l + 1];
      fine_local.resize(static_cast<size_t>(F.nfabs) * 6);
      for (int f = 0; f < F.nfabs; ++f) {
        std::memcpy(&fine_local[6 * static_cast<size_t>(f)], F.fabs[f].lo, 3 * sizeof(int32_t));
        std::memcpy(&fine_local[6 * static_cast<size_t>(f) + 3], F.fabs[f].hi, 3 * sizeof(int32_t));
      }
      fb = fine_local.data();
      nfb = F.nfabs;
    }
    if (ref_ratio[l][0] < 1 || ref_ratio[l][1] < 1 || ref_ratio[l][2] < 1)
      return report_error(AMRVR_ERR_INVALID, "refinement ratio must be positive");
    std::vector<IBox> pieces;
    std::vector<int32_t> source;
    convexify_pair(L.nfabs, coarse.data(), nfb, fb, ref_ratio[l], pieces, source);
    if (written + static_cast<int>(pieces.size()) > capacity)
      return report_error(AMRVR_ERR_INVALID, "fabs_out is too small");
    for (size_t k = 0; k < pieces.size(); ++k) {
      amrvr_fab v = L.fabs[source[k]];  // same storage, same Array4 origin and strides
      std::memcpy(v.lo, pieces[k].lo, 3 * sizeof(int32_t));
      std::memcpy(v.hi, pieces[k].hi, 3 * sizeof(int32_t));
      fabs_out[written++] = v;
    }
    nfabs_out[l] = static_cast<int32_t>(pieces.size());
  }
  return AMRVR_OK;
}
