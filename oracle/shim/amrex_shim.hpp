// oracle/shim/amrex_shim.hpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// A header-only stand-in for the handful of AMReX 26.04 symbols the reference's
// hot-path sources use, so that those sources can be compiled UNMODIFIED from
// /root/reference by oracle/Makefile into oracle/_ref/ (AMReX itself is not
// available in this image and cannot be fetched).  Everything here is written
// from the public AMReX semantics (CPU build: Real = double, serial
// ParallelFor, host Math::rsqrt = 1/sqrt); nothing is copied from AMReX or the
// reference.  "parity unpinned" caveat: AMReX's own arithmetic order for
// RealVect/SmallMatrix is restated here from memory of its documented
// behaviour; all of it feeds double->float casts or ordering keys only.
//
// Sample counting: Array4::operator() bumps a thread-local counter so the
// oracle can report "samples" (one Array4 read == one execution of the body at
// Common/VolumePainter.cpp:869-894) without touching the reference source.
#ifndef AMRVR_ORACLE_AMREX_SHIM_HPP
#define AMRVR_ORACLE_AMREX_SHIM_HPP

#include <mpi.h>

#include <algorithm>
#include <array>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <map>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <type_traits>
#include <utility>
#include <vector>

#define AMREX_SPACEDIM 3
#define AMREX_GPU_DEVICE
#define AMREX_GPU_HOST_DEVICE
#define AMREX_GPU_HOST
#define AMREX_FORCE_INLINE inline

namespace amrex {

using Real = double;
using Long = long long;

template <class T>
using Vector = std::vector<T>;
template <class T, std::size_t N>
using Array = std::array<T, N>;

// ---------------------------------------------------------------- sample counter
namespace shim {
inline thread_local long long array4_reads = 0;
}

// ---------------------------------------------------------------- IntVect / Box
struct Dim3 {
  int x, y, z;
};

class IntVect {
 public:
  IntVect() : v{0, 0, 0} {}
  explicit IntVect(int s) : v{s, s, s} {}
  IntVect(int i, int j, int k) : v{i, j, k} {}
  int& operator[](int d) { return v[d]; }
  const int& operator[](int d) const { return v[d]; }
  bool operator==(const IntVect& o) const {
    return v[0] == o.v[0] && v[1] == o.v[1] && v[2] == o.v[2];
  }

 private:
  int v[3];
};

class Box {
 public:
  Box() : lo(0), hi(-1) {}
  Box(const IntVect& l, const IntVect& h) : lo(l), hi(h) {}
  const IntVect& smallEnd() const { return lo; }
  const IntVect& bigEnd() const { return hi; }
  int length(int d) const { return hi[d] - lo[d] + 1; }
  IntVect length() const { return IntVect(length(0), length(1), length(2)); }
  Long numPts() const {
    return static_cast<Long>(length(0)) * length(1) * length(2);
  }

 private:
  IntVect lo, hi;
};

// ---------------------------------------------------------------- RealVect
class RealVect {
 public:
  RealVect() : v{0.0, 0.0, 0.0} {}
  explicit RealVect(Real s) : v{s, s, s} {}
  RealVect(Real x, Real y, Real z) : v{x, y, z} {}
  Real& operator[](int d) { return v[d]; }
  const Real& operator[](int d) const { return v[d]; }

  RealVect operator+(const RealVect& o) const {
    return RealVect(v[0] + o.v[0], v[1] + o.v[1], v[2] + o.v[2]);
  }
  RealVect operator-(const RealVect& o) const {
    return RealVect(v[0] - o.v[0], v[1] - o.v[1], v[2] - o.v[2]);
  }
  RealVect operator-() const { return RealVect(-v[0], -v[1], -v[2]); }
  RealVect operator*(Real s) const {
    return RealVect(v[0] * s, v[1] * s, v[2] * s);
  }
  RealVect operator/(Real s) const {
    return RealVect(v[0] / s, v[1] / s, v[2] / s);
  }
  RealVect& operator+=(const RealVect& o) {
    v[0] += o.v[0];
    v[1] += o.v[1];
    v[2] += o.v[2];
    return *this;
  }
  RealVect& operator-=(const RealVect& o) {
    v[0] -= o.v[0];
    v[1] -= o.v[1];
    v[2] -= o.v[2];
    return *this;
  }
  RealVect& operator*=(Real s) {
    v[0] *= s;
    v[1] *= s;
    v[2] *= s;
    return *this;
  }
  RealVect& operator/=(Real s) {
    v[0] /= s;
    v[1] /= s;
    v[2] /= s;
    return *this;
  }
  Real dotProduct(const RealVect& o) const {
    return v[0] * o.v[0] + v[1] * o.v[1] + v[2] * o.v[2];
  }
  RealVect crossProduct(const RealVect& o) const {
    return RealVect(v[1] * o.v[2] - v[2] * o.v[1],
                    v[2] * o.v[0] - v[0] * o.v[2],
                    v[0] * o.v[1] - v[1] * o.v[0]);
  }
  Real radSquared() const { return v[0] * v[0] + v[1] * v[1] + v[2] * v[2]; }
  Real vectorLength() const { return std::sqrt(radSquared()); }

 private:
  Real v[3];
};

inline RealVect operator*(Real s, const RealVect& r) { return r * s; }

// ---------------------------------------------------------------- Array4
template <class T>
struct Array4 {
  T* p = nullptr;
  Long jstride = 0;
  Long kstride = 0;
  Long nstride = 0;
  Dim3 begin{0, 0, 0};
  Dim3 end{0, 0, 0};
  int ncomp = 0;

  Array4() = default;
  Array4(T* ptr, Dim3 b, Dim3 e, int nc)
      : p(ptr),
        jstride(e.x - b.x),
        kstride(static_cast<Long>(e.x - b.x) * (e.y - b.y)),
        nstride(static_cast<Long>(e.x - b.x) * (e.y - b.y) * (e.z - b.z)),
        begin(b),
        end(e),
        ncomp(nc) {}
  template <class U,
            std::enable_if_t<std::is_same_v<std::remove_const_t<T>, U> &&
                                 std::is_const_v<T>,
                             int> = 0>
  Array4(const Array4<U>& o)
      : p(o.p),
        jstride(o.jstride),
        kstride(o.kstride),
        nstride(o.nstride),
        begin(o.begin),
        end(o.end),
        ncomp(o.ncomp) {}

  T& operator()(int i, int j, int k, int n = 0) const noexcept {
    ++shim::array4_reads;
    return p[(i - begin.x) + (j - begin.y) * jstride + (k - begin.z) * kstride +
             n * nstride];
  }
};

// ---------------------------------------------------------------- SmallMatrix
// Column-vector convention; operator* accumulates k = 0..N-1 in order from a
// zero-initialised element, as AMReX's SmallMatrix does.
template <class T, int R, int C>
class SmallMatrix {
 public:
  SmallMatrix() : a{} {}
  template <class... Ts,
            std::enable_if_t<(sizeof...(Ts) == R * C) && (R * C > 1), int> = 0>
  explicit SmallMatrix(Ts... vs) : a{static_cast<T>(vs)...} {}

  static SmallMatrix Identity() {
    SmallMatrix m;
    for (int i = 0; i < (R < C ? R : C); ++i) m(i, i) = T(1);
    return m;
  }
  T& operator()(int r, int c) { return a[r + c * R]; }
  const T& operator()(int r, int c) const { return a[r + c * R]; }
  T& operator[](int i) { return a[i]; }
  const T& operator[](int i) const { return a[i]; }

  template <int C2>
  SmallMatrix<T, R, C2> operator*(const SmallMatrix<T, C, C2>& rhs) const {
    SmallMatrix<T, R, C2> out;
    for (int j = 0; j < C2; ++j) {
      for (int i = 0; i < R; ++i) {
        T acc = T(0);
        for (int k = 0; k < C; ++k) acc += (*this)(i, k) * rhs(k, j);
        out(i, j) = acc;
      }
    }
    return out;
  }

 private:
  T a[R * C];
};

// ---------------------------------------------------------------- GpuArray / tuples
template <class T, unsigned int N>
struct GpuArray {
  T arr[N];
  T& operator[](int i) { return arr[i]; }
  const T& operator[](int i) const { return arr[i]; }
  T* data() { return arr; }
  const T* data() const { return arr; }
};

template <class... Ts>
using GpuTuple = std::tuple<Ts...>;
template <std::size_t I, class Tup>
auto get(const Tup& t) -> decltype(std::get<I>(t)) {
  return std::get<I>(t);
}
template <class... Ts>
struct TypeList {};

// ---------------------------------------------------------------- Math
namespace Math {
template <class T>
T floor(T x) {
  return std::floor(x);
}
template <class T>
T abs(T x) {
  return std::abs(x);
}
template <class T>
T rsqrt(T x) {
  return static_cast<T>(1. / std::sqrt(x));
}
template <class T>
bool isfinite(T x) {
  return std::isfinite(x);
}
}  // namespace Math

// ---------------------------------------------------------------- Gpu
namespace Gpu {
template <class T>
using DeviceVector = std::vector<T>;
template <class T>
using HostVector = std::vector<T>;
struct HostToDevice {};
struct DeviceToHost {};
inline constexpr HostToDevice hostToDevice{};
inline constexpr DeviceToHost deviceToHost{};
template <class Tag, class It, class Out>
void copyAsync(Tag, It b, It e, Out o) {
  std::copy(b, e, o);
}
inline void streamSynchronize() {}
namespace Atomic {
template <class T>
void Add(T* p, T v) {
  *p += v;
}
}  // namespace Atomic
}  // namespace Gpu

// ---------------------------------------------------------------- launches
template <class I, class F, std::enable_if_t<std::is_integral_v<I>, int> = 0>
void ParallelFor(I n, F&& f) {
  for (I i = 0; i < n; ++i) f(i);
}
template <class F>
void ParallelFor(const Box& b, F&& f) {
  for (int k = b.smallEnd()[2]; k <= b.bigEnd()[2]; ++k)
    for (int j = b.smallEnd()[1]; j <= b.bigEnd()[1]; ++j)
      for (int i = b.smallEnd()[0]; i <= b.bigEnd()[0]; ++i) f(i, j, k);
}

// ---------------------------------------------------------------- reductions
struct ReduceOpMin {
  template <class T>
  static void apply(T& d, const T& s) {
    d = std::min(d, s);
  }
  template <class T>
  static T init() {
    return std::numeric_limits<T>::max();
  }
};
struct ReduceOpMax {
  template <class T>
  static void apply(T& d, const T& s) {
    d = std::max(d, s);
  }
  template <class T>
  static T init() {
    return std::numeric_limits<T>::lowest();
  }
};
struct ReduceOpSum {
  template <class T>
  static void apply(T& d, const T& s) {
    d += s;
  }
  template <class T>
  static T init() {
    return T(0);
  }
};

namespace shim {
template <class Ops, class Tup, std::size_t... Is>
void reduce_apply(Tup& dst, const Tup& src, std::index_sequence<Is...>) {
  (std::tuple_element_t<Is, Ops>::apply(std::get<Is>(dst), std::get<Is>(src)),
   ...);
}
template <class Ops, class Tup, std::size_t... Is>
Tup reduce_init(std::index_sequence<Is...>) {
  return Tup(std::tuple_element_t<Is, Ops>::template init<
             std::tuple_element_t<Is, Tup>>()...);
}
}  // namespace shim

template <class... Ops>
class ReduceOps {
 public:
  using OpTuple = std::tuple<Ops...>;
  template <class N, class D, class F>
  void eval(N n, D& data, F&& f) {
    for (N i = 0; i < n; ++i) {
      auto r = f(i);
      shim::reduce_apply<OpTuple>(data.acc, r,
                                  std::index_sequence_for<Ops...>{});
    }
  }
};

template <class... Ts>
class ReduceData {
 public:
  using Tup = std::tuple<Ts...>;
  template <class... Ops>
  explicit ReduceData(ReduceOps<Ops...>&)
      : acc(shim::reduce_init<std::tuple<Ops...>, Tup>(
            std::index_sequence_for<Ts...>{})) {}
  template <class R>
  Tup value(R&) {
    return acc;
  }
  Tup acc;
};

// ---------------------------------------------------------------- mini MultiFab
// Just enough of BoxArray/DistributionMapping/MultiFab/MFIter for
// VolumeRenderer/SceneBuilder.cpp: a list of (box, owner) with local FAB
// storage in Fortran order, no ghost cells.
class MultiFab {
 public:
  MultiFab() = default;
  MultiFab(std::vector<Box> boxes, std::vector<int> owners, int ncomp,
           int myRank)
      : m_boxes(std::move(boxes)),
        m_owners(std::move(owners)),
        m_ncomp(ncomp),
        m_rank(myRank) {
    m_data.resize(m_boxes.size());
    for (std::size_t i = 0; i < m_boxes.size(); ++i) {
      if (m_owners[i] == m_rank) {
        m_local.push_back(static_cast<int>(i));
        m_data[i].assign(
            static_cast<std::size_t>(m_boxes[i].numPts()) * ncomp, 0.0);
      }
    }
  }
  MultiFab(MultiFab&&) = default;
  MultiFab& operator=(MultiFab&&) = default;
  MultiFab(const MultiFab&) = delete;
  MultiFab& operator=(const MultiFab&) = delete;

  MultiFab deepCopy() const {
    MultiFab c(m_boxes, m_owners, m_ncomp, m_rank);
    c.m_data = m_data;
    return c;
  }

  int size() const { return static_cast<int>(m_boxes.size()); }
  int nComp() const { return m_ncomp; }
  int localCount() const { return static_cast<int>(m_local.size()); }
  int localToGlobal(int li) const { return m_local[li]; }
  const Box& box(int gi) const { return m_boxes[gi]; }
  int owner(int gi) const { return m_owners[gi]; }
  int myRank() const { return m_rank; }
  Real* dataPtr(int gi) { return m_data[gi].data(); }
  const Real* dataPtr(int gi) const { return m_data[gi].data(); }

  Array4<Real const> const_array_global(int gi) const {
    const Box& b = m_boxes[gi];
    return Array4<Real const>(
        m_data[gi].data(),
        Dim3{b.smallEnd()[0], b.smallEnd()[1], b.smallEnd()[2]},
        Dim3{b.bigEnd()[0] + 1, b.bigEnd()[1] + 1, b.bigEnd()[2] + 1}, m_ncomp);
  }

  // MultiArray4-like view indexed by LOCAL box number.
  struct ConstArrays {
    const MultiFab* mf;
    Array4<Real const> operator[](int li) const {
      return mf->const_array_global(mf->m_local[li]);
    }
  };
  ConstArrays const_arrays() const { return ConstArrays{this}; }

  template <class MFI>
  Array4<Real const> const_array(const MFI& mfi) const {
    return const_array_global(mfi.index());
  }

 private:
  friend class MFIter;
  std::vector<Box> m_boxes;
  std::vector<int> m_owners;
  std::vector<std::vector<Real>> m_data;
  std::vector<int> m_local;
  int m_ncomp = 0;
  int m_rank = 0;
};

class MFIter {
 public:
  explicit MFIter(const MultiFab& mf) : m_mf(&mf), m_pos(0) {}
  bool isValid() const { return m_pos < m_mf->localCount(); }
  MFIter& operator++() {
    ++m_pos;
    return *this;
  }
  int index() const { return m_mf->m_local[m_pos]; }
  int LocalIndex() const { return m_pos; }
  const Box& validbox() const { return m_mf->m_boxes[index()]; }

 private:
  const MultiFab* m_mf;
  int m_pos;
};

template <class... Ops, class... Ts, class F>
std::tuple<Ts...> ParReduce(TypeList<Ops...>, TypeList<Ts...>,
                            const MultiFab& mf, const IntVect&, F&& f) {
  using Tup = std::tuple<Ts...>;
  using OpTuple = std::tuple<Ops...>;
  Tup acc = shim::reduce_init<OpTuple, Tup>(std::index_sequence_for<Ts...>{});
  for (int li = 0; li < mf.localCount(); ++li) {
    const Box& b = mf.box(mf.localToGlobal(li));
    for (int k = b.smallEnd()[2]; k <= b.bigEnd()[2]; ++k)
      for (int j = b.smallEnd()[1]; j <= b.bigEnd()[1]; ++j)
        for (int i = b.smallEnd()[0]; i <= b.bigEnd()[0]; ++i) {
          Tup r = f(li, i, j, k);
          shim::reduce_apply<OpTuple>(acc, r,
                                      std::index_sequence_for<Ts...>{});
        }
  }
  return acc;
}

namespace ParallelDescriptor {
template <class T>
struct Mpi_typemap;
template <>
struct Mpi_typemap<double> {
  static MPI_Datatype type() { return MPI_DOUBLE; }
};
template <>
struct Mpi_typemap<long long> {
  static MPI_Datatype type() { return MPI_LONG_LONG; }
};
}  // namespace ParallelDescriptor

// ---------------------------------------------------------------- plotfile I/O
// Plotfile parsing lives in un-vendored AMReX.  The shim does not parse files:
// a test registers the CONTENT of a plotfile (what PlotFileData would expose)
// under a path name, and PlotFileData(path) hands it to the reference's own
// VolumeRenderer::loadPlotFileGeometry / run().  Unregistered paths throw, as
// before.  get() deals box b to thread-rank b % nranks.
namespace shim {
struct PlotLevel {
  std::vector<Box> boxes;
  std::vector<std::vector<Real>> data;  // per box: ncomp * cells, component-major, i fastest
  Array<Real, 3> cellSize{1.0, 1.0, 1.0};
};
struct PlotContent {
  Vector<std::string> names;
  int spaceDim = 3;
  Array<Real, 3> probLo{0.0, 0.0, 0.0};
  std::vector<PlotLevel> levels;
  int refRatio = 2;
};
inline std::map<std::string, PlotContent>& plotfiles() {
  static std::map<std::string, PlotContent> registry;
  return registry;
}
}  // namespace shim

class PlotFileData {
 public:
  explicit PlotFileData(const std::string& path) {
    const auto it = shim::plotfiles().find(path);
    if (it == shim::plotfiles().end())
      throw std::runtime_error(
          "oracle shim: amrex::PlotFileData is unavailable (AMReX not vendored) and '" + path +
          "' was not registered");
    m_content = &it->second;
  }
  int spaceDim() const { return m_content->spaceDim; }
  const Vector<std::string>& varNames() const { return m_content->names; }
  int finestLevel() const { return static_cast<int>(m_content->levels.size()) - 1; }
  MultiFab get(int level, const std::string& name) {
    const shim::PlotLevel& L = m_content->levels.at(static_cast<std::size_t>(level));
    int comp = -1;
    for (std::size_t c = 0; c < m_content->names.size(); ++c)
      if (m_content->names[c] == name) comp = static_cast<int>(c);
    if (comp < 0) throw std::runtime_error("oracle shim: no such plotfile variable");
    int rank = 0, nranks = 1;
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);
    MPI_Comm_size(MPI_COMM_WORLD, &nranks);
    std::vector<int> owners(L.boxes.size());
    for (std::size_t b = 0; b < owners.size(); ++b) owners[b] = static_cast<int>(b) % nranks;
    MultiFab mf(L.boxes, owners, 1, rank);
    for (std::size_t b = 0; b < owners.size(); ++b) {
      if (owners[b] != rank) continue;
      const std::size_t cells = static_cast<std::size_t>(L.boxes[b].numPts());
      std::copy(L.data[b].begin() + static_cast<std::ptrdiff_t>(cells * comp),
                L.data[b].begin() + static_cast<std::ptrdiff_t>(cells * (comp + 1)),
                mf.dataPtr(static_cast<int>(b)));
    }
    return mf;
  }
  IntVect refRatio(int) const { return IntVect(m_content->refRatio); }
  Array<Real, 3> probLo() const { return m_content->probLo; }
  Array<Real, 3> cellSize(int level) const {
    return m_content->levels.at(static_cast<std::size_t>(level)).cellSize;
  }

 private:
  const shim::PlotContent* m_content = nullptr;
};

// convexify removes coarse cells covered by finer levels and re-cuts boxes; its
// source is un-vendored AMReX.  With ONE level there is nothing to remove, so
// the shim returns a deep copy with the same boxes (parity unpinned: real
// AMReX may also re-chop a single level's BoxArray).  For more levels a test
// may install the cuts it wants the reference to see (shim::convex_override:
// per level a list of boxes, each inside one source box); the shim then builds
// the MultiFabs on those boxes and copies the cells over, every piece staying
// on the rank of its source FAB.  Without an override multi-level input throws.
namespace shim {
struct ConvexOverride {
  bool active = false;
  std::vector<std::vector<Box>> boxes;  // [level][piece]
};
inline ConvexOverride& convex_override() {
  static ConvexOverride instance;
  return instance;
}
}  // namespace shim

inline Vector<MultiFab> convexify(const Vector<MultiFab const*>& levels,
                                  const Vector<IntVect>&) {
  const shim::ConvexOverride& cuts = shim::convex_override();
  Vector<MultiFab> out;
  if (!cuts.active) {
    if (levels.size() != 1) {
      throw std::runtime_error(
          "oracle shim: amrex::convexify for >1 level is unavailable (AMReX not "
          "vendored); pass already-convex boxes");
    }
    out.push_back(levels[0]->deepCopy());
    return out;
  }
  if (cuts.boxes.size() < levels.size())
    throw std::runtime_error("oracle shim: convexify override has too few levels");
  for (std::size_t l = 0; l < levels.size(); ++l) {
    const MultiFab& src = *levels[l];
    const std::vector<Box>& pieces = cuts.boxes[l];
    std::vector<int> owners(pieces.size(), 0), source(pieces.size(), -1);
    for (std::size_t p = 0; p < pieces.size(); ++p) {
      for (int g = 0; g < src.size() && source[p] < 0; ++g) {
        bool inside = true;
        for (int d = 0; d < 3; ++d)
          inside = inside && src.box(g).smallEnd()[d] <= pieces[p].smallEnd()[d] &&
                   pieces[p].bigEnd()[d] <= src.box(g).bigEnd()[d];
        if (inside) source[p] = g;
      }
      if (source[p] < 0) throw std::runtime_error("oracle shim: convexify piece outside every box");
      owners[p] = src.owner(source[p]);
    }
    MultiFab mf(pieces, owners, src.nComp(), src.myRank());
    for (std::size_t p = 0; p < pieces.size(); ++p) {
      if (owners[p] != src.myRank()) continue;
      const Box& sb = src.box(source[p]);
      const Box& pb = pieces[p];
      const Real* from = src.dataPtr(source[p]);
      Real* to = mf.dataPtr(static_cast<int>(p));
      const Long scells = sb.numPts(), pcells = pb.numPts();
      for (int n = 0; n < src.nComp(); ++n)
        for (int k = pb.smallEnd()[2]; k <= pb.bigEnd()[2]; ++k)
          for (int j = pb.smallEnd()[1]; j <= pb.bigEnd()[1]; ++j)
            for (int i = pb.smallEnd()[0]; i <= pb.bigEnd()[0]; ++i) {
              const Long s = (i - sb.smallEnd()[0]) +
                             static_cast<Long>(sb.length(0)) *
                                 ((j - sb.smallEnd()[1]) +
                                  static_cast<Long>(sb.length(1)) * (k - sb.smallEnd()[2]));
              const Long t = (i - pb.smallEnd()[0]) +
                             static_cast<Long>(pb.length(0)) *
                                 ((j - pb.smallEnd()[1]) +
                                  static_cast<Long>(pb.length(1)) * (k - pb.smallEnd()[2]));
              to[n * pcells + t] = from[n * scells + s];
            }
    }
    out.push_back(std::move(mf));
  }
  return out;
}

// Geometry: only what VolumeRendererApi.cpp:102-110 reads.
class Geometry {
 public:
  Geometry() = default;
  Geometry(const Array<Real, 3>& probLo, const Array<Real, 3>& cellSize)
      : m_lo(probLo), m_dx(cellSize) {}
  Array<Real, 3> ProbLoArray() const { return m_lo; }
  Array<Real, 3> CellSizeArray() const { return m_dx; }

 private:
  Array<Real, 3> m_lo{0.0, 0.0, 0.0};
  Array<Real, 3> m_dx{1.0, 1.0, 1.0};
};

namespace FileSystem {
inline bool Exists(const std::string& path) { return shim::plotfiles().count(path) != 0; }
}  // namespace FileSystem

}  // namespace amrex

#endif  // AMRVR_ORACLE_AMREX_SHIM_HPP
