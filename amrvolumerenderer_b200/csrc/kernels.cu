// kernels.cu -- sm_100a kernels of the hot path.  Compiled with -fmad=false: the
// reference's default x86-64 build never contracts a*b+c, and position, cell
// index and blend arithmetic must round the same way.
// Division and sqrt stay IEEE (-prec-div/-prec-sqrt defaults, no fast-math).
//
//   brick_stats_kernel          K2  min / max / min-positive / finite count (HBM: 8 B/cell)
//   brick_histogram_kernel      f4  histogram of the transformed scalar (HBM: 8 B/cell)
//   brick_quantise_batch_kernel K5  FAB doubles -> colour-table index bytes, a group of
//                                   bricks per launch (HBM: 9 B/cell)
//   raymarch_bricks_kernel      K1  all (box, 16x16 pixel tile) pairs of a rank, one launch,
//                                   persistent CTAs pulling tiles from a counter
//   fold_bin_kernel             K6  per 8-row band: the ordered list of layers touching it
//   fold_layers_kernel          K6  per-pixel run-structured depth-swap 'over' of a rank's rows; the
//                                   fragments of those rows lie in the rank's OWN pool (the march of
//                                   whichever GPU took the tile stored them there over NVLink);
//                                   <true>: overlay + RGB8 epilogue (into the root's frame over NVLink)
//   pack_slices_kernel          C6  row slices into per-destination send buffers (NCCL path)
//   push_slices_kernel          C6  the same slices stored straight into the destination GPU's receive pool
//   peer_wait_kernel                closes a sharded frame: waits for every GPU's fold flag, time-out
//   peer_barrier_kernel             flag barrier across the GPUs of one box, with time-out (set-up paths)
//   peer_copy_kernel                load/store copy between peers (amrvr_peer_bandwidth, the NVLink probe)
//   post_piece_kernel               overlay + bytes of a piece into the root's RGB8 frame
//   downsample_kernel           K8  box filter
//   overlay_kernel                  wireframe coverage on the float frame
//   quantize_rgb8_kernel            float -> byte, rows flipped
//   order_rank_kernel           K7  single-CTA LSD radix sort of box depth keys
//   order_pairs_kernel          K7  all-pairs face-adjacency test
//   order_count_kernel          K7  per-box count of boxes in front (priority of the Kahn peel)
#include <cuda_runtime.h>

#include <algorithm>
#include <cfloat>
#include <climits>
#include <cmath>

#include "internal.hpp"

namespace amrvr {

namespace {

constexpr int kTile = 16;          // march tile edge in pixels
constexpr int kMarchThreads = 256; // 8 warps, each an 8x4 pixel patch

__device__ __forceinline__ float clamp01(float v) {
  return (v < 0.0f) ? 0.0f : ((v > 1.0f) ? 1.0f : v);
}

// ------------------------------------------------------------------ K2 / K5

struct ValuePath {  // frame constants of the per-cell value path
  int log_scale, normalize;
  double positive_floor, norm_min, inv_norm_span;
  float range_min, inv_range, clip_start;
  int apply_clip;
};

// reference: Common/VolumeTypes.hpp:33-67 + Common/VolumePainter.cpp:75-105, 870-883
__device__ __forceinline__ unsigned char cell_to_index(double raw, const ValuePath& p) {
  double v = isfinite(raw) ? raw : 0.0;
  if (p.log_scale) {
    if (!(v > 0.0) || v < p.positive_floor) v = p.positive_floor;
    v = log(v);
  }
  if (p.normalize) {
    v = (v - p.norm_min) * p.inv_norm_span;
    v = (v < 0.0) ? 0.0 : ((v > 1.0) ? 1.0 : v);
  }
  float s = static_cast<float>(v);
  if (p.apply_clip) {  // soft roll-off between clip_start and 1
    const float end = (p.clip_start < 1.0f) ? 1.0f : p.clip_start;
    float c = (s < 0.0f) ? 0.0f : ((s > end) ? end : s);
    if ((end > p.clip_start + 1e-5f) && (c > p.clip_start)) {
      if (!(c < end)) {
        c = end;
      } else {
        const float n = (c - p.clip_start) / (end - p.clip_start);
        const float smooth = n + n * n - n * n * n;
        c = p.clip_start + (end - p.clip_start) * smooth;
      }
    }
    s = c;
  }
  float t = (s - p.range_min) * p.inv_range;
  t = (t < 0.0f) ? 0.0f : t;
  t = (t > 1.0f) ? 1.0f : t;
  int idx = static_cast<int>(t * 255.0f);
  idx = (idx < 0) ? 0 : ((idx > 255) ? 255 : idx);
  return static_cast<unsigned char>(idx);
}

// One thread per 4 consecutive cells of a row (rows are nx cells, i fastest);
// blockIdx.y picks the brick.  A 64^3 brick is 2 MB -- alone it cannot fill the
// machine for longer than a launch takes, so the upload path quantises a whole
// group of freshly copied FABs with one launch.
__global__ void __launch_bounds__(256)
brick_quantise_batch_kernel(const QuantJob* __restrict__ jobs, ValuePath p) {
  const QuantJob J = jobs[blockIdx.y];
  const int nx = J.nx, ny = J.ny;
  if (J.jstride == nx && J.kstride == static_cast<long long>(nx) * ny &&
      ((reinterpret_cast<uintptr_t>(J.src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(J.dst) & 3) == 0)) {
    // dense FAB (no ghost cells): the brick is one contiguous run of cells, no
    // row arithmetic at all -- four cells per thread per step, two steps in flight
    const long long ncells = static_cast<long long>(nx) * ny * J.nz;
    const long long nq = ncells >> 2;
    const long long stride = static_cast<long long>(gridDim.x) * blockDim.x;
    const double2* src2 = reinterpret_cast<const double2*>(J.src);
    uchar4* dst4 = reinterpret_cast<uchar4*>(J.dst);
    long long q = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
    for (; q + stride < nq; q += 2 * stride) {
      const double2 a = __ldcs(src2 + 2 * q), b = __ldcs(src2 + 2 * q + 1);
      const double2 c = __ldcs(src2 + 2 * (q + stride)), d = __ldcs(src2 + 2 * (q + stride) + 1);
      dst4[q] = make_uchar4(cell_to_index(a.x, p), cell_to_index(a.y, p), cell_to_index(b.x, p),
                            cell_to_index(b.y, p));
      dst4[q + stride] = make_uchar4(cell_to_index(c.x, p), cell_to_index(c.y, p),
                                     cell_to_index(d.x, p), cell_to_index(d.y, p));
    }
    for (; q < nq; q += stride) {
      const double2 a = __ldcs(src2 + 2 * q), b = __ldcs(src2 + 2 * q + 1);
      dst4[q] = make_uchar4(cell_to_index(a.x, p), cell_to_index(a.y, p), cell_to_index(b.x, p),
                            cell_to_index(b.y, p));
    }
    if (blockIdx.x == 0 && threadIdx.x < (ncells & 3))
      J.dst[4 * nq + threadIdx.x] = cell_to_index(J.src[4 * nq + threadIdx.x], p);
    return;
  }
  const int quads_per_row = (nx + 3) >> 2;
  const long long nquads = static_cast<long long>(quads_per_row) * ny * J.nz;
  for (long long q = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; q < nquads;
       q += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int qi = static_cast<int>(q % quads_per_row);
    const long long row = q / quads_per_row;
    const int j = static_cast<int>(row % ny);
    const int k = static_cast<int>(row / ny);
    const double* src = J.src + j * J.jstride + k * J.kstride + 4 * qi;
    unsigned char* dst = J.dst + (static_cast<long long>(k) * ny + j) * nx + 4 * qi;
    const int left = nx - 4 * qi;
    if (left >= 4 && ((reinterpret_cast<uintptr_t>(src) & 15) == 0) &&
        ((reinterpret_cast<uintptr_t>(dst) & 3) == 0)) {
      const double2 a = __ldcs(reinterpret_cast<const double2*>(src));
      const double2 b = __ldcs(reinterpret_cast<const double2*>(src) + 1);
      uchar4 o;
      o.x = cell_to_index(a.x, p);
      o.y = cell_to_index(a.y, p);
      o.z = cell_to_index(b.x, p);
      o.w = cell_to_index(b.y, p);
      *reinterpret_cast<uchar4*>(dst) = o;
    } else {
      for (int c = 0; c < left && c < 4; ++c) dst[c] = cell_to_index(src[c], p);
    }
  }
}

struct Stats4 {
  double lo, hi, lo_pos;
  long long finite;
};

__device__ __forceinline__ void stats_merge(Stats4& a, const Stats4& b) {
  a.lo = fmin(a.lo, b.lo);
  a.hi = fmax(a.hi, b.hi);
  a.lo_pos = fmin(a.lo_pos, b.lo_pos);
  a.finite += b.finite;
}

__device__ __forceinline__ Stats4 stats_block_reduce(Stats4 s) {
  __shared__ Stats4 part[32];
  for (int off = 16; off > 0; off >>= 1) {
    Stats4 o;
    o.lo = __shfl_down_sync(0xffffffffu, s.lo, off);
    o.hi = __shfl_down_sync(0xffffffffu, s.hi, off);
    o.lo_pos = __shfl_down_sync(0xffffffffu, s.lo_pos, off);
    o.finite = __shfl_down_sync(0xffffffffu, s.finite, off);
    stats_merge(s, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) part[warp] = s;
  __syncthreads();
  if (warp == 0) {
    const int nwarps = (blockDim.x + 31) >> 5;
    Stats4 t = (lane < nwarps) ? part[lane] : Stats4{INFINITY, -INFINITY, INFINITY, 0};
    for (int off = 16; off > 0; off >>= 1) {
      Stats4 o;
      o.lo = __shfl_down_sync(0xffffffffu, t.lo, off);
      o.hi = __shfl_down_sync(0xffffffffu, t.hi, off);
      o.lo_pos = __shfl_down_sync(0xffffffffu, t.lo_pos, off);
      o.finite = __shfl_down_sync(0xffffffffu, t.finite, off);
      stats_merge(t, o);
    }
    s = t;
  }
  return s;
}

// reference: VolumeRenderer/SceneBuilder.cpp:53-94 (non-finite cells are skipped)
__global__ void __launch_bounds__(256)
brick_stats_kernel(const double* __restrict__ data, long long jstride, long long kstride, int nx,
                   int ny, int nz, double* __restrict__ partials) {
  Stats4 s{INFINITY, -INFINITY, INFINITY, 0};
  const long long ncells = static_cast<long long>(nx) * ny * nz;
  for (long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; c < ncells;
       c += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int i = static_cast<int>(c % nx);
    const long long row = c / nx;
    const int j = static_cast<int>(row % ny);
    const int k = static_cast<int>(row / ny);
    const double v = __ldcs(data + i + j * jstride + k * kstride);
    if (isfinite(v)) {
      s.lo = fmin(s.lo, v);
      s.hi = fmax(s.hi, v);
      if (v > 0.0) s.lo_pos = fmin(s.lo_pos, v);
      ++s.finite;
    }
  }
  s = stats_block_reduce(s);
  if (threadIdx.x == 0) {
    double* o = partials + 4 * blockIdx.x;
    o[0] = s.lo;
    o[1] = s.hi;
    o[2] = s.lo_pos;
    o[3] = __longlong_as_double(s.finite);
  }
}

__global__ void __launch_bounds__(256)
brick_stats_final_kernel(const double* __restrict__ partials, int nparts, double* __restrict__ out4) {
  Stats4 s{INFINITY, -INFINITY, INFINITY, 0};
  for (int p = threadIdx.x; p < nparts; p += blockDim.x) {
    Stats4 o{partials[4 * p], partials[4 * p + 1], partials[4 * p + 2],
             __double_as_longlong(partials[4 * p + 3])};
    stats_merge(s, o);
  }
  s = stats_block_reduce(s);
  if (threadIdx.x == 0) {
    out4[0] = s.lo;
    out4[1] = s.hi;
    out4[2] = s.lo_pos;
    out4[3] = __longlong_as_double(s.finite);
  }
}

// f4.  reference: VolumeRenderer/SceneBuilder.cpp:445-577 (per-cell body :503-539).
// Per-CTA histogram in shared memory (bins <= kHistSmemBins), flushed with one
// 64-bit atomic per non-empty bin; larger bin counts go straight to global atomics.
constexpr int kHistSmemBins = 8192;

__device__ __forceinline__ int histogram_bin(double raw, const ValuePath& p, float range_min,
                                             float range_max, float inverse_width, int bin_count) {
  double v = isfinite(raw) ? raw : 0.0;  // applyScalarTransform, VolumeTypes.hpp:33-67
  if (p.log_scale) {
    if (!(v > 0.0) || v < p.positive_floor) v = p.positive_floor;
    v = log(v);
  }
  if (p.normalize) {
    v = (v - p.norm_min) * p.inv_norm_span;
    v = (v < 0.0) ? 0.0 : ((v > 1.0) ? 1.0 : v);
  }
  float value = static_cast<float>(v);
  if (value < range_min) value = range_min;
  else if (value > range_max) value = range_max;
  float normalized = (value - range_min) * inverse_width;
  if (normalized < 0.0f) normalized = 0.0f;
  else if (normalized > 1.0f) normalized = 1.0f;
  int index = static_cast<int>(normalized * static_cast<float>(bin_count));
  if (index >= bin_count) index = bin_count - 1;
  else if (index < 0) index = 0;
  return index;
}

__global__ void __launch_bounds__(256)
brick_histogram_kernel(const double* __restrict__ data, long long jstride, long long kstride, int nx,
                       int ny, int nz, const __grid_constant__ ValuePath p, float range_min,
                       float range_max, float inverse_width, int bin_count,
                       unsigned long long* __restrict__ counts) {
  extern __shared__ unsigned s_bins[];
  const bool in_smem = bin_count <= kHistSmemBins;
  if (in_smem) {
    for (int b = threadIdx.x; b < bin_count; b += blockDim.x) s_bins[b] = 0u;
    __syncthreads();
  }
  const long long ncells = static_cast<long long>(nx) * ny * nz;
  for (long long c = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x; c < ncells;
       c += static_cast<long long>(gridDim.x) * blockDim.x) {
    const int i = static_cast<int>(c % nx);
    const long long row = c / nx;
    const int j = static_cast<int>(row % ny);
    const int k = static_cast<int>(row / ny);
    const double v = __ldcs(data + i + j * jstride + k * kstride);
    const int bin = histogram_bin(v, p, range_min, range_max, inverse_width, bin_count);
    if (in_smem) atomicAdd(&s_bins[bin], 1u);  // < 2^32 cells per CTA by the grid-stride bound
    else atomicAdd(&counts[bin], 1ULL);
  }
  if (in_smem) {
    __syncthreads();
    for (int b = threadIdx.x; b < bin_count; b += blockDim.x)
      if (s_bins[b] != 0u) atomicAdd(&counts[b], static_cast<unsigned long long>(s_bins[b]));
  }
}

// ------------------------------------------------------------------------ K1

// One CTA = one 16x16 pixel tile of one box's screen footprint; one thread = one
// (ray, box) pair, marched with the reference's serial float recurrence
// distance += step (Common/VolumePainter.cpp:735-922).  Cooperation is across
// rays only: the transfer table is staged once per CTA in shared memory, warps
// are 8x4 pixel patches so the bytes a warp gathers from the brick share L1
// lines, and stores are whole 128 B row segments.
//
// Every sample is evaluated with the reference's arithmetic (no FMA, true
// division unless the cell size is a power of two).  What is removed is work
// whose outcome is known in advance:
//   * the table is stored kTableCopies times, entry e of copy c at float4 slot
//     e*kTableCopies + c, and lane l reads copy l%8: the eight lanes of one
//     LDS.128 phase then always hit eight different 16 B bank groups;
//   * floor(v) for v >= 0 (the sample is inside the box) is a round-down
//     multiply by 2^-149: the subnormal result's bit pattern IS the integer,
//     exact for v < 2^23.  When the cell size is a power of two, (p - lo)/cell
//     is itself exact, so one round-down multiply by 2^-149/cell gives the same
//     integer.  This keeps the loop off the quarter-rate conversion pipe;
//   * between t_safe0 and t_safe1 -- the slab interval of the box shrunk by
//     MarchBox::margin, a bound on every rounding error of the position -- the
//     sample is provably inside the box and its cell provably in range, so the
//     six inside tests and the three index clamps are skipped there.  The few
//     samples before and after that interval run the checked body.

constexpr int kTableCopies = 8;
constexpr float kTinyOne = 1.401298464324817e-45f;  // 2^-149: bits(floor(v) * 2^-149) == floor(v)

__device__ __forceinline__ unsigned smem_u32(const void* p) {
  return static_cast<unsigned>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ float4 lds_f4(unsigned addr) {
  float4 v;
  asm("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
  return v;
}

__device__ __forceinline__ float pin_reg(float v) {
  asm volatile("mov.f32 %0, %1;" : "=f"(v) : "f"(v));
  return v;
}
__device__ __forceinline__ const unsigned char* pin_ptr(const unsigned char* p) {
  asm volatile("mov.b64 %0, %1;" : "=l"(p) : "l"(p));
  return p;
}

struct RayOut {
  float r, g, b, a, depth;
  unsigned nsamples;
};

template <bool kCount, bool kRecip>
__device__ __forceinline__ RayOut march_ray(const MarchFrame& f, const MarchBox& B,
                                            const unsigned char* __restrict__ bricks,
                                            unsigned table_addr, int px, int py, bool valid) {
  const float lox = B.lo[0], loy = B.lo[1], loz = B.lo[2];
  const float hix = B.hi[0], hiy = B.hi[1], hiz = B.hi[2];
  // held in registers: re-reading the constant bank costs an issue slot per use
  const float ex = pin_reg(f.eye[0]), ey = pin_reg(f.eye[1]), ez = pin_reg(f.eye[2]);

  // ray through the pixel centre (:738-766)
  const float ndc_x = (static_cast<float>(px) + 0.5f) * f.inv_w * 2.0f - 1.0f;
  const float ndc_y = (static_cast<float>(py) + 0.5f) * f.inv_h * 2.0f - 1.0f;
  const float plane_x = ndc_x * f.tan_half * f.aspect;
  const float plane_y = ndc_y * f.tan_half;
  float dx = f.fwd[0] + plane_x * f.right[0] + plane_y * f.up[0];
  float dy = f.fwd[1] + plane_x * f.right[1] + plane_y * f.up[1];
  float dz = f.fwd[2] + plane_x * f.right[2] + plane_y * f.up[2];
  const float len_sq = dx * dx + dy * dy + dz * dz;
  // host rsqrt of the reference is 1.0/sqrt(x) evaluated in double on a float
  const float len = (len_sq > 0.0f)
                        ? (1.0f / static_cast<float>(1.0 / static_cast<double>(sqrtf(len_sq))))
                        : 0.0f;
  if (len > 0.0f) {
    const float inv = 1.0f / len;
    dx *= inv;
    dy *= inv;
    dz *= inv;
  }

  // slab test (:768-809) and, with the same reciprocals, the slab interval of
  // the box shrunk by B.margin on every face
  float tmin = -INFINITY, tmax = INFINITY;
  float ts0 = -INFINITY, ts1 = INFINITY;
  const float margin = B.margin;
#define AMRVR_SLAB(o, d, lo, hi)                                        \
  if (fabsf(d) < 1e-8f) {                                               \
    if (o < lo || o > hi) {                                             \
      tmin = INFINITY;                                                  \
      tmax = -INFINITY;                                                 \
    }                                                                   \
    if (!(o >= lo + margin && o <= hi - margin)) {                      \
      ts0 = INFINITY;                                                   \
      ts1 = -INFINITY;                                                  \
    }                                                                   \
  } else {                                                              \
    const float inv = 1.0f / d;                                         \
    float t1 = (lo - o) * inv, t2 = (hi - o) * inv;                     \
    if (t1 > t2) {                                                      \
      const float t = t1;                                               \
      t1 = t2;                                                          \
      t2 = t;                                                           \
    }                                                                   \
    tmin = (tmin > t1) ? tmin : t1;                                     \
    tmax = (tmax < t2) ? tmax : t2;                                     \
    const float s_near = (((d > 0.0f) ? lo + margin : hi - margin) - o) * inv; \
    const float s_far = (((d > 0.0f) ? hi - margin : lo + margin) - o) * inv;  \
    ts0 = fmaxf(ts0, s_near);                                           \
    ts1 = fminf(ts1, s_far);                                            \
  }
  AMRVR_SLAB(ex, dx, lox, hix)
  AMRVR_SLAB(ey, dy, loy, hiy)
  AMRVR_SLAB(ez, dz, loz, hiz)
#undef AMRVR_SLAB
  if (!(margin > 0.0f) || !(ts0 <= ts1)) {  // no provably-inside stretch on this ray
    ts0 = INFINITY;
    ts1 = -INFINITY;
  }

  RayOut out{0.0f, 0.0f, 0.0f, 0.0f, INFINITY, 0u};
  if (!(valid && tmax >= tmin)) return out;

  float acc_r = 0.0f, acc_g = 0.0f, acc_b = 0.0f, acc_a = 0.0f;
  unsigned nsamples = 0;
  const float step = B.step;
  float distance = tmin + B.eps;
  if (distance < 0.0f) distance = B.eps;
  float pxs = ex + dx * distance, pys = ey + dy * distance, pzs = ez + dz * distance;

  const unsigned nx = static_cast<unsigned>(B.n[0]), ny = static_cast<unsigned>(B.n[1]);
  const unsigned ix_max = nx - 1u, iy_max = ny - 1u, iz_max = static_cast<unsigned>(B.n[2]) - 1u;
  const unsigned char* brick = pin_ptr(bricks + B.brick_off);
  const float cx = kRecip ? B.inv_cell_tiny[0] : B.cell[0];
  const float cy = kRecip ? B.inv_cell_tiny[1] : B.cell[1];
  const float cz = kRecip ? B.inv_cell_tiny[2] : B.cell[2];

  auto advance = [&]() {
    distance += step;
    pxs = ex + dx * distance;
    pys = ey + dy * distance;
    pzs = ez + dz * distance;
  };
  auto fetch = [&](unsigned cell) {
    // spelled out so that the byte arrives zero-extended
    unsigned idx;
    asm volatile("ld.global.nc.u8 %0, [%1];" : "=r"(idx) : "l"(brick + cell));
    return idx;
  };
  // predicated gather: no access when !wanted, and idx then keeps its previous
  // (valid) value, so a table read through it stays in range
  auto fetch_if = [&](unsigned& idx, unsigned cell, bool wanted) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %2, 0;\n\t"
        "@p ld.global.nc.u8 %0, [%1];\n\t}"
        : "+r"(idx)
        : "l"(brick + cell), "r"(static_cast<int>(wanted)));
  };
  auto table_entry = [&](unsigned idx) {
    unsigned slot;  // the table address is one multiply-add
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(slot) : "r"(idx), "r"(16u * kTableCopies), "r"(table_addr));
    return lds_f4(slot);
  };
  auto blend = [&](const float4& e) {  // :884-894
    const float a = e.w * (1.0f - acc_a);
    acc_r += e.x * a;
    acc_g += e.y * a;
    acc_b += e.z * a;
    acc_a += a;
    if (kCount) ++nsamples;
  };
  auto shade = [&](unsigned jk, unsigned ib) { blend(table_entry(fetch(jk * nx + ib))); };  // jk = k*ny + j
  auto cell_of = [&](float p, float lo, float c) {
    // reference: floor((p - lo) / cell), :846-848
    return __float_as_uint(kRecip ? __fmul_rd(p - lo, c) : __fmul_rd((p - lo) / c, kTinyOne));
  };
  auto checked = [&]() {
    if (!(pxs < lox || pxs > hix || pys < loy || pys > hiy || pzs < loz || pzs > hiz)) {
      // inside: (p - lo) >= 0, so floor() >= 0 and only the upper clamp can bind
      const unsigned ib = min(cell_of(pxs, lox, cx), ix_max);
      const unsigned jb = min(cell_of(pys, loy, cy), iy_max);
      const unsigned kb = min(cell_of(pzs, loz, cz), iz_max);
      shade(kb * ny + jb, ib);
    }
    advance();
  };

  // The reference first steps forward until the sample is inside (:827-835) and
  // then marches (:837-900); skipping outside samples inside one loop is the
  // same sequence of samples.
  const float t_head = fminf(ts0, tmax);
  while (distance < t_head && acc_a < 1.0f) checked();
  // Provably-inside stretch, software-pipelined two deep.  The position of a
  // sample is a function of `distance` alone, never of the shading, so while
  // sample A is blended the table entry of sample B is being read from shared
  // memory and the brick byte of sample C is being gathered from L1/L2.  The
  // sequence of samples and every operation on them are those of the plain
  // loop; speculative requests are either predicated off or go to in-range
  // addresses and their values are dropped when the ray saturates first.
  const float t_body = fminf(ts1, tmax);
  if (distance < t_body && acc_a < 1.0f) {
    auto cell_at = [&](float d) {
      const float qx = ex + dx * d, qy = ey + dy * d, qz = ez + dz * d;
      return (cell_of(qz, loz, cz) * ny + cell_of(qy, loy, cy)) * nx + cell_of(qx, lox, cx);
    };
    unsigned idx_a = 0u, idx_b = 0u, idx_c = 0u;
    fetch_if(idx_a, cell_at(distance), true);
    float d_b = distance + step, d_c = 0.0f;
    bool more_b = d_b < t_body, more_c = false;
    fetch_if(idx_b, cell_at(d_b), more_b);
    float4 e_a = table_entry(idx_a), e_b;
    // One stage: request the byte two samples ahead, read the table entry one
    // sample ahead, blend the current sample.  Returns true when the ray leaves
    // the loop; d_next is then the plain loop's final `distance`.  The loop body
    // is two stages with the roles of the register sets swapped, so nothing is
    // moved between them.
    auto stage = [&](const float4& e_cur, float4& e_next, unsigned idx_next, unsigned& idx_after,
                     float d_next, float& d_after, bool more_next, bool& more_after) {
      d_after = d_next + step;
      more_after = more_next && (d_after < t_body);
      fetch_if(idx_after, cell_at(d_after), more_after);
      e_next = table_entry(idx_next);
      blend(e_cur);
      return !more_next || !(acc_a < 1.0f);
    };
    float d_exit;
    for (;;) {
      if (stage(e_a, e_b, idx_b, idx_c, d_b, d_c, more_b, more_c)) {
        d_exit = d_b;
        break;
      }
      if (stage(e_b, e_a, idx_c, idx_b, d_c, d_b, more_c, more_b)) {
        d_exit = d_c;
        break;
      }
    }
    d_b = d_exit;
    distance = d_b;  // the plain loop leaves the state one step past the last shaded sample
    pxs = ex + dx * distance;
    pys = ey + dy * distance;
    pzs = ez + dz * distance;
  }
  while (distance < tmax && acc_a < 1.0f) checked();

  acc_r = (acc_r > 1.0f) ? 1.0f : acc_r;
  acc_g = (acc_g > 1.0f) ? 1.0f : acc_g;
  acc_b = (acc_b > 1.0f) ? 1.0f : acc_b;
  acc_a = (acc_a > 1.0f) ? 1.0f : acc_a;
  float depth = INFINITY;
  if (acc_a > 0.0f) {  // entry depth along the view axis (:912-921)
    const float qx = ex + dx * tmin, qy = ey + dy * tmin, qz = ez + dz * tmin;
    depth = (qx - ex) * f.fwd[0] + (qy - ey) * f.fwd[1] + (qz - ez) * f.fwd[2];
  }
  out.r = acc_r;
  out.g = acc_g;
  out.b = acc_b;
  out.a = acc_a;
  out.depth = depth;
  out.nsamples = nsamples;
  return out;
}

// Persistent CTAs: the grid is (SM count x resident CTAs per SM) and every CTA
// pulls 16x16 tiles from a queue counter until none is left.  Tiles of one item
// are consecutive and items are sorted by transfer table, so the box parameters
// in shared memory are refilled per tile but the table only when the launch
// moves on to the next AMR level.
//
// Queues (MarchQueues): one for a single-GPU march.  A frame sharded over the
// GPUs of a box has one queue per GPU -- the tiles of the 16-row frame strips
// that GPU folds (QueueItem, internal.hpp); every GPU holds every brick, so any
// GPU can march any tile.  A CTA drains its own GPU's queue first (the fragments
// stay in local HBM) and then takes tiles from the other queues, heaviest first,
// storing those fragments straight into the pool of the GPU that folds them --
// posted NVLink stores from an issue-bound kernel, the exchange of
// DirectSendBase::PostSends/PostReceives (DirectSendBase.cpp:76-177) overlapped
// with the march tile by tile.  The claim of the next tile is an atomic on the
// owning GPU's counter, issued by the first warp that is through with the current
// tile, so its round trip (NVLink for a remote queue) and the look-up of the
// tile's item hide behind the warps still marching.  "hints" keep CTAs from
// probing queues another CTA already found drained.  The last CTA to finish
// raises this rank's flag on every GPU: that is what the folds wait for.
#ifndef AMRVR_MARCH_MIN_BLOCKS
#define AMRVR_MARCH_MIN_BLOCKS 4
#endif
#ifndef AMRVR_MARCH_TIMELINE
#define AMRVR_MARCH_TIMELINE 1  // thread 0 of every CTA times itself on %globaltimer (amrvr_frame_stats::march_*_ns)
#endif

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ unsigned ld_volatile_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void st_volatile_u32(unsigned* p, unsigned v) {
  asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

template <bool kCount>
__global__ void __launch_bounds__(kMarchThreads, AMRVR_MARCH_MIN_BLOCKS)
raymarch_bricks_kernel(const __grid_constant__ MarchFrame f, const __grid_constant__ MarchQueues Q,
                       const MarchBox* __restrict__ boxes,
                       const unsigned char* __restrict__ bricks,
                       const float4* __restrict__ tables) {
  __shared__ float4 s_table[AMRVR_COLOR_TABLE_SIZE * kTableCopies];
  __shared__ MarchBox s_box[2];                      // mailbox: the box parameters of the slot's tile
  __shared__ int s_tile[2], s_which[2], s_queue[2];  // mailbox: tile id, its box (item), its queue
  __shared__ int s_taken[2];                         // mailbox slot: a warp has taken on filling it

  // ---- thread 0 only; its bookkeeping lives in shared memory so that the march
  // loop keeps its registers.  The try order is the own queue, then steal_order[].
  __shared__ int s_cur;                          // position in the try order
  __shared__ unsigned long long s_pushed, s_stolen;
  auto queue_at = [&](int k) { return (k == 0) ? Q.mine : Q.steal_order[k - 1]; };
  auto mark_drained = [&](int q) {
    if (Q.nranks <= 0) return;
    if (ld_volatile_u32(Q.ctl[Q.me] + kCtlHints + q) == Q.epoch) return;
    if (q == Q.mine) {  // tell every GPU: nothing left to take here
      for (int r = 0; r < Q.nranks; ++r) st_volatile_u32(Q.ctl[r] + kCtlHints + q, Q.epoch);
    } else {
      st_volatile_u32(Q.ctl[Q.me] + kCtlHints + q, Q.epoch);
    }
  };
  auto known_drained = [&](int q) {
    return Q.nranks > 0 && q != Q.mine && ld_volatile_u32(Q.ctl[Q.me] + kCtlHints + q) == Q.epoch;
  };
  auto tile_of = [&](int q, unsigned t) {  // -1: the queue is drained
    return (t < static_cast<unsigned>(Q.tile_begin[q + 1] - Q.tile_begin[q]))
               ? Q.tile_begin[q] + static_cast<int>(t)
               : -1;
  };
  auto claim_blocking = [&]() {  // walks the try order; -1 when every queue is drained
    int cur = s_cur, tile = -1;
    while (cur < Q.nqueues) {
      const int q = queue_at(cur);
      if (!known_drained(q)) {
        tile = tile_of(q, atomicAdd(Q.counter[q], 1u));
        if (tile >= 0) break;
        mark_drained(q);
      }
      ++cur;
    }
    s_cur = cur;
    return tile;
  };
  auto publish = [&](int slot, int tile) {
    int lo = 0, q = 0;
    if (tile >= 0) {  // last item whose first tile is <= tile
      int hi = f.nboxes - 1;
      if (Q.queue_items != nullptr) {  // striped frame: the items of the queue the tile came from
        q = queue_at(s_cur);
        const QueueItem* items = Q.queue_items + static_cast<size_t>(q) * f.nboxes;
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (__ldg(&items[mid].tile_base) <= tile) lo = mid; else hi = mid - 1;
        }
      } else {
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (__ldg(&boxes[mid].tile_base) <= tile) lo = mid; else hi = mid - 1;
        }
      }
      if (s_cur > 0) ++s_stolen;
    }
    s_tile[slot] = tile;
    s_which[slot] = lo;
    s_queue[slot] = q;
  };

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned table_addr = smem_u32(s_table) + static_cast<unsigned>(lane % kTableCopies) * 16u;
  int have_table = -1;
  unsigned nsamples = 0;
  // thread 0 keeps the launch's time line (Q.stats[4..7]): CTA busy time, first start, last end, table staging
  __shared__ unsigned long long s_t0, s_table_ns, s_stage_t;  // in shared memory: the march loop keeps its registers

  // One warp fills mailbox slot `slot`: lane 0 claims a tile and looks its item up (an
  // atomic, possibly on another GPU, and a binary search: several dependent round trips),
  // then the warp copies the item's box parameters into the slot's MarchBox -- so the CTA
  // finds everything for its next tile in shared memory after ONE barrier.
  auto fill_slot = [&](int slot) {
    if (lane == 0) publish(slot, claim_blocking());
    __syncwarp();
    const int tile = *reinterpret_cast<volatile int*>(&s_tile[slot]);
    if (tile < 0) return;
    const int which = *reinterpret_cast<volatile int*>(&s_which[slot]);
    const int queue = *reinterpret_cast<volatile int*>(&s_queue[slot]);
    static_assert(sizeof(MarchBox) == 128 && offsetof(MarchBox, layer_off) == 112,
                  "words 28..31 (layer_off, dest, ry_first) and tile_base are the item's");
    constexpr int kTileBaseWord = offsetof(MarchBox, tile_base) / 4;
    const int* src = reinterpret_cast<const int*>(boxes + which);
    int* dst = reinterpret_cast<int*>(&s_box[slot]);
    if (Q.queue_items == nullptr) {
      dst[lane] = __ldg(src + lane);
    } else {
      // striped frame: the box, then what (queue, box) adds to it
      if (lane < 28 && lane != kTileBaseWord) dst[lane] = __ldg(src + lane);
      if (lane == 31) {
        const QueueItem* item = Q.queue_items + static_cast<size_t>(queue) * f.nboxes + which;
        s_box[slot].tile_base = __ldg(&item->tile_base);
        s_box[slot].ry_first = __ldg(&item->ry_first);
        s_box[slot].layer_off = __ldg(&item->layer_off);
        s_box[slot].dest = queue;
      }
    }
  };

  if (threadIdx.x == 0) {
    s_t0 = AMRVR_MARCH_TIMELINE ? global_timer_ns() : 0ull;
    s_table_ns = 0;
    s_cur = 0;
    s_pushed = s_stolen = 0;
    s_taken[0] = s_taken[1] = 0;
  }
  __syncthreads();
  if (warp == 0) fill_slot(0);
  for (int it = 0;; ++it) {
    __syncthreads();  // mailbox slot it&1 is filled; everyone is done with the previous tile
    const int tile = s_tile[it & 1];
    if (tile < 0) break;
    const MarchBox& B = s_box[it & 1];
    const int table = B.table;
    if (table != have_table) {  // uniform across the CTA; once per AMR level and queue
      if (AMRVR_MARCH_TIMELINE && threadIdx.x == 0) s_stage_t = global_timer_ns();
      // consecutive threads write consecutive 16 B slots: conflict-free stores
      const float4* src4 = tables + table * AMRVR_COLOR_TABLE_SIZE;
      for (int slot = threadIdx.x; slot < AMRVR_COLOR_TABLE_SIZE * kTableCopies; slot += kMarchThreads)
        s_table[slot] = __ldg(src4 + (slot / kTableCopies));
      have_table = table;
      __syncthreads();
      if (AMRVR_MARCH_TIMELINE && threadIdx.x == 0) s_table_ns += global_timer_ns() - s_stage_t;
    }

    // tile row ty of an item starts at frame row ry_first + ty * strip_stride (the next strip
    // of the same GPU when the frame is striped) and is row 16 * ty of the stored layer
    const int local_tile = tile - B.tile_base;
    const int tx = local_tile % B.tiles_x, ty = local_tile / B.tiles_x;
    const int x = tx * kTile + (warp & 1) * 8 + (lane & 7);
    const int y_in = (warp >> 1) * 4 + (lane >> 3);
    const int y = ty * kTile + y_in;                               // row of the layer
    const int frame_y = B.ry_first + ty * f.strip_stride + y_in;   // row of the frame
    const int rw = B.rw;
    // footprints are padded to 4-pixel column multiples: columns past the frame are not rays;
    // a strip may begin above the footprint or end below it
    const bool valid = x < rw && B.rx0 + x < f.width && frame_y >= B.ry0 && frame_y < B.ry0 + B.rh;

    const RayOut r = B.exact_recip
                         ? march_ray<kCount, true>(f, B, bricks, table_addr, B.rx0 + x, frame_y, valid)
                         : march_ray<kCount, false>(f, B, bricks, table_addr, B.rx0 + x, frame_y, valid);

    // host epilogue of the reference (:939-955)
    float4 out = make_float4(clamp01(r.r), clamp01(r.g), clamp01(r.b), clamp01(r.a));
    float depth = r.depth;
    if (!isfinite(depth) || out.w <= 0.0f) depth = INFINITY;
    if (valid) {
      float* row = Q.pool[B.dest] + B.layer_off + static_cast<size_t>(y) * rw * 5;
      reinterpret_cast<float4*>(row)[x] = out;
      row[4 * rw + x] = depth;
    }
    if (kCount) nsamples += r.nsamples;

    // The FIRST warp through with its patch fills the other mailbox slot while the rest are
    // still marching; with thread 0 doing it after its own ray, every tile whose slowest
    // patch was warp 0's paid for the claim and the look-up in full.
    int taker = 0;
    if (lane == 0) taker = (atomicExch(&s_taken[(it + 1) & 1], 1) == 0) ? 1 : 0;
    taker = __shfl_sync(0xffffffffu, taker, 0);
    if (taker) {
      if (lane == 0) {
        if (Q.nranks > 0 && B.dest != Q.me) {  // fragment bytes of this tile that left the GPU
          const int w = min(min(kTile, rw - tx * kTile), f.width - B.rx0 - tx * kTile);
          const int top = B.ry_first + ty * f.strip_stride;
          const int h = min(top + kTile, B.ry0 + B.rh) - max(top, B.ry0);
          if (w > 0 && h > 0) s_pushed += 20ull * static_cast<unsigned>(w * h);
        }
        s_taken[it & 1] = 0;  // free for iteration it + 2; nobody looks at it before the next barrier
      }
      fill_slot((it + 1) & 1);
    }
  }
  if (kCount) {
    for (int off = 16; off > 0; off >>= 1) nsamples += __shfl_down_sync(0xffffffffu, nsamples, off);
    if (lane == 0 && nsamples != 0)
      atomicAdd(Q.stats, static_cast<unsigned long long>(nsamples));
  }
  if (threadIdx.x == 0) {
    if (s_pushed != 0) atomicAdd(Q.stats + 1, s_pushed);
    if (s_stolen != 0) atomicAdd(Q.stats + 2, s_stolen);
#if AMRVR_MARCH_TIMELINE
    const unsigned long long cta_t1 = global_timer_ns();
    atomicAdd(Q.stats + 4, cta_t1 - s_t0);
    atomicMax(Q.stats + 5, ~s_t0);  // ~ of the earliest start
    atomicMax(Q.stats + 6, cta_t1);
    if (s_table_ns != 0) atomicAdd(Q.stats + 7, s_table_ns);
    if (blockIdx.x == 0) reinterpret_cast<unsigned*>(Q.stats)[7] = gridDim.x;  // beside the u32 tile counter
#endif
    if (Q.nranks > 0) {
      // every thread's fragment stores precede the barrier the loop was left through
      __threadfence_system();
      unsigned* done = Q.ctl[Q.me] + kCtlDoneMarch;
      if (atomicAdd(done, 1u) + 1u == gridDim.x) {
        st_volatile_u32(done, 0u);
        __threadfence_system();
        for (int r = 0; r < Q.nranks; ++r)
          st_volatile_u32(Q.ctl[r] + Q.signal_row * kMaxPeers + Q.me, Q.epoch);
      }
    }
  }
}

// ------------------------------------------------------------------------ K6

struct Px5 {
  float r, g, b, a, d;
};

// reference: Common/ImageRGBAFloatColorDepthSort.hpp:13-27
__device__ __forceinline__ Px5 blend_depth_swap(const Px5& top, const Px5& bottom) {
  const bool top_front = top.d <= bottom.d;
  const Px5& fr = top_front ? top : bottom;
  const Px5& bk = top_front ? bottom : top;
  const float t = 1.0f - fr.a;
  Px5 o;
  o.r = fr.r + bk.r * t;
  o.g = fr.g + bk.g * t;
  o.b = fr.b + bk.b * t;
  o.a = fr.a + bk.a * t;
  o.d = (bottom.d < top.d) ? bottom.d : top.d;
  return o;
}

// Wireframe overlay of one pixel (VolumeRenderer/VolumeRenderer.cpp:228-333, same
// arithmetic as overlay_kernel) followed by the byte rule of Common/Color.hpp:66-91.
// `candidates`: bit e set when edge e can reach this pixel (callers that know
// nothing pass all ones); edges are visited in index order either way.
__device__ __forceinline__ void overlay_and_quantise(int px, int py, float c0, float c1, float c2,
                                                     const OverlayEdges* edges, float pixel_radius,
                                                     unsigned char* out3, unsigned candidates = ~0u) {
  if (edges != nullptr) {
    if (edges->count < 32) candidates &= (1u << edges->count) - 1u;
    for (unsigned m = candidates; m != 0u; m &= m - 1u) {
      const int e = __ffs(m) - 1;
      if (px < edges->x0[e] || px > edges->x1[e] || py < edges->y0[e] || py > edges->y1[e]) continue;
      float coverage;
      if (edges->degenerate[e]) {
        coverage = 1.0f;
      } else {
        const float ax = edges->ax[e], ay = edges->ay[e], bx = edges->bx[e], by = edges->by[e];
        const float ex = bx - ax, ey = by - ay;
        const float len_sq = ex * ex + ey * ey;
        const float cx = static_cast<float>(px) + 0.5f, cy = static_cast<float>(py) + 0.5f;
        const float apx = cx - ax, apy = cy - ay;
        float t = (apx * ex + apy * ey) / len_sq;
        t = (t < 0.0f) ? 0.0f : ((1.0f < t) ? 1.0f : t);
        const float qx = ax + (bx - ax) * t, qy = ay + (by - ay) * t;
        const float ddx = cx - qx, ddy = cy - qy;
        const float dist = sqrtf(ddx * ddx + ddy * ddy);
        coverage = (pixel_radius + 0.5f - dist) * 0.6f;
        coverage = (coverage < 0.0f) ? 0.0f : ((1.0f < coverage) ? 1.0f : coverage);
        if (coverage <= 0.0f) continue;
      }
      const float a = coverage;
      const float s = 1.0f * a;
      c0 = s + c0 * (1.0f - a);
      c1 = s + c1 * (1.0f - a);
      c2 = s + c2 * (1.0f - a);
    }
  }
  const float c[3] = {c0, c1, c2};
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    const int v = static_cast<int>(c[k] * 256.f);
    out3[k] = static_cast<unsigned char>((v < 0) ? 0 : ((v > 255) ? 255 : v));
  }
}


constexpr int kBinThreads = 256;  // fold_bin_kernel
constexpr int kFoldThreads = 128; // fold_layers_kernel: 4 warps, each 32 pixels x 2 rows
constexpr int kFoldTileW = 32;
constexpr int kFoldBandRows = 8;  // one CTA tile = 32 x 8 pixels; one band = 8 frame rows
constexpr int kFoldDepth = 2;     // (register variant) list entries in flight per thread (x 2 pixels each)
constexpr int kFoldWarpList = 64; // per-warp list of matched layers (shared memory)
#ifndef AMRVR_FOLD_RING
#define AMRVR_FOLD_RING 4         // > 0: fragments travel through a per-thread ring of cp.async copies this deep
                                  // (C4 fold on one B200: 0.0925 ms without, 0.0905 / 0.0823 / 0.0864 ms at depth 6 / 4 / 3)
#endif
#ifndef AMRVR_FOLD_MIN_BLOCKS
#if AMRVR_FOLD_RING > 0
#define AMRVR_FOLD_MIN_BLOCKS 6   // 28 KB of shared memory per CTA, <= 85 registers
#else
#define AMRVR_FOLD_MIN_BLOCKS 8   // 64 registers: 32 warps per SM keep the HBM pipe full
#endif
#endif
constexpr int kFoldRing = AMRVR_FOLD_RING > 0 ? AMRVR_FOLD_RING : 1;

// cp.async: global -> shared without a register in between.  A thread only ever
// reads back what it copied itself, so cp.async.wait_group is all the
// synchronisation the ring needs.
__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_4(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem)), "l"(gmem) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int kPending>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(kPending) : "memory");
}

// A layer as one warp sees it: the address of the warp's first row inside the
// layer (never dereferenced when that row lies outside the slice), the column
// range, which of the warp's two rows the slice holds, the run id.  32 bytes.
struct FoldEntry {
  const float* row;  // row y0 of the layer: base + (y0 - row0) * w * 5
  int x0, w;
  int rows;          // bit 0: row y0 is in the slice, bit 1: row y0 + 1, bit 2: interleaved layout
  int run;
  int pad[2];
};

// The fully transparent pixel (DirectSendBase.cpp:427-432, 450-455).  It is a
// bit-exact identity of blend_depth_swap on either side for finite colours:
// x + 0 * (1 - a) == x and 0 + x * 1 == x, and min(d, inf) == d.  The fold
// therefore starts runs and the accumulator from it instead of carrying "have
// a value yet" flags -- every lane of a warp then runs the same control flow.
__device__ __forceinline__ Px5 transparent_px() { return Px5{0.f, 0.f, 0.f, 0.f, INFINITY}; }

// One fragment of a layer at column cx of the warp's row r (0 or 1).  Loads
// bypass L1 (ld.global.cg): a fragment is read once, and in a sharded frame it
// may have arrived over NVLink after this kernel started.
__device__ __forceinline__ Px5 fold_load(const FoldEntry& E, int r, int cx) {
  if (E.rows & 4) {  // the reference's own layout: 5 floats per pixel
    const float* q = E.row + (static_cast<long long>(r) * E.w + cx) * 5;
    return Px5{__ldcg(q), __ldcg(q + 1), __ldcg(q + 2), __ldcg(q + 3), __ldcg(q + 4)};
  }
  const float* rowp = E.row + static_cast<size_t>(r) * E.w * 5;
  const float4 c = __ldcg(reinterpret_cast<const float4*>(rowp) + cx);
  return Px5{c.x, c.y, c.z, c.w, __ldcg(rowp + 4 * E.w + cx)};
}

// First frame row of band `band` of the rows a rank folds.  Contiguous rows (strip_p == 1):
// first_row + 8 * band.  Striped frame (strip_p = P GPUs, first_row = 16 * rank): the rank's
// strips are 16 rows every 16 * P rows, two bands each.
__device__ __forceinline__ int fold_band_row(int first_row, int strip_p, int band) {
  return first_row + (band >> 1) * 2 * kFoldBandRows * strip_p + (band & 1) * kFoldBandRows;
}

// Pass 1: one CTA per band of 8 frame rows filters the globally ordered layer
// table down to the layers whose rows touch the band (ordered compaction, so
// the list keeps the blend order).  band_lists[band * nlayers + i], band_counts[band].
__global__ void __launch_bounds__(kBinThreads)
fold_bin_kernel(const FoldLayer* __restrict__ layers, int nlayers, int first_row, int strip_p,
                int* __restrict__ band_lists, int* __restrict__ band_counts,
                int2* __restrict__ band_xspan) {
  __shared__ int s_warp_count[kBinThreads / 32];
  __shared__ int s_xmin, s_xmax;
  if (threadIdx.x == 0) {
    s_xmin = INT_MAX;
    s_xmax = INT_MIN;
  }
  __syncthreads();
  int xmin = INT_MAX, xmax = INT_MIN;  // union of the x ranges of the band's layers
  const int band = blockIdx.x;
  const int y0 = fold_band_row(first_row, strip_p, band), y1 = y0 + kFoldBandRows;  // [y0, y1)
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int* list = band_lists + static_cast<size_t>(band) * nlayers;
  int written = 0;
  for (int chunk = 0; chunk < nlayers; chunk += kBinThreads) {
    const int l = chunk + threadIdx.x;
    bool keep = false;
    if (l < nlayers) {
      const int row0 = layers[l].row0, row1 = layers[l].row1;
      keep = row0 < y1 && row1 > y0;
      if (keep) {
        xmin = min(xmin, layers[l].x0);
        xmax = max(xmax, layers[l].x0 + layers[l].w);
      }
    }
    const unsigned ballot = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp_count[warp] = __popc(ballot);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kBinThreads / 32; ++w) {
      const int c = s_warp_count[w];
      if (w < warp) before += c;
      total += c;
    }
    if (keep) list[written + before + __popc(ballot & ((1u << lane) - 1u))] = l;
    written += total;
    __syncthreads();
  }
  for (int o = 16; o > 0; o >>= 1) {
    xmin = min(xmin, __shfl_xor_sync(0xffffffffu, xmin, o));
    xmax = max(xmax, __shfl_xor_sync(0xffffffffu, xmax, o));
  }
  if (lane == 0 && xmin < xmax) {
    atomicMin(&s_xmin, xmin);
    atomicMax(&s_xmax, xmax);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    band_counts[band] = written;
    band_xspan[band] = make_int2(s_xmin, s_xmax);  // empty band: (INT_MAX, INT_MIN)
  }
}

struct FoldArgs {
  const FoldLayer* layers;
  int nlayers;
  const int* band_lists;
  const int* band_counts;
  const int2* band_xspan;
  int first_row, bands, width, height;
  int strip_p, strip_g;  // striped frame: GPUs and this rank (layers then hold the rank's strips, 16 rows each); else 1, 0
  int pixel_begin, pixel_end;
  float* piece5;
  const OverlayEdges* edges;
  unsigned char* rgb8;
  FoldSync sync;  // nranks == 0: no hand-shake
};

// Pass 2.  Persistent CTAs walk the 32 x 8 pixel tiles of this rank's piece
// (pixels outside [pixel_begin, pixel_end) are masked: a piece may start and end
// inside a row).  A warp owns 32 columns x 2 rows of the tile and works alone --
// no CTA barrier anywhere: it scans its band's ordered layer list 32 entries at
// a time, keeps the layers whose rectangle reaches its 64 pixels (one ballot
// per chunk, ordered append to a warp-private list in shared memory) and folds
// them, kFoldDepth layers x 2 pixels of loads in flight per thread: consecutive
// layers of one run are blended first, then the run result is blended into the
// accumulator -- DirectSendBase::composeLayered, per pixel
// (DirectSendBase.cpp:400-446).  In a sharded frame all fragments of the piece
// already lie in this GPU's pool (the marches pushed them), so the fold reads
// local HBM only; it first waits for every rank's "march finished" flag.
// kBytes: the epilogue applies the overlay and the byte rule and writes RGB8
// rows (top row = y_max, SavePPM order) -- into the root's frame, across NVLink
// when the root is another GPU -- instead of 20 B per pixel.
template <bool kBytes>
__global__ void __launch_bounds__(kFoldThreads, AMRVR_FOLD_MIN_BLOCKS)
fold_layers_kernel(const __grid_constant__ FoldArgs A) {
  __shared__ __align__(16) FoldEntry s_list[kFoldThreads / 32][kFoldWarpList];
#if AMRVR_FOLD_RING > 0
  // ring slot of a thread: RGBA of its two pixels, then their depths (40 bytes)
  __shared__ __align__(16) float4 s_ring_rgba[kFoldRing][2][kFoldThreads];
  __shared__ float s_ring_depth[kFoldRing][2][kFoldThreads];
#endif
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (A.sync.nranks > 0) {
    // every march that may have written into this pool has finished
    if (warp == 0) {
      unsigned* mine = static_cast<unsigned*>(A.sync.ctl.p[A.sync.me]);
      if (lane < A.sync.nranks) {
        const unsigned* flag = mine + A.sync.wait_row * kMaxPeers + lane;
        const unsigned long long t0 = global_timer_ns();
        while (static_cast<int>(ld_volatile_u32(flag) - A.sync.epoch) < 0) {
          __nanosleep(32);
          if (global_timer_ns() - t0 > A.sync.timeout_ns) {
            st_volatile_u32(mine + kCtlError, 1u);
            break;
          }
        }
      }
      __threadfence_system();
      // nobody claims tiles any more: the march queue counter is free for the next frame
      if (blockIdx.x == 0 && lane == 0) st_volatile_u32(mine + kCtlTileCounter, 0u);
    }
    __syncthreads();
  }

  FoldEntry* list = s_list[warp];
  const int tiles_x = (A.width + kFoldTileW - 1) / kFoldTileW;
  const int ntiles = tiles_x * A.bands;
  for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const int band = tile / tiles_x;
    const int tile_x0 = (tile - band * tiles_x) * kFoldTileW;
    const int y0 = fold_band_row(A.first_row, A.strip_p, band) + 2 * warp, y1 = y0 + 1;
    const int px = tile_x0 + lane;
    const long long p0 = static_cast<long long>(y0) * A.width + px;
    const long long p1 = p0 + A.width;
    const bool live0 = px < A.width && p0 >= A.pixel_begin && p0 < A.pixel_end;
    const bool live1 = px < A.width && p1 >= A.pixel_begin && p1 < A.pixel_end;
    const int x_last = min(tile_x0 + kFoldTileW, A.width) - 1;

    // run being pre-blended and accumulator of the two pixels; the run id is the same
    // in every lane (all lanes walk the same list), so run changes never diverge
    Px5 run0 = transparent_px(), run1 = run0, acc0 = run0, acc1 = run0;
    int cur_run = -1;
    const int* band_list = A.band_lists + static_cast<size_t>(band) * A.nlayers;
    // no layer of this band reaches the tile's columns: nothing to scan
    const int2 xspan = __ldg(&A.band_xspan[band]);
    const int count = (xspan.x <= x_last && xspan.y > tile_x0) ? __ldg(&A.band_counts[band]) : 0;

    int nlist = 0;
    for (int chunk = 0; chunk < count; chunk += 32) {
      const int i = chunk + lane;
      bool keep = false;
      FoldLayer L;
      if (i < count) {
        L = A.layers[__ldg(&band_list[i])];
        keep = L.row0 <= y1 && L.row1 > y0 && L.x0 <= x_last && L.x0 + L.w > tile_x0;
      }
      const unsigned ballot = __ballot_sync(0xffffffffu, keep);
      if (keep) {
        FoldEntry E;
        long long layer_row = static_cast<long long>(y0) - L.row0;
        // striped: the layer stores this rank's strips one after the other, 16 rows each
        if (A.strip_p > 1) layer_row = stripe_layer_row(A.strip_p, A.strip_g, L.row0, y0);
        E.row = L.base + layer_row * L.w * 5;
        E.x0 = L.x0;
        E.w = L.w;
        E.rows = ((y0 >= L.row0 && y0 < L.row1) ? 1 : 0) | ((y1 >= L.row0 && y1 < L.row1) ? 2 : 0) |
                 (L.interleaved ? 4 : 0);
        E.run = L.run;
        E.pad[0] = E.pad[1] = 0;
        list[nlist + __popc(ballot & ((1u << lane) - 1u))] = E;
      }
      nlist += __popc(ballot);
      __syncwarp();
      if (nlist <= kFoldWarpList - 32 && chunk + 32 < count) continue;  // room for another chunk
#if AMRVR_FOLD_RING > 0
      // fold the list through the ring: entry e + kFoldRing is requested when entry e has
      // been blended, so kFoldRing layers x 2 pixels x 20 B per thread are in flight and no
      // register waits for them; one (possibly empty) group per entry keeps the count fixed
      auto covers = [&](const FoldEntry& E, bool& in0, bool& in1, int& cx) {
        cx = px - E.x0;
        const bool inx = cx >= 0 && cx < E.w;
        in0 = live0 && inx && (E.rows & 1);
        in1 = live1 && inx && (E.rows & 2);
      };
      auto request = [&](int e) {
        if (e < nlist) {
          const FoldEntry E = list[e];  // same address in every lane: broadcast
          bool in0, in1;
          int cx;
          covers(E, in0, in1, cx);
          const int slot = e % kFoldRing;
#pragma unroll
          for (int r = 0; r < 2; ++r) {
            if (!(r == 0 ? in0 : in1)) continue;
            float4* dst = &s_ring_rgba[slot][r][threadIdx.x];
            float* dst_d = &s_ring_depth[slot][r][threadIdx.x];
            if (E.rows & 4) {  // the reference's own layout: 5 floats per pixel, 4-byte aligned only
              const float* q = E.row + (static_cast<long long>(r) * E.w + cx) * 5;
#pragma unroll
              for (int k = 0; k < 4; ++k) cp_async_4(reinterpret_cast<float*>(dst) + k, q + k);
              cp_async_4(dst_d, q + 4);
            } else {
              const float* rowp = E.row + static_cast<size_t>(r) * E.w * 5;
              cp_async_16(dst, reinterpret_cast<const float4*>(rowp) + cx);
              cp_async_4(dst_d, rowp + 4 * E.w + cx);
            }
          }
        }
        cp_async_commit();
      };
#pragma unroll
      for (int e = 0; e < kFoldRing; ++e) request(e);
      for (int e = 0; e < nlist; ++e) {
        cp_async_wait<kFoldRing - 1>();
        const FoldEntry E = list[e];
        bool in0, in1;
        int cx;
        covers(E, in0, in1, cx);
        const int slot = e % kFoldRing;
        Px5 f0 = transparent_px(), f1 = f0;
        if (in0) {
          const float4 c = s_ring_rgba[slot][0][threadIdx.x];
          f0 = Px5{c.x, c.y, c.z, c.w, s_ring_depth[slot][0][threadIdx.x]};
        }
        if (in1) {
          const float4 c = s_ring_rgba[slot][1][threadIdx.x];
          f1 = Px5{c.x, c.y, c.z, c.w, s_ring_depth[slot][1][threadIdx.x]};
        }
        if (E.run != cur_run) {  // uniform: the run closes (DirectSendBase.cpp:441-445)
          acc0 = blend_depth_swap(acc0, run0);
          acc1 = blend_depth_swap(acc1, run1);
          run0 = run1 = transparent_px();
          cur_run = E.run;
        }
        run0 = blend_depth_swap(run0, f0);  // :413-426
        run1 = blend_depth_swap(run1, f1);
        request(e + kFoldRing);  // after the slot's values were used
      }
      cp_async_wait<0>();
#else
      // fold the list: all loads of a batch are issued before its first blend
      for (int e0 = 0; e0 < nlist; e0 += kFoldDepth) {
        Px5 f0[kFoldDepth], f1[kFoldDepth];
        int runid[kFoldDepth];
#pragma unroll
        for (int k = 0; k < kFoldDepth; ++k) {
          f0[k] = transparent_px();
          f1[k] = transparent_px();
          runid[k] = INT_MIN;  // past the end of the list: blends a transparent pixel, closes no run
          if (e0 + k < nlist) {
            const FoldEntry E = list[e0 + k];  // same address in every lane: broadcast
            runid[k] = E.run;
            const int cx = px - E.x0;
            const bool inx = cx >= 0 && cx < E.w;
            if (live0 && inx && (E.rows & 1)) f0[k] = fold_load(E, 0, cx);
            if (live1 && inx && (E.rows & 2)) f1[k] = fold_load(E, 1, cx);
          }
        }
#pragma unroll
        for (int k = 0; k < kFoldDepth; ++k) {
          if (runid[k] != cur_run && runid[k] != INT_MIN) {  // uniform: the run closes (DirectSendBase.cpp:441-445)
            acc0 = blend_depth_swap(acc0, run0);
            acc1 = blend_depth_swap(acc1, run1);
            run0 = run1 = transparent_px();
            cur_run = runid[k];
          }
          run0 = blend_depth_swap(run0, f0[k]);  // :413-426
          run1 = blend_depth_swap(run1, f1[k]);
        }
      }
#endif
      nlist = 0;
      __syncwarp();  // the list is rewritten by the next chunk
    }
    const Px5 a0 = blend_depth_swap(acc0, run0), a1 = blend_depth_swap(acc1, run1);

    // the warp's list doubles as its output staging (2 rows x 32 pixels x 20 B = 1280 B)
    if (!kBytes || A.piece5 != nullptr) {
      float* s_out = reinterpret_cast<float*>(list);  // 160 floats per row
      const Px5 a[2] = {a0, a1};
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float* o = s_out + 160 * r + lane * 5;
        o[0] = a[r].r;
        o[1] = a[r].g;
        o[2] = a[r].b;
        o[3] = a[r].a;
        o[4] = a[r].d;
      }
      __syncwarp();
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        // up to 32 pixels x 5 floats, contiguous in the piece
        const long long row_first = static_cast<long long>(y0 + r) * A.width + tile_x0;
        const long long seg_lo = max(static_cast<long long>(A.pixel_begin) - row_first, 0LL);
        const long long seg_hi = min(min(static_cast<long long>(A.pixel_end) - row_first,
                                         static_cast<long long>(A.width - tile_x0)),
                                     static_cast<long long>(kFoldTileW));
        const float* src = s_out + 160 * r;
        float* dst = A.piece5 + (row_first - A.pixel_begin) * 5;
        if (A.strip_p > 1) {  // the piece holds this rank's strips one after the other
          const long long local_row = static_cast<long long>(band >> 1) * 16 + (band & 1) * kFoldBandRows + 2 * warp + r;
          dst = A.piece5 + (local_row * A.width + tile_x0) * 5;
        }
#pragma unroll
        for (int j = 0; j < 5; ++j) {
          const int fl = lane + 32 * j;  // float within the row segment
          const int pix = fl / 5;        // pixel within the segment
          if (pix >= seg_lo && pix < seg_hi) dst[fl] = src[fl];
        }
      }
      __syncwarp();
    }
    if (kBytes) {
      unsigned char* s_bytes = reinterpret_cast<unsigned char*>(list);  // 96 B per row
      // which of the (at most 12) wireframe edges can reach these 32 x 2 pixels at all:
      // most see none, and the per-pixel loop then has nothing to visit
      unsigned edge_mask = 0u;
      if (A.edges != nullptr) {
        const int ecount = __ldg(&A.edges->count);
        const bool reach = lane < ecount && lane < 12 &&
                           !(x_last < __ldg(&A.edges->x0[lane]) || tile_x0 > __ldg(&A.edges->x1[lane]) ||
                             y1 < __ldg(&A.edges->y0[lane]) || y0 > __ldg(&A.edges->y1[lane]));
        edge_mask = __ballot_sync(0xffffffffu, reach);
      }
      overlay_and_quantise(px, y0, a0.r, a0.g, a0.b, A.edges, 0.5f, s_bytes + lane * 3, edge_mask);
      overlay_and_quantise(px, y1, a1.r, a1.g, a1.b, A.edges, 0.5f, s_bytes + 96 + lane * 3, edge_mask);
      __syncwarp();
      const int npx = min(A.width - tile_x0, kFoldTileW);  // pixels of this row segment
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const int y = y0 + r;
        if (y >= A.height) continue;
        const long long first = static_cast<long long>(y) * A.width + tile_x0;  // pixel index of lane 0
        unsigned char* row = A.rgb8 + 3 * (static_cast<size_t>(A.height - 1 - y) * A.width + tile_x0);
        const bool whole = first >= A.pixel_begin && first + npx <= A.pixel_end;
        if (whole && (A.width & 3) == 0 && (npx & 3) == 0) {
          // tile_x0 is a multiple of 32 and the width of 4: whole aligned words
          if (lane < (npx * 3) / 4)
            reinterpret_cast<unsigned*>(row)[lane] = reinterpret_cast<const unsigned*>(s_bytes + 96 * r)[lane];
        } else if (r == 0 ? live0 : live1) {
          for (int k = 0; k < 3; ++k) row[lane * 3 + k] = s_bytes[96 * r + lane * 3 + k];
        }
      }
      __syncwarp();
    }
  }

  if (A.sync.nranks > 0) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence_system();  // this CTA's bytes / pixels are out before the count
      unsigned* mine = static_cast<unsigned*>(A.sync.ctl.p[A.sync.me]);
      unsigned* done = mine + kCtlDoneFold;
      if (atomicAdd(done, 1u) + 1u == gridDim.x) {
        st_volatile_u32(done, 0u);
        __threadfence_system();
        for (int r = 0; r < A.sync.nranks; ++r)
          st_volatile_u32(static_cast<unsigned*>(A.sync.ctl.p[r]) + A.sync.signal_row * kMaxPeers + A.sync.me,
                          A.sync.epoch);
      }
    }
  }
}

// ------------------------------------------------------------------------ C6

__global__ void __launch_bounds__(256)
pack_slices_kernel(const CopyJob* __restrict__ jobs, const float* __restrict__ src,
                   float* __restrict__ dst) {
  const CopyJob job = jobs[blockIdx.x];
  const float4* s = reinterpret_cast<const float4*>(src + job.src_off);
  float4* d = reinterpret_cast<float4*>(dst + job.dst_off);
  const unsigned n4 = job.count >> 2;
  for (unsigned i = threadIdx.x; i < n4; i += blockDim.x) d[i] = s[i];
}

// Fused pack + push: the same row slices, stored straight into the receive pool of the rank whose piece
// needs them (CopyJob::pad = destination rank, dst_off = offset inside THAT rank's pool) over NVLink.
__global__ void __launch_bounds__(256)
push_slices_kernel(const CopyJob* __restrict__ jobs, const float* __restrict__ src,
                   const __grid_constant__ PeerPtrs recv) {
  const CopyJob job = jobs[blockIdx.x];
  const float4* s = reinterpret_cast<const float4*>(src + job.src_off);
  float4* d = reinterpret_cast<float4*>(static_cast<float*>(recv.p[job.pad]) + job.dst_off);
  const unsigned n4 = job.count >> 2;
  unsigned i = threadIdx.x;
  for (; i + 3 * 256 < n4; i += 4 * 256) {  // four 16 B loads in flight before the first remote store
    const float4 a = __ldcs(s + i), b = __ldcs(s + i + 256), c = __ldcs(s + i + 512), e = __ldcs(s + i + 768);
    d[i] = a;
    d[i + 256] = b;
    d[i + 512] = c;
    d[i + 768] = e;
  }
  for (; i < n4; i += 256) d[i] = __ldcs(s + i);
}

// ------------------------------------------------------------------ K8, post

// reference: VolumeRenderer/VolumeRenderer.cpp:479-528 (sum order dy-major, dx-minor)
__global__ void __launch_bounds__(256)
downsample_kernel(const float* __restrict__ src5, int width, int height, int block,
                  float* __restrict__ dst5) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= width * height) return;
  const int x = p % width, y = p / width;
  const int src_w = width * block;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  for (int dy = 0; dy < block; ++dy)
    for (int dx = 0; dx < block; ++dx) {
      const float* s = src5 + 5 * (static_cast<size_t>(y * block + dy) * src_w + (x * block + dx));
      s0 += s[0];
      s1 += s[1];
      s2 += s[2];
      s3 += s[3];
    }
  const float inv = 1.0f / static_cast<float>(block * block);
  float* d = dst5 + 5 * static_cast<size_t>(p);
  d[0] = s0 * inv;
  d[1] = s1 * inv;
  d[2] = s2 * inv;
  d[3] = s3 * inv;
  d[4] = INFINITY;
}

// reference: VolumeRenderer/VolumeRenderer.cpp:228-333.  One thread per pixel walks
// the 12 edges in order, so overlapping edges blend in the reference's order.
__global__ void __launch_bounds__(256)
overlay_kernel(const __grid_constant__ OverlayEdges edges, float pixel_radius, int width, int height,
               float* __restrict__ image5) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= width * height) return;
  const int px = p % width, py = p / width;
  float* b = image5 + 5 * static_cast<size_t>(p);
  float c0 = b[0], c1 = b[1], c2 = b[2], c3 = b[3], c4 = b[4];
  bool touched = false;
  for (int e = 0; e < edges.count; ++e) {
    if (px < edges.x0[e] || px > edges.x1[e] || py < edges.y0[e] || py > edges.y1[e]) continue;
    float coverage;
    if (edges.degenerate[e]) {
      coverage = 1.0f;
    } else {
      const float ax = edges.ax[e], ay = edges.ay[e], bx = edges.bx[e], by = edges.by[e];
      const float ex = bx - ax, ey = by - ay;
      const float len_sq = ex * ex + ey * ey;
      const float cx = static_cast<float>(px) + 0.5f, cy = static_cast<float>(py) + 0.5f;
      const float apx = cx - ax, apy = cy - ay;
      float t = (apx * ex + apy * ey) / len_sq;
      t = (t < 0.0f) ? 0.0f : ((1.0f < t) ? 1.0f : t);
      const float qx = ax + (bx - ax) * t, qy = ay + (by - ay) * t;
      const float ddx = cx - qx, ddy = cy - qy;
      const float dist = sqrtf(ddx * ddx + ddy * ddy);
      coverage = (pixel_radius + 0.5f - dist) * 0.6f;
      coverage = (coverage < 0.0f) ? 0.0f : ((1.0f < coverage) ? 1.0f : coverage);
      if (coverage <= 0.0f) continue;
    }
    const float a = coverage;  // already in [0,1]
    const float s = 1.0f * a;  // white line
    c0 = s + c0 * (1.0f - a);
    c1 = s + c1 * (1.0f - a);
    c2 = s + c2 * (1.0f - a);
    c3 = a + c3 * (1.0f - a);
    c4 = -FLT_MAX;
    touched = true;
  }
  if (touched) {
    b[0] = c0;
    b[1] = c1;
    b[2] = c2;
    b[3] = c3;
    b[4] = c4;
  }
}

// reference: Common/Color.hpp:66-91 byte rule, Common/SavePPM.cpp:25-32 row order
__global__ void __launch_bounds__(256)
quantize_rgb8_kernel(const float* __restrict__ image5, int width, int height,
                     unsigned char* __restrict__ rgb8) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= width * height) return;
  const int x = p % width, out_row = p / width;
  const float* s = image5 + 5 * (static_cast<size_t>(height - 1 - out_row) * width + x);
  unsigned char* d = rgb8 + 3 * static_cast<size_t>(p);
#pragma unroll
  for (int c = 0; c < 3; ++c) {
    const int v = static_cast<int>(s[c] * 256.f);
    d[c] = static_cast<unsigned char>((v < 0) ? 0 : ((v > 255) ? 255 : v));
  }
}

// ------------------------------------------------------- peer-memory exchange



__global__ void __launch_bounds__(32)
peer_barrier_kernel(const __grid_constant__ PeerPtrs flags, int me, int nranks, int which,
                    unsigned epoch, unsigned long long timeout_ns, unsigned* error_flag) {
  const int q = threadIdx.x;
  if (q >= nranks) return;
  // everything this GPU wrote before (layers, pieces, bytes) is visible system-wide first
  __threadfence_system();
  volatile unsigned* theirs = static_cast<unsigned*>(flags.p[q]) + which * kMaxPeers + me;
  *theirs = epoch;
  volatile unsigned* mine = static_cast<unsigned*>(flags.p[me]) + which * kMaxPeers + q;
  const unsigned long long t0 = global_timer_ns();
  while (static_cast<int>(*mine - epoch) < 0) {
    __nanosleep(64);
    if (global_timer_ns() - t0 > timeout_ns) {
      *error_flag = 1u;
      break;
    }
  }
  __threadfence_system();
}

// Wait half of the barrier: the signals were raised by the last CTA of a fold (or
// march) on every rank.
__global__ void __launch_bounds__(32)
peer_wait_kernel(const __grid_constant__ PeerPtrs ctl, int me, int nranks, int row, unsigned epoch,
                 unsigned long long timeout_ns) {
  const int q = threadIdx.x;
  unsigned* mine = static_cast<unsigned*>(ctl.p[me]);
  if (q < nranks) {
    const unsigned* flag = mine + row * kMaxPeers + q;
    const unsigned long long t0 = global_timer_ns();
    while (static_cast<int>(ld_volatile_u32(flag) - epoch) < 0) {
      __nanosleep(32);
      if (global_timer_ns() - t0 > timeout_ns) {
        st_volatile_u32(mine + kCtlError, 1u);
        break;
      }
    }
  }
  __threadfence_system();
}

// NVLink probe: a plain grid-stride copy, 16 B per thread per step, four loads in
// flight; src or dst may be another GPU's memory.
__global__ void __launch_bounds__(256)
peer_copy_kernel(const uint4* __restrict__ src, uint4* __restrict__ dst, size_t n16) {
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  size_t i = blockIdx.x * static_cast<size_t>(blockDim.x) + threadIdx.x;
  for (; i + 3 * stride < n16; i += 4 * stride) {
    const uint4 a = src[i], b = src[i + stride], c = src[i + 2 * stride], d = src[i + 3 * stride];
    dst[i] = a;
    dst[i + stride] = b;
    dst[i + 2 * stride] = c;
    dst[i + 3 * stride] = d;
  }
  for (; i < n16; i += stride) dst[i] = src[i];
}

// reference: VolumeRenderer/VolumeRenderer.cpp:228-333 (overlay, same arithmetic as
// overlay_kernel) then Common/Color.hpp:66-91 + SavePPM row order, per pixel of a
// piece; the three bytes go to the root's frame.
__global__ void __launch_bounds__(256)
post_piece_kernel(const float* __restrict__ piece5, int pixel_begin, int pixel_end, int width,
                  int height, const OverlayEdges* __restrict__ edges_ptr, float pixel_radius,
                  unsigned char* __restrict__ rgb8_root) {
  __shared__ OverlayEdges s_edges;
  __shared__ __align__(16) unsigned char s_bytes[256 * 3];
  const bool overlay = edges_ptr != nullptr;
  if (overlay) {
    const int words = sizeof(OverlayEdges) / 4;
    for (int i = threadIdx.x; i < words; i += blockDim.x)
      reinterpret_cast<int*>(&s_edges)[i] = reinterpret_cast<const int*>(edges_ptr)[i];
    __syncthreads();
  }
  // the CTA's 256 pixels start at a multiple of 4, so that (width % 4 == 0) every
  // group of 4 pixels = 12 bytes = 3 aligned words lies in one row of the frame
  const int p0 = (pixel_begin & ~3) + blockIdx.x * 256;
  const int p = p0 + threadIdx.x;
  const bool live = p >= pixel_begin && p < pixel_end;
  if (live) {
    const float* b = piece5 + 5 * static_cast<size_t>(p - pixel_begin);
    const int px = p % width, py = p / width;
    overlay_and_quantise(px, py, b[0], b[1], b[2], overlay ? &s_edges : nullptr, pixel_radius,
                         s_bytes + threadIdx.x * 3);
  }
  __syncthreads();
  if ((width & 3) == 0) {
    // 192 words; word w belongs to the 4-pixel group w / 3 (one row, bottom-up row order)
    if (threadIdx.x < 192) {
      const int w = threadIdx.x, g = w / 3;
      const int pg = p0 + 4 * g;
      if (pg < pixel_end && pg + 3 >= pixel_begin) {
        const int px = pg % width, py = pg / width;
        unsigned char* row = rgb8_root + 3 * (static_cast<size_t>(height - 1 - py) * width + px);
        if (pg >= pixel_begin && pg + 3 < pixel_end) {
          reinterpret_cast<unsigned*>(row)[w - 3 * g] = reinterpret_cast<const unsigned*>(s_bytes)[w];
        } else {  // a piece boundary cuts the group: bytes of the live pixels only
          for (int k = 0; k < 4; ++k) {
            const int byte = 4 * (w - 3 * g) + k, q = pg + byte / 3;
            if (q >= pixel_begin && q < pixel_end) row[byte] = s_bytes[12 * g + byte];
          }
        }
      }
    }
  } else if (live) {
    const int px = p % width, py = p / width;
    unsigned char* d = rgb8_root + 3 * (static_cast<size_t>(height - 1 - py) * width + px);
    for (int k = 0; k < 3; ++k) d[k] = s_bytes[threadIdx.x * 3 + k];
  }
}

// ------------------------------------------------------------------------ K7

__device__ __forceinline__ unsigned ordered_bits(float v) {
  if (v == 0.0f) v = 0.0f;  // -0 and +0 compare equal in the reference comparator
  const unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int kSortThreads = 256;

// Single-CTA least-significant-digit radix sort (4-bit digits) of the
// (minDepth, maxDepth, owner) keys; the stable passes keep equal keys in index
// order, which is the last tie-break of Common/VisibilityOrdering.cpp:238-257.
// Non-finite depth intervals were already collapsed to (+inf,+inf) and so sort
// last.  priority[i] = final position of box i.
__global__ void __launch_bounds__(kSortThreads)
order_rank_kernel(const OrderBox* __restrict__ boxes, int n, uint4* __restrict__ buf_a,
                  uint4* __restrict__ buf_b, int owner_bits, int* __restrict__ priority) {
  __shared__ unsigned s_hist[16 * kSortThreads];
  __shared__ unsigned s_warp[kSortThreads / 32];
  const int t = threadIdx.x;
  const int per = (n + kSortThreads - 1) / kSortThreads;
  const int begin = min(t * per, n), end = min(begin + per, n);

  for (int i = t; i < n; i += kSortThreads) {
    const OrderBox b = boxes[i];
    buf_a[i] = make_uint4(static_cast<unsigned>(b.owner), ordered_bits(b.max_depth),
                          ordered_bits(b.min_depth), static_cast<unsigned>(i));
  }
  __syncthreads();

  uint4* in = buf_a;
  uint4* out = buf_b;
  const int owner_passes = (owner_bits + 3) / 4;
  for (int pass = 0; pass < owner_passes + 16; ++pass) {
    int word, shift;  // x = owner (least significant), y = max depth, z = min depth
    if (pass < owner_passes) {
      word = 0;
      shift = 4 * pass;
    } else {
      const int q = pass - owner_passes;
      word = 1 + (q >> 3);
      shift = 4 * (q & 7);
    }
    unsigned local[16];
#pragma unroll
    for (int d = 0; d < 16; ++d) local[d] = 0;
    for (int i = begin; i < end; ++i) {
      const uint4 k = in[i];
      const unsigned w = (word == 0) ? k.x : ((word == 1) ? k.y : k.z);
      const unsigned digit = (w >> shift) & 15u;
#pragma unroll
      for (int d = 0; d < 16; ++d) local[d] += (digit == static_cast<unsigned>(d));
    }
#pragma unroll
    for (int d = 0; d < 16; ++d) s_hist[d * kSortThreads + t] = local[d];
    __syncthreads();
    // exclusive scan of the 16*256 counters in (digit, thread) order
    unsigned sum = 0;
    unsigned vals[16];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      vals[e] = s_hist[t * 16 + e];
      sum += vals[e];
    }
    unsigned incl = sum;
    const int lane = t & 31, warp = t >> 5;
    for (int off = 1; off < 32; off <<= 1) {
      const unsigned o = __shfl_up_sync(0xffffffffu, incl, off);
      if (lane >= off) incl += o;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    unsigned base = incl - sum;
    for (int w = 0; w < warp; ++w) base += s_warp[w];
#pragma unroll
    for (int e = 0; e < 16; ++e) {
      s_hist[t * 16 + e] = base;
      base += vals[e];
    }
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 16; ++d) local[d] = s_hist[d * kSortThreads + t];
    for (int i = begin; i < end; ++i) {
      const uint4 k = in[i];
      const unsigned w = (word == 0) ? k.x : ((word == 1) ? k.y : k.z);
      const unsigned digit = (w >> shift) & 15u;
      unsigned pos = 0;
#pragma unroll
      for (int d = 0; d < 16; ++d)
        if (digit == static_cast<unsigned>(d)) pos = local[d]++;
      out[pos] = k;
    }
    __syncthreads();
    uint4* tmp = in;
    in = out;
    out = tmp;
  }
  for (int i = t; i < n; i += kSortThreads) priority[in[i].w] = i;
}

// The same total order by counting: priority[i] = number of boxes whose key
// (minDepth, maxDepth, owner, index) is smaller than box i's.  The keys are
// distinct (the index is part of them), so the counts are a permutation -- the
// one a stable sort by (minDepth, maxDepth, owner) produces.  CTA (bx, by)
// compares 256 boxes against 256 others held in shared memory; n = 4096 is
// 256 CTAs of 64 K key comparisons each, no passes and no scans, against the
// 17 dependent passes of the single-CTA radix sort.
__global__ void __launch_bounds__(256)
order_count_kernel(const OrderBox* __restrict__ boxes, int n, int* __restrict__ priority) {
  __shared__ uint4 s_key[256];
  const int j0 = blockIdx.y * 256;
  {
    const int j = j0 + threadIdx.x;
    uint4 k = make_uint4(0xffffffffu, 0xffffffffu, 0xffffffffu, 0xffffffffu);
    if (j < n) {
      const OrderBox b = boxes[j];
      k = make_uint4(ordered_bits(b.min_depth), ordered_bits(b.max_depth),
                     static_cast<unsigned>(b.owner), static_cast<unsigned>(j));
    }
    s_key[threadIdx.x] = k;
  }
  __syncthreads();
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= n) return;
  const OrderBox me = boxes[i];
  const unsigned a0 = ordered_bits(me.min_depth), a1 = ordered_bits(me.max_depth);
  const unsigned a2 = static_cast<unsigned>(me.owner), a3 = static_cast<unsigned>(i);
  const int m = min(256, n - j0);
  int before = 0;
  for (int t = 0; t < m; ++t) {
    const uint4 k = s_key[t];
    const bool less = (k.x != a0) ? (k.x < a0)
                                  : ((k.y != a1) ? (k.y < a1) : ((k.z != a2) ? (k.z < a2) : (k.w < a3)));
    before += less ? 1 : 0;
  }
  if (before != 0) atomicAdd(&priority[i], before);
}

__device__ __forceinline__ float fmax_std(float a, float b) { return (a < b) ? b : a; }
__device__ __forceinline__ float fmin_std(float a, float b) { return (b < a) ? b : a; }

__device__ __forceinline__ bool faces_touch(float a, float b) {
  const float scale = fmax_std(fmax_std(1.0f, fabsf(a)), fabsf(b));
  return fabsf(a - b) <= 1e-5f * scale;
}
__device__ __forceinline__ bool spans_overlap_dev(float a0, float a1, float b0, float b1) {
  const float lo = fmax_std(a0, b0), hi = fmin_std(a1, b1);
  float scale = 1.0f;
  scale = fmax_std(scale, fabsf(a0));
  scale = fmax_std(scale, fabsf(a1));
  scale = fmax_std(scale, fabsf(b0));
  scale = fmax_std(scale, fabsf(b1));
  scale = fmax_std(scale, fabsf(lo));
  scale = fmax_std(scale, fabsf(hi));
  return (hi - lo) > 1e-5f * scale;
}

// All pairs i<j, three axes: face-adjacent boxes whose cross-sections overlap
// get a front->back edge chosen by the sign of the view direction on that axis
// (Common/VisibilityOrdering.cpp:261-317).  Edges are appended unordered; the
// host sorts them back into the reference's insertion order.
__global__ void __launch_bounds__(256)
order_pairs_kernel(const OrderBox* __restrict__ boxes, int n, float dir0, float dir1, float dir2,
                   OrderEdge* __restrict__ edges, int capacity, int* __restrict__ count) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y;
  if (j >= n || j <= i) return;
  const OrderBox a = boxes[i];
  const OrderBox b = boxes[j];
  const float dir[3] = {dir0, dir1, dir2};
#pragma unroll
  for (int axis = 0; axis < 3; ++axis) {
    const int u = (axis + 1) % 3, v = (axis + 2) % 3;
    if (!spans_overlap_dev(a.lo[u], a.hi[u], b.lo[u], b.hi[u])) continue;
    if (!spans_overlap_dev(a.lo[v], a.hi[v], b.lo[v], b.hi[v])) continue;
    const float along = dir[axis];
    int forward = -1;
    if (faces_touch(a.hi[axis], b.lo[axis])) {
      if (along > 1e-6f) forward = 0;
      else if (along < -1e-6f) forward = 1;
    } else if (faces_touch(b.hi[axis], a.lo[axis])) {
      if (along > 1e-6f) forward = 1;
      else if (along < -1e-6f) forward = 0;
    }
    if (forward >= 0) {
      const int slot = atomicAdd(count, 1);
      if (slot < capacity) edges[slot] = OrderEdge{i, j, axis, forward};
    }
  }
}

inline cudaStream_t as_stream(void* s) { return static_cast<cudaStream_t>(s); }

ValuePath make_value_path(const amrvr_scalar_transform& t, float range_min, float range_max) {
  ValuePath p;
  p.log_scale = t.log_scale_input;
  p.normalize = t.normalize_to_unit_range;
  p.positive_floor = t.positive_floor;
  p.norm_min = t.normalization_min;
  p.inv_norm_span = t.inverse_normalization_span;
  p.range_min = range_min;
  p.inv_range = (range_max != range_min) ? 1.0f / (range_max - range_min) : 1.0f;  // :719-722
  const float cs = (range_max < 0.0f) ? 0.0f : ((1.0f < range_max) ? 1.0f : range_max);
  p.clip_start = cs;
  p.apply_clip = 1.0f > cs + 1e-5f;
  return p;
}

}  // namespace

// ------------------------------------------------------------------ launchers

int launch_brick_stats(const double* data, int64_t jstride, int64_t kstride, int nx, int ny,
                       int nz, double* partials, int blocks, void* stream, int64_t* launches) {
  brick_stats_kernel<<<blocks, 256, 0, as_stream(stream)>>>(data, jstride, kstride, nx, ny, nz,
                                                            partials);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_brick_histogram(const double* data, int64_t jstride, int64_t kstride, int nx, int ny,
                           int nz, const amrvr_scalar_transform& t, float range_min,
                           float range_max, int bin_count, unsigned long long* counts_dev,
                           void* stream, int64_t* launches) {
  const long long ncells = static_cast<long long>(nx) * ny * nz;
  if (ncells <= 0) return 0;
  ValuePath p{};
  p.log_scale = t.log_scale_input;
  p.normalize = t.normalize_to_unit_range;
  p.positive_floor = t.positive_floor;
  p.norm_min = t.normalization_min;
  p.inv_norm_span = t.inverse_normalization_span;
  const float width = range_max - range_min;
  const int blocks = static_cast<int>(std::min<long long>(148 * 8, (ncells + 255) / 256));
  const size_t smem = bin_count <= kHistSmemBins ? static_cast<size_t>(bin_count) * sizeof(unsigned) : 0;
  brick_histogram_kernel<<<blocks, 256, smem, as_stream(stream)>>>(
      data, jstride, kstride, nx, ny, nz, p, range_min, range_max, 1.0f / width, bin_count, counts_dev);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_brick_stats_final(const double* partials, int nparts, double* out4, void* stream,
                             int64_t* launches) {
  brick_stats_final_kernel<<<1, 256, 0, as_stream(stream)>>>(partials, nparts, out4);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_brick_quantise_batch(const QuantJob* jobs_dev, int njobs, long long max_cells,
                                const amrvr_scalar_transform& t, float range_min, float range_max,
                                void* stream, int64_t* launches) {
  if (njobs <= 0 || max_cells <= 0) return 0;
  // a quarter cell-quad per thread, grid-stride; x sized so the whole launch is about
  // 148 SMs x 8 resident CTAs x a few waves
  long long bx = (max_cells / 4 + 255) / 256;
  const long long cap = std::max<long long>(1, (148LL * 32) / njobs);
  if (bx > cap) bx = cap;
  for (int first = 0; first < njobs; first += 65535) {
    const int n = std::min(65535, njobs - first);
    brick_quantise_batch_kernel<<<dim3(static_cast<unsigned>(bx), static_cast<unsigned>(n)), 256, 0,
                                  as_stream(stream)>>>(jobs_dev + first,
                                                       make_value_path(t, range_min, range_max));
    ++*launches;
  }
  return static_cast<int>(cudaGetLastError());
}

int march_grid_size(bool count_samples) {
  static int cached[2] = {0, 0};
  int& grid = cached[count_samples ? 1 : 0];
  if (grid == 0) {
    int device = 0, sms = 148, per_sm = 1;
    cudaGetDevice(&device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (count_samples)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, raymarch_bricks_kernel<true>,
                                                    kMarchThreads, 0);
    else
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, raymarch_bricks_kernel<false>,
                                                    kMarchThreads, 0);
    grid = sms * (per_sm > 0 ? per_sm : 1);
  }
  return grid;
}

int launch_raymarch(const MarchFrame& frame, const MarchQueues& queues, const MarchBox* boxes_dev,
                    const uint8_t* bricks, const float* tables, bool count_samples, void* stream,
                    int64_t* launches) {
  if (frame.ntiles <= 0 && queues.nranks <= 0) return 0;
  // a sharded frame launches even without tiles of its own: the CTAs help with the
  // other queues and the last one raises this rank's "march finished" flag
  int grid = march_grid_size(count_samples);
  if (queues.nranks <= 0) grid = std::min(grid, frame.ntiles);
  if (count_samples) {
    raymarch_bricks_kernel<true><<<grid, kMarchThreads, 0, as_stream(stream)>>>(
        frame, queues, boxes_dev, bricks, reinterpret_cast<const float4*>(tables));
  } else {
    raymarch_bricks_kernel<false><<<grid, kMarchThreads, 0, as_stream(stream)>>>(
        frame, queues, boxes_dev, bricks, reinterpret_cast<const float4*>(tables));
  }
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

namespace {
// first row and number of 8-row bands of the rows a rank folds
void fold_rows(int width, int pixel_begin, int pixel_end, const FoldStrips* strips, int* first_row,
               int* bands) {
  if (strips != nullptr && strips->p > 1) {
    const int total = (strips->height + 15) / 16;
    const int mine = (total > strips->g) ? (total - 1 - strips->g) / strips->p + 1 : 0;
    *first_row = 16 * strips->g;
    *bands = 2 * mine;
    return;
  }
  *first_row = pixel_begin / width;
  *bands = (pixel_end > pixel_begin) ? ((pixel_end - 1) / width - *first_row) / kFoldBandRows + 1 : 0;
}
}  // namespace

size_t fold_scratch_bytes(int nlayers, int width, int pixel_begin, int pixel_end, const FoldStrips* strips) {
  if (width <= 0) return 0;
  int first_row = 0, bands = 0;
  fold_rows(width, pixel_begin, pixel_end, strips, &first_row, &bands);
  return static_cast<size_t>(bands) * (static_cast<size_t>(std::max(nlayers, 1)) + 3) * sizeof(int);
}

namespace {
struct FoldScratch {
  int first_row, bands;
  int2* band_xspan;
  int* band_counts;
  int* band_lists;
};
FoldScratch fold_scratch_layout(void* scratch, int width, int pixel_begin, int pixel_end,
                                const FoldStrips* strips) {
  FoldScratch L;
  fold_rows(width, pixel_begin, pixel_end, strips, &L.first_row, &L.bands);
  L.band_xspan = static_cast<int2*>(scratch);
  L.band_counts = reinterpret_cast<int*>(L.band_xspan + L.bands);
  L.band_lists = L.band_counts + L.bands;
  return L;
}
int fold_grid_size(bool bytes) {
  static int cached[2] = {0, 0};
  int& grid = cached[bytes ? 1 : 0];
  if (grid == 0) {
    int device = 0, sms = 148, per_sm = 1;
    cudaGetDevice(&device);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    if (bytes)
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fold_layers_kernel<true>, kFoldThreads, 0);
    else
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fold_layers_kernel<false>, kFoldThreads, 0);
    grid = sms * (per_sm > 0 ? per_sm : 1);
  }
  return grid;
}
}  // namespace

int launch_fold_bin(const FoldLayer* layers_dev, int nlayers, int width, int pixel_begin,
                    int pixel_end, void* scratch, void* stream, int64_t* launches,
                    const FoldStrips* strips) {
  const bool striped = strips != nullptr && strips->p > 1;
  if (!striped && pixel_end <= pixel_begin) return 0;
  const FoldScratch L = fold_scratch_layout(scratch, width, pixel_begin, pixel_end, strips);
  if (L.bands <= 0) return 0;
  fold_bin_kernel<<<L.bands, kBinThreads, 0, as_stream(stream)>>>(layers_dev, nlayers, L.first_row,
                                                                  striped ? strips->p : 1, L.band_lists,
                                                                  L.band_counts, L.band_xspan);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_fold(const FoldLayer* layers_dev, int nlayers, int width,
                int pixel_begin, int pixel_end, float* piece5, void* scratch, void* stream,
                int64_t* launches, const FoldBytes* bytes, const FoldSync* sync,
                const FoldStrips* strips) {
  const bool striped = strips != nullptr && strips->p > 1;
  if (striped) {  // whole rows; rows past the frame are masked by the pixel range
    pixel_begin = 0;
    pixel_end = width * strips->height;
  }
  const int npix = pixel_end - pixel_begin;
  const bool sharded = sync != nullptr && sync->nranks > 0;
  if (npix <= 0 && !sharded) return 0;
  FoldArgs A{};
  A.layers = layers_dev;
  A.nlayers = nlayers;
  A.width = width;
  A.pixel_begin = pixel_begin;
  A.pixel_end = pixel_end;
  A.strip_p = striped ? strips->p : 1;
  A.strip_g = striped ? strips->g : 0;
  if (npix > 0) {
    const FoldScratch L = fold_scratch_layout(scratch, width, pixel_begin, pixel_end, strips);
    A.first_row = L.first_row;
    A.bands = L.bands;
    A.band_xspan = L.band_xspan;
    A.band_counts = L.band_counts;
    A.band_lists = L.band_lists;
  }
  A.piece5 = piece5;
  if (sharded) A.sync = *sync;
  const bool as_bytes = bytes != nullptr && bytes->rgb8 != nullptr;
  const int tiles = A.bands * ((width + kFoldTileW - 1) / kFoldTileW);
  // an empty piece of a sharded frame still takes part in the hand-shake
  const int grid = std::max(1, std::min(fold_grid_size(as_bytes), tiles));
  if (as_bytes) {
    A.edges = bytes->edges;
    A.height = bytes->height;
    A.rgb8 = bytes->rgb8;
    fold_layers_kernel<true><<<grid, kFoldThreads, 0, as_stream(stream)>>>(A);
  } else {
    A.height = 0;
    fold_layers_kernel<false><<<grid, kFoldThreads, 0, as_stream(stream)>>>(A);
  }
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_peer_wait(const PeerPtrs& ctl, int me, int nranks, int row, uint32_t epoch,
                     unsigned long long timeout_ns, void* stream, int64_t* launches) {
  if (nranks > kMaxPeers || row < 0 || row >= kFlagRows) return static_cast<int>(cudaErrorInvalidValue);
  peer_wait_kernel<<<1, 32, 0, as_stream(stream)>>>(ctl, me, nranks, row, epoch, timeout_ns);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_peer_copy(const void* src, void* dst, size_t bytes, int blocks, void* stream,
                     int64_t* launches) {
  if (bytes < 16) return 0;
  peer_copy_kernel<<<blocks, 256, 0, as_stream(stream)>>>(static_cast<const uint4*>(src),
                                                          static_cast<uint4*>(dst), bytes / 16);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_pack(const CopyJob* jobs_dev, int njobs, const float* src, float* dst, void* stream,
                int64_t* launches) {
  if (njobs <= 0) return 0;
  pack_slices_kernel<<<njobs, 256, 0, as_stream(stream)>>>(jobs_dev, src, dst);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_push(const CopyJob* jobs_dev, int njobs, const float* src, const PeerPtrs& recv, void* stream,
                int64_t* launches) {
  if (njobs <= 0) return 0;
  push_slices_kernel<<<njobs, 256, 0, as_stream(stream)>>>(jobs_dev, src, recv);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_peer_barrier(const PeerPtrs& flags, int me, int nranks, int which, uint32_t epoch,
                        unsigned long long timeout_ns, uint32_t* error_flag, void* stream,
                        int64_t* launches) {
  if (nranks > kMaxPeers || which < 0 || which >= kFlagRows) return static_cast<int>(cudaErrorInvalidValue);
  peer_barrier_kernel<<<1, 32, 0, as_stream(stream)>>>(flags, me, nranks, which, epoch, timeout_ns,
                                                       error_flag);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_post_piece(const float* piece5, int pixel_begin, int pixel_end, int width, int height,
                      const OverlayEdges* edges, uint8_t* rgb8_root, void* stream,
                      int64_t* launches) {
  const int n = pixel_end - pixel_begin;
  if (n <= 0) return 0;
  const int first = pixel_begin & ~3;
  post_piece_kernel<<<(pixel_end - first + 255) / 256, 256, 0, as_stream(stream)>>>(
      piece5, pixel_begin, pixel_end, width, height, edges, 0.5f, rgb8_root);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_downsample(const float* src5, int width, int height, int sqrt_aa, float* dst5,
                      void* stream, int64_t* launches) {
  const int n = width * height;
  if (n <= 0) return 0;
  downsample_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(src5, width, height,
                                                                    sqrt_aa > 1 ? sqrt_aa : 1, dst5);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_overlay(const OverlayEdges& edges, int sqrt_aa, int width, int height, float* image5,
                   void* stream, int64_t* launches) {
  const int n = width * height;
  if (n <= 0 || edges.count <= 0) return 0;
  const float radius = 0.5f * static_cast<float>(sqrt_aa > 1 ? sqrt_aa : 1);
  overlay_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(edges, radius, width, height,
                                                                 image5);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_quantize_rgb8(const float* image5, int width, int height, uint8_t* rgb8, void* stream,
                         int64_t* launches) {
  const int n = width * height;
  if (n <= 0) return 0;
  quantize_rgb8_kernel<<<(n + 255) / 256, 256, 0, as_stream(stream)>>>(image5, width, height, rgb8);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_order_rank(const OrderBox* boxes_dev, int n, void* scratch_a, void* scratch_b,
                      int owner_bits, int32_t* priority_dev, void* stream, int64_t* launches) {
  if (n <= 0) return 0;
  order_rank_kernel<<<1, kSortThreads, 0, as_stream(stream)>>>(
      boxes_dev, n, static_cast<uint4*>(scratch_a), static_cast<uint4*>(scratch_b), owner_bits,
      priority_dev);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_order_count(const OrderBox* boxes_dev, int n, int32_t* priority_dev, void* stream,
                       int64_t* launches) {
  if (n <= 0) return 0;
  const unsigned blocks = static_cast<unsigned>((n + 255) / 256);
  order_count_kernel<<<dim3(blocks, blocks), 256, 0, as_stream(stream)>>>(boxes_dev, n, priority_dev);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

int launch_order_pairs(const OrderBox* boxes_dev, int n, float dx, float dy, float dz,
                       OrderEdge* edges_dev, int capacity, int32_t* count_dev, void* stream,
                       int64_t* launches) {
  if (n <= 1) return 0;
  dim3 grid((n + 255) / 256, n);
  order_pairs_kernel<<<grid, 256, 0, as_stream(stream)>>>(boxes_dev, n, dx, dy, dz, edges_dev,
                                                          capacity, count_dev);
  ++*launches;
  return static_cast<int>(cudaGetLastError());
}

}  // namespace amrvr
