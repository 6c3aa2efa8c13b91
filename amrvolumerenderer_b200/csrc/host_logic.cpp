// host_logic.cpp -- the sequential, per-frame host side of the hot path: transfer
// table, camera frame, per-box sampling constants, screen footprints, layer
// order, visibility-graph driver.  Everything here is O(boxes) or O(256) work
// the reference also does on the host; the per-pixel and per-cell work lives in
// kernels.cu.  Compiled without FP contraction so float expressions round
// exactly like the reference's default x86-64 build.
#include "internal.hpp"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <queue>
#if defined(__x86_64__) && defined(__GNUC__)
#include <immintrin.h>
#endif

namespace amrvr {

namespace {

constexpr float kPiF = 3.14159265358979323846f;

inline float clampf(float v, float lo, float hi) {
  return (v < lo) ? lo : ((hi < v) ? hi : v);
}
inline float fmax2(float a, float b) { return (a < b) ? b : a; }  // std::max
inline float fmin2(float a, float b) { return (b < a) ? b : a; }  // std::min

struct V3 {
  double x, y, z;
};
inline V3 sub(const double* a, const double* b) {
  return {a[0] - b[0], a[1] - b[1], a[2] - b[2]};
}
inline V3 cross(const V3& a, const V3& b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline double dot(const V3& a, const V3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// reference: Common/CameraUtils.hpp:16-22 (fallback is always (0,0,-1))
inline V3 unit_or(const V3& v, const V3& fallback) {
  const double len = std::sqrt(dot(v, v));
  if (len > 0.0 && std::isfinite(len)) return {v.x / len, v.y / len, v.z / len};
  return fallback;
}

struct Frame3 {
  V3 f, r, u;
};
// forward / right / true-up, all in double (Common/VolumePainter.cpp:631-641,
// Common/CameraUtils.hpp:27-39)
Frame3 camera_frame(const amrvr_camera& cam) {
  Frame3 fr;
  fr.f = unit_or(sub(cam.look_at, cam.eye), V3{0.0, 0.0, -1.0});
  const V3 up{cam.up[0], cam.up[1], cam.up[2]};
  V3 r = cross(fr.f, up);
  const double rl = std::sqrt(dot(r, r));
  if (rl > 0.0 && std::isfinite(rl)) {
    r.x /= rl;
    r.y /= rl;
    r.z /= rl;
  } else {
    r = V3{1.0, 0.0, 0.0};
  }
  fr.r = r;
  fr.u = cross(fr.r, fr.f);
  return fr;
}

}  // namespace

void make_camera_basis(const amrvr_camera& cam, CameraBasis& out) {
  const Frame3 fr = camera_frame(cam);
  const double f[3] = {fr.f.x, fr.f.y, fr.f.z};
  const double r[3] = {fr.r.x, fr.r.y, fr.r.z};
  const double u[3] = {fr.u.x, fr.u.y, fr.u.z};
  for (int d = 0; d < 3; ++d) {
    out.fwd_d[d] = f[d];
    out.right_d[d] = r[d];
    out.up_d[d] = u[d];
    out.fwd[d] = static_cast<float>(f[d]);
    out.right[d] = static_cast<float>(r[d]);
    out.up[d] = static_cast<float>(u[d]);
    out.eye[d] = static_cast<float>(cam.eye[d]);
  }
}

// ---------------------------------------------------------------------------
// Transfer table.  Two piecewise functions over the scalar range: a colour
// curve (RGB nodes; interpolated in RGB for the built-in map, in CIELAB for a
// user map) and an opacity curve whose node alphas were already corrected for
// the box's step length.  Reference: Common/VolumePainter.cpp:107-516.
// ---------------------------------------------------------------------------
namespace {

struct ColorNode {
  float v, r, g, b;
};
struct OpacityNode {
  float v, a, mid, sharp;
};

struct TransferCurves {
  bool lab = false;
  std::vector<ColorNode> color;
  std::vector<OpacityNode> opacity;

  template <class Node>
  static void put(std::vector<Node>& nodes, const Node& n) {
    auto it = std::lower_bound(nodes.begin(), nodes.end(), n,
                               [](const Node& a, const Node& b) { return a.v < b.v; });
    if (it != nodes.end() && it->v == n.v) {
      *it = n;  // same abscissa replaces (:127-151)
    } else {
      nodes.insert(it, n);
    }
  }

  // :153-200 -- stretch all abscissae from their current span to [lo, hi]
  void stretch_to(float lo, float hi) {
    bool any = false;
    float cur_lo = 0.0f, cur_hi = 0.0f;
    auto see = [&](float v) {
      if (!any) {
        cur_lo = cur_hi = v;
        any = true;
      } else {
        cur_lo = fmin2(cur_lo, v);
        cur_hi = fmax2(cur_hi, v);
      }
    };
    for (const auto& c : color) see(c.v);
    for (const auto& o : opacity) see(o.v);
    const float old_span = cur_hi - cur_lo;
    const float new_span = hi - lo;
    if (!(old_span > 0.0f) || !(new_span > 0.0f)) return;
    for (auto& c : color) c.v = lo + new_span * ((c.v - cur_lo) / old_span);
    for (auto& o : opacity) o.v = lo + new_span * ((o.v - cur_lo) / old_span);
  }
};

// :107-125 -- opacity correction for the step length, in double
float step_corrected_alpha(float base, float alpha_scale, float norm) {
  const float s = clampf(base * alpha_scale, 0.0f, 1.0f);
  if (norm <= 0.0f || s <= 0.0f) return 0.0f;
  if (s >= 1.0f) return 1.0f;
  const double remaining = std::pow(1.0 - static_cast<double>(s), static_cast<double>(norm));
  float a = static_cast<float>(1.0 - remaining);
  if (!std::isfinite(a)) a = s;
  return clampf(a, 0.0f, 1.0f);
}

// sRGB <-> CIELAB (D65), float, with the reference's constants (:202-320)
struct Lab {
  float L, a, b;
};
inline float gamma_expand(float c) {
  return (c > 0.04045f) ? std::pow((c + 0.055f) / 1.055f, 2.4f) : c / 12.92f;
}
inline float gamma_compress(float c) {
  return (c > 0.0031308f) ? 1.055f * std::pow(c, 1.0f / 2.4f) - 0.055f : 12.92f * c;
}
inline float lab_curve(float t) {
  return (t > 0.008856f) ? std::pow(t, 1.0f / 3.0f) : (7.787f * t) + 16.0f / 116.0f;
}
inline float lab_curve_inv(float t) {
  const float cube = std::pow(t, 3.0f);
  return (cube > 0.008856f) ? cube : (t - 16.0f / 116.0f) / 7.787f;
}
Lab to_lab(float r8, float g8, float b8) {
  const float r = gamma_expand(r8), g = gamma_expand(g8), b = gamma_expand(b8);
  const float X = r * 0.4124f + g * 0.3576f + b * 0.1805f;
  const float Y = r * 0.2126f + g * 0.7152f + b * 0.0722f;
  const float Z = r * 0.0193f + g * 0.1192f + b * 0.9505f;
  const float fx = lab_curve(X / 0.9505f);
  const float fy = lab_curve(Y / 1.0f);
  const float fz = lab_curve(Z / 1.089f);
  return {(116.0f * fy) - 16.0f, 500.0f * (fx - fy), 200.0f * (fy - fz)};
}
void from_lab(const Lab& c, float* rgb) {
  float fy = (c.L + 16.0f) / 116.0f;
  float fx = c.a / 500.0f + fy;
  float fz = fy - c.b / 200.0f;
  float X = lab_curve_inv(fx), Y = lab_curve_inv(fy), Z = lab_curve_inv(fz);
  X *= 0.9505f;
  Y *= 1.0f;
  Z *= 1.089f;
  float r = X * 3.2406f + Y * -1.5372f + Z * -0.4986f;
  float g = X * -0.9689f + Y * 1.8758f + Z * 0.0415f;
  float b = X * 0.0557f + Y * -0.2040f + Z * 1.0570f;
  r = gamma_compress(r);
  g = gamma_compress(g);
  b = gamma_compress(b);
  const float peak = fmax2(r, fmax2(g, b));
  if (peak > 1.0f) {
    r /= peak;
    g /= peak;
    b /= peak;
  }
  rgb[0] = fmax2(r, 0.0f);
  rgb[1] = fmax2(g, 0.0f);
  rgb[2] = fmax2(b, 0.0f);
}

// colour curve at `value`, clamping outside the node span (:331-379)
void eval_color(const TransferCurves& tc, float value, float* rgb) {
  if (!std::isfinite(value)) {
    rgb[0] = tc.lab ? 1.0f : 0.25f;
    rgb[1] = rgb[2] = 0.0f;
    return;
  }
  if (tc.color.empty()) {
    rgb[0] = rgb[1] = rgb[2] = 0.0f;
    return;
  }
  const ColorNode& first = tc.color.front();
  const ColorNode& last = tc.color.back();
  const ColorNode* edge = nullptr;
  if (value < first.v) {
    edge = &first;
  } else if (value > last.v) {
    edge = &last;
  } else if (value == first.v) {
    edge = &first;
  } else if (value == last.v) {
    edge = &last;
  }
  if (edge == nullptr) {
    for (size_t i = 1; i < tc.color.size(); ++i) {
      const ColorNode& hi = tc.color[i];
      if (!(hi.v >= value)) continue;
      const ColorNode& lo = tc.color[i - 1];
      const float span = hi.v - lo.v;
      const float w = (span > 0.0f) ? (value - lo.v) / span : 0.0f;
      if (tc.lab) {
        const Lab a = to_lab(lo.r, lo.g, lo.b);
        const Lab b = to_lab(hi.r, hi.g, hi.b);
        const Lab m{a.L + (b.L - a.L) * w, a.a + (b.a - a.a) * w, a.b + (b.b - a.b) * w};
        from_lab(m, rgb);
      } else {
        rgb[0] = lo.r + (hi.r - lo.r) * w;
        rgb[1] = lo.g + (hi.g - lo.g) * w;
        rgb[2] = lo.b + (hi.b - lo.b) * w;
      }
      return;
    }
    edge = &last;
  }
  rgb[0] = edge->r;
  rgb[1] = edge->g;
  rgb[2] = edge->b;
}

// opacity curve at `value` (:381-440): piecewise with midpoint / sharpness
float eval_opacity(const TransferCurves& tc, float value) {
  if (!std::isfinite(value) || tc.opacity.empty()) return 1.0f;
  const OpacityNode& first = tc.opacity.front();
  const OpacityNode& last = tc.opacity.back();
  if (value <= first.v) return first.a;
  if (value >= last.v) return last.a;
  for (size_t i = 1; i < tc.opacity.size(); ++i) {
    const OpacityNode& hi = tc.opacity[i];
    if (!(hi.v >= value)) continue;
    const OpacityNode& lo = tc.opacity[i - 1];
    const float span = hi.v - lo.v;
    float w = (span > 0.0f) ? (value - lo.v) / span : 0.0f;
    if (w < lo.mid) {
      w = 0.5f * w / lo.mid;
    } else {
      w = 0.5f + 0.5f * (w - lo.mid) / (1.0f - lo.mid);
    }
    if (lo.sharp == 1.0f) return (w < 0.5f) ? lo.a : hi.a;
    if (lo.sharp == 0.0f) return lo.a + (hi.a - lo.a) * w;
    if (w < 0.5f) {
      w = 0.5f * std::pow(w * 2.0f, 1.0f + 10.0f * lo.sharp);
    } else if (w > 0.5f) {
      w = 1.0f - 0.5f * std::pow((1.0f - w) * 2.0f, 1.0f + 10.0f * lo.sharp);
    }
    const float w2 = w * w, w3 = w2 * w;
    const float h00 = 2.0f * w3 - 3.0f * w2 + 1.0f;
    const float h01 = -2.0f * w3 + 3.0f * w2;
    const float h10 = w3 - 2.0f * w2 + w;
    const float h11 = w3 - w2;
    const float tangent = (1.0f - lo.sharp) * (hi.a - lo.a);
    float a = h00 * lo.a + h01 * hi.a + h10 * tangent + h11 * tangent;
    a = fmax2(a, fmin2(lo.a, hi.a));
    a = fmin2(a, fmax2(lo.a, hi.a));
    return a;
  }
  return last.a;
}

}  // namespace

void build_color_table(float alpha_scale, float norm_factor, float range_min,
                       float range_max, const amrvr_colormap_pt* pts, int npts,
                       float* rgba256) {
  TransferCurves tc;
  const float span = range_max - range_min;
  if (pts != nullptr && npts > 0) {  // user map (:449-466)
    tc.lab = true;
    for (int i = 0; i < npts; ++i) {
      const amrvr_colormap_pt& p = pts[i];
      TransferCurves::put(tc.color, ColorNode{p.value, clampf(p.red, 0.0f, 1.0f),
                                              clampf(p.green, 0.0f, 1.0f),
                                              clampf(p.blue, 0.0f, 1.0f)});
      TransferCurves::put(
          tc.opacity,
          OpacityNode{p.value, step_corrected_alpha(p.alpha, alpha_scale, norm_factor),
                      0.5f, 0.0f});
    }
  } else {  // built-in jet + opacity ramp (:467-500)
    static const ColorNode kJet[] = {
        {0.0f, 0.0f, 0.0f, 0.5625f},    {0.111111f, 0.0f, 0.0f, 1.0f},
        {0.3650795f, 0.0f, 1.0f, 1.0f}, {0.4920635f, 0.5f, 1.0f, 0.5f},
        {0.6190475f, 1.0f, 1.0f, 0.0f}, {0.873016f, 1.0f, 0.0f, 0.0f},
        {1.0f, 0.5f, 0.0f, 0.0f}};
    for (const ColorNode& n : kJet) TransferCurves::put(tc.color, n);
    static const float kRampAt[] = {0.0f, 0.15f, 0.35f, 0.6f, 0.85f, 1.0f};
    static const float kRampAlpha[] = {0.05f, 0.15f, 0.22f, 0.3f, 0.38f, 0.5f};
    for (int i = 0; i < 6; ++i) {
      TransferCurves::put(
          tc.opacity,
          OpacityNode{kRampAt[i] * span + range_min,
                      step_corrected_alpha(kRampAlpha[i], alpha_scale, norm_factor),
                      0.5f, 0.0f});
    }
    tc.stretch_to(range_min, range_max);
  }
  for (int i = 0; i < AMRVR_COLOR_TABLE_SIZE; ++i) {  // :507-514
    const float t = static_cast<float>(i) / static_cast<float>(AMRVR_COLOR_TABLE_SIZE - 1);
    const float value = range_min + span * t;
    eval_color(tc, value, rgba256 + 4 * i);
    rgba256[4 * i + 3] = eval_opacity(tc, value);
  }
}

void box_sampling(const double* min_corner, const double* max_corner,
                  const int32_t* dims, float reference_sample_distance,
                  float* step_out, float* norm_out) {
  float finest = FLT_MAX;
  for (int d = 0; d < 3; ++d) {
    double edge = 0.0;
    if (dims[d] > 0) edge = (max_corner[d] - min_corner[d]) / static_cast<double>(dims[d]);
    const float e = static_cast<float>(edge);
    if (e > 0.0f && e < finest && std::isfinite(e)) finest = e;
  }
  if (!(finest > 0.0f && std::isfinite(finest))) {
    // paint() is handed the box itself as `bounds` fallback (:593-598)
    double shortest = max_corner[0] - min_corner[0];
    for (int d = 1; d < 3; ++d) shortest = std::min(shortest, max_corner[d] - min_corner[d]);
    finest = fmax2(1e-4f, static_cast<float>(shortest) * 0.01f);
  }
  const float step = fmax2(finest * 0.5f, 1e-5f);
  float ref = reference_sample_distance;
  if (!(ref > 0.0f && std::isfinite(ref))) ref = step;
  float norm = step / ref;
  if (!std::isfinite(norm)) norm = 1.0f;
  *step_out = step;
  *norm_out = fmax2(norm, 0.0f);
}

float box_depth_hint(const amrvr_box_record& rec, const amrvr_camera& cam) {
  const V3 dir = unit_or(sub(cam.look_at, cam.eye), V3{0.0, 0.0, -1.0});
  float nearest = INFINITY;
  for (int c = 0; c < 8; ++c) {
    const double p[3] = {(c & 1) ? rec.max_corner[0] : rec.min_corner[0],
                         (c & 2) ? rec.max_corner[1] : rec.min_corner[1],
                         (c & 4) ? rec.max_corner[2] : rec.min_corner[2]};
    const float along = static_cast<float>(dot(sub(p, cam.eye), dir));
    nearest = fmin2(nearest, along);
  }
  return nearest;
}

void layer_order(int nglobal, const amrvr_box_record* recs, const amrvr_camera& cam,
                 float* hints, int* order) {
  std::vector<float> depth(static_cast<size_t>(std::max(nglobal, 0)));
  for (int i = 0; i < nglobal; ++i) {
    depth[i] = box_depth_hint(recs[i], cam);
    if (hints != nullptr) hints[i] = depth[i];
    order[i] = i;
  }
  std::sort(order, order + std::max(nglobal, 0), [&](int a, int b) {
    if (depth[a] == depth[b]) {
      if (recs[a].owner == recs[b].owner) return recs[a].local_index < recs[b].local_index;
      return recs[a].owner < recs[b].owner;
    }
    return depth[a] < depth[b];
  });
}

void stripe_plan(int p, int nboxes, const int32_t* ry0, const int32_t* rh, const int32_t* rw,
                 QueueItem* qitems, int32_t* tile_begin, uint64_t* pool_floats) {
  // per box, once: first strip k0, k0 % p, and how its nk strips deal out over p queues
  // (queue q gets base, plus one when it is among the first `rem` after k0 % p) -- the
  // same numbers as stripe_first_strip / stripe_count, without a division per (queue, box)
  struct PerBox {
    int32_t k0, r0, base, rem, tiles_x;
    uint64_t strip_floats;
  };
  std::vector<PerBox> pb(static_cast<size_t>(std::max(nboxes, 0)));
  for (int b = 0; b < nboxes; ++b) {
    const int k0 = ry0[b] >> 4, k1 = (ry0[b] + rh[b] - 1) >> 4;
    const int nk = (rh[b] > 0) ? k1 - k0 + 1 : 0;
    pb[static_cast<size_t>(b)] = PerBox{k0, k0 % p, nk / p, nk % p, (rw[b] + 15) / 16,
                                        static_cast<uint64_t>(16) * rw[b] * 5};
  }
  int tiles = 0;
  for (int q = 0; q < p; ++q) {
    tile_begin[q] = tiles;
    uint64_t floats = 0;
    QueueItem* row = qitems + static_cast<size_t>(q) * nboxes;
    for (int b = 0; b < nboxes; ++b) {
      const PerBox& B = pb[static_cast<size_t>(b)];
      int a = q - B.r0;  // how many strips after k0 the queue's first one lies
      if (a < 0) a += p;
      const int ns = B.base + (a < B.rem ? 1 : 0);
      row[b].tile_base = tiles;
      row[b].ry_first = 16 * (B.k0 + a);
      row[b].layer_off = floats;
      tiles += ns * B.tiles_x;
      floats += static_cast<uint64_t>(ns) * B.strip_floats;  // whole strips, 16 rows each
    }
    pool_floats[q] = floats;
  }
  tile_begin[p] = tiles;
}

void piece_range(int num_pixels, int piece, int num_pieces, int* begin, int* end) {
  const int each = num_pixels / num_pieces;
  *begin = each * piece;
  *end = (piece < num_pieces - 1) ? (*begin + each) : num_pixels;
}

float reference_sample_distance(int nglobal, const amrvr_box_record* recs,
                                const double* bounds_min, const double* bounds_max) {
  float coarsest = 0.0f;
  for (int i = 0; i < nglobal; ++i) {
    float finest = FLT_MAX;
    for (int d = 0; d < 3; ++d) {
      double edge = 0.0;  // Real / float promotes to double (:1146-1154)
      if (recs[i].dims[d] > 0)
        edge = (recs[i].max_corner[d] - recs[i].min_corner[d]) /
               static_cast<float>(recs[i].dims[d]);
      if (edge > 0.0f && edge < finest && std::isfinite(edge)) finest = static_cast<float>(edge);
    }
    if (finest > 0.0f && std::isfinite(finest)) coarsest = fmax2(coarsest, finest);
  }
  if (!(coarsest > 0.0f && std::isfinite(coarsest))) {
    float shortest = FLT_MAX;
    for (int d = 0; d < 3; ++d) {
      const float len = static_cast<float>(bounds_max[d] - bounds_min[d]);
      if (len > 0.0f && std::isfinite(len)) shortest = fmin2(shortest, len);
    }
    if (!(shortest > 0.0f && std::isfinite(shortest))) shortest = 1.0f;
    coarsest = fmax2(1e-4f, shortest * 0.01f);
  }
  return fmax2(coarsest * 0.5f, 1e-5f);
}

// ---------------------------------------------------------------------------
// Screen footprint
// ---------------------------------------------------------------------------
namespace {
Rect footprint_core(const double* min_corner, const double* max_corner, const int32_t* dims,
                    const CameraBasis& basis, double tan_half, double aspect, int width, int height);
}  // namespace

Rect box_footprint(const double* min_corner, const double* max_corner,
                   const int32_t* dims, const CameraBasis& basis,
                   const amrvr_camera& cam, int width, int height) {
  const float aspect_f = static_cast<float>(width) / static_cast<float>(std::max(height, 1));
  const double tan_half = std::tan(cam.fov_y_degrees * 0.5f * kPiF / 180.0f);
  return footprint_core(min_corner, max_corner, dims, basis, tan_half, aspect_f, width, height);
}

// Footprints of a whole box table: the camera constants are evaluated once.
void box_footprints(int n, const amrvr_box_record* recs, const CameraBasis& basis,
                    const amrvr_camera& cam, int width, int height, Rect* out) {
  const float aspect_f = static_cast<float>(width) / static_cast<float>(std::max(height, 1));
  const double tan_half = std::tan(cam.fov_y_degrees * 0.5f * kPiF / 180.0f);
  for (int g = 0; g < n; ++g)
    out[g] = footprint_core(recs[g].min_corner, recs[g].max_corner, recs[g].dims, basis, tan_half,
                            aspect_f, width, height);
}

namespace {
Rect footprint_core(const double* min_corner, const double* max_corner, const int32_t* dims,
                    const CameraBasis& basis, double tan_half, double aspect, int width, int height) {
  Rect full{0, 0, (width + 3) & ~3, height};
  if (dims[0] <= 0 || dims[1] <= 0 || dims[2] <= 0 || width <= 0 || height <= 0)
    return Rect{0, 0, 0, 0};
  // the march sees float-truncated corners and a float eye
  double lo[3], hi[3], eye[3];
  double diag2 = 0.0;
  for (int d = 0; d < 3; ++d) {
    lo[d] = static_cast<float>(min_corner[d]);
    hi[d] = static_cast<float>(max_corner[d]);
    eye[d] = basis.eye[d];
    diag2 += (hi[d] - lo[d]) * (hi[d] - lo[d]);
  }
  const double guard = 1e-3 * std::sqrt(diag2) + 1e-6;
  double px_lo = INFINITY, px_hi = -INFINITY, py_lo = INFINITY, py_hi = -INFINITY;
  for (int c = 0; c < 8; ++c) {
    const double p[3] = {(c & 1) ? hi[0] : lo[0], (c & 2) ? hi[1] : lo[1],
                         (c & 4) ? hi[2] : lo[2]};
    double depth = 0.0, xr = 0.0, yu = 0.0;
    for (int d = 0; d < 3; ++d) {
      const double rel = p[d] - eye[d];
      depth += rel * basis.fwd[d];
      xr += rel * basis.right[d];
      yu += rel * basis.up[d];
    }
    if (!(depth > guard) || !std::isfinite(depth)) return full;  // corner at/behind the eye plane
    const double ndc_x = xr / (depth * tan_half * aspect);
    const double ndc_y = yu / (depth * tan_half);
    if (!std::isfinite(ndc_x) || !std::isfinite(ndc_y)) return full;
    const double px = (ndc_x + 1.0) * 0.5 * width - 0.5;
    const double py = (ndc_y + 1.0) * 0.5 * height - 0.5;
    px_lo = std::min(px_lo, px);
    px_hi = std::max(px_hi, px);
    py_lo = std::min(py_lo, py);
    py_hi = std::max(py_hi, py);
  }
  // clamp in double before converting (far off-screen boxes)
  const double big = 1e9;
  px_lo = std::max(px_lo, -big);
  py_lo = std::max(py_lo, -big);
  px_hi = std::min(px_hi, big);
  py_hi = std::min(py_hi, big);
  long x_lo = static_cast<long>(std::floor(px_lo)) - 1;
  long x_hi = static_cast<long>(std::ceil(px_hi)) + 1;
  long y_lo = static_cast<long>(std::floor(py_lo)) - 1;
  long y_hi = static_cast<long>(std::ceil(py_hi)) + 1;
  x_lo = std::max<long>(x_lo, 0);
  y_lo = std::max<long>(y_lo, 0);
  x_hi = std::min<long>(x_hi, width - 1);
  y_hi = std::min<long>(y_hi, height - 1);
  if (x_lo > x_hi || y_lo > y_hi) return Rect{0, 0, 0, 0};
  Rect r;
  r.x0 = static_cast<int32_t>(x_lo) & ~3;
  r.y0 = static_cast<int32_t>(y_lo);
  r.w = (static_cast<int32_t>(x_hi) + 1 - r.x0 + 3) & ~3;
  r.h = static_cast<int32_t>(y_hi - y_lo + 1);
  return r;
}
}  // namespace

// ---------------------------------------------------------------------------
// Visibility ordering
// ---------------------------------------------------------------------------
void make_order_matrices(const amrvr_camera& cam, float aspect, OrderMatrices& m) {
  const Frame3 fr = camera_frame(cam);
  const V3 eye{cam.eye[0], cam.eye[1], cam.eye[2]};
  float* v = m.view;
  std::memset(v, 0, sizeof(m.view));
  // Common/CameraUtils.hpp:41-60 writes basis vectors down the columns and the
  // translation into the LAST ROW; reproduce the element positions, not the
  // textbook matrix.
  v[0 * 4 + 0] = static_cast<float>(fr.r.x);
  v[1 * 4 + 0] = static_cast<float>(fr.r.y);
  v[2 * 4 + 0] = static_cast<float>(fr.r.z);
  v[3 * 4 + 0] = static_cast<float>(-dot(fr.r, eye));
  v[0 * 4 + 1] = static_cast<float>(fr.u.x);
  v[1 * 4 + 1] = static_cast<float>(fr.u.y);
  v[2 * 4 + 1] = static_cast<float>(fr.u.z);
  v[3 * 4 + 1] = static_cast<float>(-dot(fr.u, eye));
  v[0 * 4 + 2] = static_cast<float>(-fr.f.x);
  v[1 * 4 + 2] = static_cast<float>(-fr.f.y);
  v[2 * 4 + 2] = static_cast<float>(-fr.f.z);
  v[3 * 4 + 2] = static_cast<float>(dot(fr.f, eye));
  v[3 * 4 + 3] = 1.0f;

  float* p = m.proj;  // Common/VisibilityOrdering.cpp:34-59
  std::memset(p, 0, sizeof(m.proj));
  const float near_p = cam.near_plane, far_p = cam.far_plane;
  const float half = near_p * std::tan(cam.fov_y_degrees * kPiF / 180.0f * 0.5f);
  const float left = -half * aspect, right = half * aspect;
  const float bottom = -half, top = half;
  p[0 * 4 + 0] = 2.0f * near_p / (right - left);
  p[1 * 4 + 1] = 2.0f * near_p / (top - bottom);
  p[0 * 4 + 2] = (right + left) / (right - left);
  p[1 * 4 + 2] = (top + bottom) / (top - bottom);
  p[2 * 4 + 2] = -(far_p + near_p) / (far_p - near_p);
  p[3 * 4 + 2] = -1.0f;
  p[2 * 4 + 3] = -(2.0f * far_p * near_p) / (far_p - near_p);
  p[3 * 4 + 3] = 0.0f;

  m.dir[0] = static_cast<float>(fr.f.x);
  m.dir[1] = static_cast<float>(fr.f.y);
  m.dir[2] = static_cast<float>(fr.f.z);
}

namespace {
inline void mul4(const float* a, const float* x, float* y) {
  for (int r = 0; r < 4; ++r) {
    float s = 0.0f;
    for (int k = 0; k < 4; ++k) s += a[r * 4 + k] * x[k];
    y[r] = s;
  }
}
}  // namespace

void order_box_depth_range(const OrderMatrices& m, OrderBox& b) {
  float zmin = INFINITY, zmax = -INFINITY;
  for (int c = 0; c < 8; ++c) {
    const float h[4] = {(c & 1) ? b.hi[0] : b.lo[0], (c & 2) ? b.hi[1] : b.lo[1],
                        (c & 4) ? b.hi[2] : b.lo[2], 1.0f};
    float eye_space[4], clip[4];
    mul4(m.view, h, eye_space);
    for (int r = 2; r < 4; ++r) {  // only z and w of the clip-space point are used
      float s = 0.0f;
      for (int k = 0; k < 4; ++k) s += m.proj[r * 4 + k] * eye_space[k];
      clip[r] = s;
    }
    if (clip[3] != 0.0f) {
      const float z = clip[2] / clip[3];
      zmin = fmin2(zmin, z);
      zmax = fmax2(zmax, z);
    }
  }
  if (!std::isfinite(zmin) || !std::isfinite(zmax)) zmin = zmax = INFINITY;
  b.min_depth = zmin;
  b.max_depth = zmax;
}

// All boxes of the table.  On x86 the eight corners of a box go through the two
// matrices four at a time (SSE2): every lane performs the scalar code's
// operations in the scalar code's order (no contraction, zeros included), so
// each corner's z is bit-identical; only the min/max reduction is re-associated,
// which cannot change a finite or infinite result.
void order_box_depth_ranges(const OrderMatrices& m, OrderBox* boxes, int n) {
#if defined(__x86_64__) && defined(__GNUC__)
  const __m128 zero = _mm_setzero_ps(), one = _mm_set1_ps(1.0f);
  const __m128 inf = _mm_set1_ps(INFINITY), ninf = _mm_set1_ps(-INFINITY);
  for (int i = 0; i < n; ++i) {
    OrderBox& b = boxes[i];
    const __m128 X = _mm_setr_ps(b.lo[0], b.hi[0], b.lo[0], b.hi[0]);  // corners c = 0..3 (and 4..7)
    const __m128 Y = _mm_setr_ps(b.lo[1], b.lo[1], b.hi[1], b.hi[1]);
    const __m128 Z[2] = {_mm_set1_ps(b.lo[2]), _mm_set1_ps(b.hi[2])};
    __m128 txy[4];
    for (int r = 0; r < 4; ++r)
      txy[r] = _mm_add_ps(_mm_add_ps(zero, _mm_mul_ps(_mm_set1_ps(m.view[r * 4 + 0]), X)),
                          _mm_mul_ps(_mm_set1_ps(m.view[r * 4 + 1]), Y));
    __m128 lo_acc = inf, hi_acc = ninf;
    for (int half = 0; half < 2; ++half) {
      __m128 e[4];
      for (int r = 0; r < 4; ++r)
        e[r] = _mm_add_ps(_mm_add_ps(txy[r], _mm_mul_ps(_mm_set1_ps(m.view[r * 4 + 2]), Z[half])),
                          _mm_mul_ps(_mm_set1_ps(m.view[r * 4 + 3]), one));
      __m128 clip[2];
      for (int r = 2; r < 4; ++r) {
        __m128 sum = zero;
        for (int k = 0; k < 4; ++k) sum = _mm_add_ps(sum, _mm_mul_ps(_mm_set1_ps(m.proj[r * 4 + k]), e[k]));
        clip[r - 2] = sum;
      }
      const __m128 z = _mm_div_ps(clip[0], clip[1]);
      const __m128 use = _mm_cmpneq_ps(clip[1], zero);  // clip.w != 0 (true for NaN, as in C++)
      const __m128 lo_new = _mm_min_ps(z, lo_acc), hi_new = _mm_max_ps(z, hi_acc);
      lo_acc = _mm_or_ps(_mm_and_ps(use, lo_new), _mm_andnot_ps(use, lo_acc));
      hi_acc = _mm_or_ps(_mm_and_ps(use, hi_new), _mm_andnot_ps(use, hi_acc));
    }
    float lo4[4], hi4[4];
    _mm_storeu_ps(lo4, lo_acc);
    _mm_storeu_ps(hi4, hi_acc);
    float zmin = INFINITY, zmax = -INFINITY;
    for (int k = 0; k < 4; ++k) {
      zmin = fmin2(zmin, lo4[k]);
      zmax = fmax2(zmax, hi4[k]);
    }
    if (!std::isfinite(zmin) || !std::isfinite(zmax)) zmin = zmax = INFINITY;
    b.min_depth = zmin;
    b.max_depth = zmax;
  }
#else
  for (int i = 0; i < n; ++i) order_box_depth_range(m, boxes[i]);
#endif
}

namespace {

// Common/VisibilityOrdering.cpp:211-228
inline bool touching(float a, float b) {
  const float scale = fmax2(fmax2(1.0f, std::fabs(a)), std::fabs(b));
  return std::fabs(a - b) <= 1e-5f * scale;
}
inline bool spans_overlap(float a0, float a1, float b0, float b1) {
  const float lo = fmax2(a0, b0), hi = fmin2(a1, b1);
  float scale = 1.0f;
  for (float v : {a0, a1, b0, b1, lo, hi}) scale = fmax2(scale, std::fabs(v));
  return (hi - lo) > 1e-5f * scale;
}

// total order of Common/VisibilityOrdering.cpp:238-257
inline bool box_precedes(const std::vector<OrderBox>& boxes, int a, int b) {
  const OrderBox& x = boxes[a];
  const OrderBox& y = boxes[b];
  const bool xf = std::isfinite(x.min_depth), yf = std::isfinite(y.min_depth);
  if (xf != yf) return xf;
  if (x.min_depth != y.min_depth) return x.min_depth < y.min_depth;
  if (x.max_depth != y.max_depth) return x.max_depth < y.max_depth;
  if (x.owner != y.owner) return x.owner < y.owner;
  return a < b;
}

struct HostOrderBackend final : OrderBackend {
  void rank_boxes(const std::vector<OrderBox>& boxes, std::vector<int32_t>& priority) override {
    const int n = static_cast<int>(boxes.size());
    std::vector<int> idx(n);
    for (int i = 0; i < n; ++i) idx[i] = i;
    std::sort(idx.begin(), idx.end(), [&](int a, int b) { return box_precedes(boxes, a, b); });
    priority.assign(n, 0);
    for (int p = 0; p < n; ++p) priority[idx[p]] = p;
  }
  void pair_edges(const std::vector<OrderBox>& boxes, const float* dir,
                  std::vector<OrderEdge>& edges) override {
    order_pair_edges_host(boxes, dir, edges);
  }
};

}  // namespace

OrderBackend* host_order_backend() {
  static HostOrderBackend backend;
  return &backend;
}

namespace {

// Exact test of one pair (i < j), all three axes: VisibilityOrdering.cpp:261-317.
inline void pair_edges_exact(const std::vector<OrderBox>& boxes, const float* dir, int i, int j,
                             unsigned axes, std::vector<OrderEdge>& edges) {
  const OrderBox& a = boxes[i];
  const OrderBox& b = boxes[j];
  for (int axis = 0; axis < 3; ++axis) {
    if (((axes >> axis) & 1u) == 0u) continue;  // the filter proved this axis carries no edge
    const int u = (axis + 1) % 3, v = (axis + 2) % 3;
    const float along = dir[axis];
    // a below b on this axis: looking up the axis (along > 0) puts a in front.
    int forward = -1;
    if (touching(a.hi[axis], b.lo[axis])) {
      if (along > 1e-6f) forward = 0;        // reference adds edge j -> i
      else if (along < -1e-6f) forward = 1;  // edge i -> j
    } else if (touching(b.hi[axis], a.lo[axis])) {
      if (along > 1e-6f) forward = 1;
      else if (along < -1e-6f) forward = 0;
    }
    if (forward < 0) continue;
    if (!spans_overlap(a.lo[u], a.hi[u], b.lo[u], b.hi[u])) continue;
    if (!spans_overlap(a.lo[v], a.hi[v], b.lo[v], b.hi[v])) continue;
    edges.push_back(OrderEdge{i, j, axis, forward});
  }
}

// Conservative filter for "pair (i, j) may carry an edge", eight (AVX2), four
// (SSE2) or one pair at a time.  An edge on axis A needs the faces to touch on A
// (|a.hi - b.lo| or |b.hi - a.lo| <= 1e-5 * scale, and scale <= `largest`, so
// <= tol) and the spans on the other two axes to overlap by MORE than
// 1e-5 * scale > 0; the filter keeps every pair for which that can hold and the
// exact test above decides.  soa = lo0 lo1 lo2 hi0 hi1 hi2, `stride` floats each
// (n boxes, then 8 columns of padding placed far away from everything).
// cand[k] = j | (axes << 28): bit a of axes set when axis a may carry the edge (n < 2^28).
typedef void (*PairFilterFn)(const float* soa, int stride, int n, int i, float tol, int* cand, int* ncand);

inline unsigned pair_may_touch(const float* soa, int stride, int i, int j, float tol) {
  bool ov[3], touch[3];
  for (int d = 0; d < 3; ++d) {
    const float alo = soa[d * stride + i], ahi = soa[(3 + d) * stride + i];
    const float blo = soa[d * stride + j], bhi = soa[(3 + d) * stride + j];
    ov[d] = (fmin2(ahi, bhi) - fmax2(alo, blo)) > 0.0f;
    touch[d] = std::fabs(ahi - blo) <= tol || std::fabs(bhi - alo) <= tol;
  }
  return (touch[0] && ov[1] && ov[2] ? 1u : 0u) | (touch[1] && ov[2] && ov[0] ? 2u : 0u) |
         (touch[2] && ov[0] && ov[1] ? 4u : 0u);
}

[[maybe_unused]] void pair_filter_scalar(const float* soa, int stride, int n, int i, float tol,
                                         int* cand, int* ncand) {
  int k = 0;
  for (int j = i + 1; j < n; ++j) {
    const unsigned axes = pair_may_touch(soa, stride, i, j, tol);
    if (axes != 0u) cand[k++] = j | static_cast<int>(axes << 28);
  }
  *ncand = k;
}

#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("avx2"))) void pair_filter_avx2(const float* soa, int stride, int n, int i,
                                                      float tol, int* cand, int* ncand) {
  const __m256 absmask = _mm256_castsi256_ps(_mm256_set1_epi32(0x7fffffff));
  const __m256 vtol = _mm256_set1_ps(tol), zero = _mm256_setzero_ps();
  __m256 alo[3], ahi[3];
  for (int d = 0; d < 3; ++d) {
    alo[d] = _mm256_set1_ps(soa[d * stride + i]);
    ahi[d] = _mm256_set1_ps(soa[(3 + d) * stride + i]);
  }
  int k = 0, j = i + 1;
  for (; j < n; j += 8) {  // columns n .. n+7 are padding that never matches
    __m256 ov[3], touch[3];
    for (int d = 0; d < 3; ++d) {
      const __m256 blo = _mm256_loadu_ps(soa + d * stride + j), bhi = _mm256_loadu_ps(soa + (3 + d) * stride + j);
      ov[d] = _mm256_cmp_ps(_mm256_sub_ps(_mm256_min_ps(ahi[d], bhi), _mm256_max_ps(alo[d], blo)), zero,
                            _CMP_GT_OQ);
      const __m256 g0 = _mm256_and_ps(_mm256_sub_ps(ahi[d], blo), absmask);
      const __m256 g1 = _mm256_and_ps(_mm256_sub_ps(bhi, alo[d]), absmask);
      touch[d] = _mm256_cmp_ps(_mm256_min_ps(g0, g1), vtol, _CMP_LE_OQ);
    }
    const unsigned m0 = static_cast<unsigned>(
        _mm256_movemask_ps(_mm256_and_ps(touch[0], _mm256_and_ps(ov[1], ov[2]))));
    const unsigned m1 = static_cast<unsigned>(
        _mm256_movemask_ps(_mm256_and_ps(touch[1], _mm256_and_ps(ov[2], ov[0]))));
    const unsigned m2 = static_cast<unsigned>(
        _mm256_movemask_ps(_mm256_and_ps(touch[2], _mm256_and_ps(ov[0], ov[1]))));
    unsigned bits = m0 | m1 | m2;
    while (bits != 0u) {
      const int b = __builtin_ctz(bits);
      const unsigned axes = ((m0 >> b) & 1u) | (((m1 >> b) & 1u) << 1) | (((m2 >> b) & 1u) << 2);
      cand[k++] = (j + b) | static_cast<int>(axes << 28);
      bits &= bits - 1u;
    }
  }
  *ncand = k;
}

void pair_filter_sse2(const float* soa, int stride, int n, int i, float tol, int* cand, int* ncand) {
  const __m128 absmask = _mm_castsi128_ps(_mm_set1_epi32(0x7fffffff));
  const __m128 vtol = _mm_set1_ps(tol), zero = _mm_setzero_ps();
  __m128 alo[3], ahi[3];
  for (int d = 0; d < 3; ++d) {
    alo[d] = _mm_set1_ps(soa[d * stride + i]);
    ahi[d] = _mm_set1_ps(soa[(3 + d) * stride + i]);
  }
  int k = 0, j = i + 1;
  for (; j < n; j += 4) {  // columns n .. n+7 are padding that never matches
    __m128 ov[3], touch[3];
    for (int d = 0; d < 3; ++d) {
      const __m128 blo = _mm_loadu_ps(soa + d * stride + j), bhi = _mm_loadu_ps(soa + (3 + d) * stride + j);
      ov[d] = _mm_cmpgt_ps(_mm_sub_ps(_mm_min_ps(ahi[d], bhi), _mm_max_ps(alo[d], blo)), zero);
      const __m128 g0 = _mm_and_ps(_mm_sub_ps(ahi[d], blo), absmask);
      const __m128 g1 = _mm_and_ps(_mm_sub_ps(bhi, alo[d]), absmask);
      touch[d] = _mm_cmple_ps(_mm_min_ps(g0, g1), vtol);
    }
    const unsigned m0 =
        static_cast<unsigned>(_mm_movemask_ps(_mm_and_ps(touch[0], _mm_and_ps(ov[1], ov[2]))));
    const unsigned m1 =
        static_cast<unsigned>(_mm_movemask_ps(_mm_and_ps(touch[1], _mm_and_ps(ov[2], ov[0]))));
    const unsigned m2 =
        static_cast<unsigned>(_mm_movemask_ps(_mm_and_ps(touch[2], _mm_and_ps(ov[0], ov[1]))));
    unsigned bits = m0 | m1 | m2;
    while (bits != 0u) {
      const int b = __builtin_ctz(bits);
      const unsigned axes = ((m0 >> b) & 1u) | (((m1 >> b) & 1u) << 1) | (((m2 >> b) & 1u) << 2);
      cand[k++] = (j + b) | static_cast<int>(axes << 28);
      bits &= bits - 1u;
    }
  }
  *ncand = k;
}

PairFilterFn pick_pair_filter() {
  __builtin_cpu_init();
  return __builtin_cpu_supports("avx2") ? pair_filter_avx2 : pair_filter_sse2;
}
#else
PairFilterFn pick_pair_filter() { return pair_filter_scalar; }
#endif

}  // namespace

void order_pair_edges_host(const std::vector<OrderBox>& boxes, const float* dir,
                           std::vector<OrderEdge>& edges) {
  edges.clear();
  const int n = static_cast<int>(boxes.size());
  static const PairFilterFn filter = pick_pair_filter();
  // The filter is valid when every coordinate is finite and lo <= hi (NaNs and
  // inverted boxes make min/max/compare behave differently from the scalar
  // reference test); anything else takes the exact test on every pair.
  float largest = 1.0f;
  bool well_formed = true;
  for (const OrderBox& b : boxes)
    for (int d = 0; d < 3; ++d) {
      largest = fmax2(largest, fmax2(std::fabs(b.lo[d]), std::fabs(b.hi[d])));
      well_formed = well_formed && (b.lo[d] <= b.hi[d]) && std::isfinite(b.lo[d]) && std::isfinite(b.hi[d]);
    }
  if (!well_formed || !std::isfinite(largest) || n >= (1 << 28)) {
    for (int i = 0; i < n; ++i)
      for (int j = i + 1; j < n; ++j) pair_edges_exact(boxes, dir, i, j, 7u, edges);
    return;
  }
  const float tol = 1.001e-5f * largest;
  thread_local std::vector<float> soa;
  thread_local std::vector<int> cand;
  const int stride = n + 8;
  soa.assign(static_cast<size_t>(stride) * 6, 3.0e38f);  // padding: touches and overlaps nothing
  cand.resize(static_cast<size_t>(n) + 8);
  for (int i = 0; i < n; ++i)
    for (int d = 0; d < 3; ++d) {
      soa[static_cast<size_t>(d) * stride + i] = boxes[i].lo[d];
      soa[static_cast<size_t>(3 + d) * stride + i] = boxes[i].hi[d];
    }
  for (int i = 0; i < n; ++i) {
    int ncand = 0;
    filter(soa.data(), stride, n, i, tol, cand.data(), &ncand);
    for (int c = 0; c < ncand; ++c)
      pair_edges_exact(boxes, dir, i, cand[c] & 0x0fffffff, static_cast<unsigned>(cand[c]) >> 28, edges);
  }
}

namespace {

// Adjacency in compressed rows: the targets of node v are adj[start[v] ..
// start[v + 1]) in the reference's insertion order.
struct Graph {
  int n = 0;
  std::vector<int> start, adj, indegree;
  int degree(int v) const { return start[v + 1] - start[v]; }
  int target(int v, int k) const { return adj[start[v] + k]; }
};

// adjacency lists in the reference's insertion order (:261-317): edges arrive
// sorted by (i, j, axis); duplicates (same from->to) are dropped.  Two edges
// with the same endpoints always belong to the same pair (i, j), so after the
// sort a duplicate is one of the at most two edges just before it.
void build_graph(int n, std::vector<OrderEdge>& edges, Graph& g) {
  auto by_pair = [](const OrderEdge& a, const OrderEdge& b) {
    if (a.i != b.i) return a.i < b.i;
    if (a.j != b.j) return a.j < b.j;
    return a.axis < b.axis;
  };
  if (!std::is_sorted(edges.begin(), edges.end(), by_pair)) {
    // the host pair test emits edges in order; the device kernel appends them as
    // its threads find them.  Bucket by i (counting sort), then order each
    // bucket -- a box has a handful of neighbours -- by (j, axis) with an
    // insertion sort.
    thread_local std::vector<int> head, at;
    thread_local std::vector<OrderEdge> sorted;
    bool in_range = true;
    for (const OrderEdge& e : edges) in_range = in_range && e.i >= 0 && e.i < n;
    if (in_range) {
      head.assign(static_cast<size_t>(n) + 1, 0);
      for (const OrderEdge& e : edges) ++head[static_cast<size_t>(e.i) + 1];
      for (int v = 0; v < n; ++v) head[v + 1] += head[v];
      at.assign(head.begin(), head.end() - 1);
      sorted.resize(edges.size());
      for (const OrderEdge& e : edges) {
        // insert into the filled part of bucket e.i, keeping it ordered
        const int first = head[e.i];
        int pos = at[e.i]++;
        while (pos > first && by_pair(e, sorted[pos - 1])) {
          sorted[pos] = sorted[pos - 1];
          --pos;
        }
        sorted[pos] = e;
      }
      edges.swap(sorted);
    } else {
      std::sort(edges.begin(), edges.end(), by_pair);
    }
  }
  const size_t m = edges.size();
  thread_local std::vector<unsigned char> keep;
  keep.assign(m, 1);
  g.n = n;
  g.start.assign(static_cast<size_t>(n) + 1, 0);
  g.indegree.assign(static_cast<size_t>(n), 0);
  for (size_t e = 0; e < m; ++e) {
    const OrderEdge& E = edges[e];
    for (size_t back = 1; back <= 2 && back <= e; ++back) {
      const OrderEdge& F = edges[e - back];
      if (F.i != E.i || F.j != E.j) break;
      if (keep[e - back] && F.forward == E.forward) keep[e] = 0;
    }
    if (!keep[e]) continue;
    const int from = E.forward ? E.i : E.j, to = E.forward ? E.j : E.i;
    ++g.start[static_cast<size_t>(from) + 1];
    ++g.indegree[to];
  }
  for (int v = 0; v < n; ++v) g.start[v + 1] += g.start[v];
  g.adj.resize(static_cast<size_t>(g.start[n]));
  thread_local std::vector<int> fill;
  fill.assign(g.start.begin(), g.start.end() - 1);
  for (size_t e = 0; e < m; ++e) {
    if (!keep[e]) continue;
    const OrderEdge& E = edges[e];
    const int from = E.forward ? E.i : E.j, to = E.forward ? E.j : E.i;
    g.adj[fill[from]++] = to;
  }
}

// Kahn's algorithm taking, at each step, the ready box that comes first in the
// compareBoxes order (:360-401 re-sorts the ready list and pops the front).
// priority is a permutation of 0..n-1, so the ready set is a bitmap indexed by
// priority and "the front" is its lowest set bit.
bool topo_order(const Graph& g, const std::vector<int32_t>& priority, std::vector<int>& order,
                std::vector<int>& residual) {
  const int n = g.n;
  residual = g.indegree;
  const int words = (n + 63) / 64;
  thread_local std::vector<uint64_t> ready;
  thread_local std::vector<int> node_at;
  ready.assign(static_cast<size_t>(words), 0);
  node_at.resize(static_cast<size_t>(n));
  for (int i = 0; i < n; ++i) node_at[priority[i]] = i;
  for (int i = 0; i < n; ++i)
    if (residual[i] == 0) ready[priority[i] >> 6] |= 1ull << (priority[i] & 63);
  order.clear();
  order.reserve(n);
  int low = 0;  // no bit is set below word `low`... unless a push lowers it again
  for (;;) {
    while (low < words && ready[low] == 0) ++low;
    if (low == words) break;
    const int bit = __builtin_ctzll(ready[low]);
    ready[low] &= ready[low] - 1;
    const int node = node_at[low * 64 + bit];
    order.push_back(node);
    for (int k = g.start[node]; k < g.start[node + 1]; ++k) {
      const int next = g.adj[k];
      if (--residual[next] == 0) {
        const int p = priority[next];
        ready[p >> 6] |= 1ull << (p & 63);
        if ((p >> 6) < low) low = p >> 6;
      }
    }
  }
  return static_cast<int>(order.size()) == n;
}

// First cycle a depth-first walk meets, nodes in walk order (:403-443).
// Iterative (the reference recurses); visiting order is identical.
std::vector<int> first_cycle(const Graph& g, const std::vector<int>& residual) {
  const int n = g.n;
  std::vector<int> state(n, 0), parent(n, -1);
  std::vector<size_t> cursor(n, 0);
  std::vector<int> stack;
  for (int root = 0; root < n; ++root) {
    if (!(residual[root] > 0) || state[root] != 0) continue;
    stack.assign(1, root);
    state[root] = 1;
    while (!stack.empty()) {
      const int node = stack.back();
      if (cursor[node] < static_cast<size_t>(g.degree(node))) {
        const int next = g.target(node, static_cast<int>(cursor[node]++));
        if (state[next] == 0) {
          parent[next] = node;
          state[next] = 1;
          stack.push_back(next);
        } else if (state[next] == 1) {
          std::vector<int> cycle{next};
          for (int cur = node; cur != next && cur != -1; cur = parent[cur]) cycle.push_back(cur);
          std::reverse(cycle.begin(), cycle.end());
          return cycle;
        }
      } else {
        state[node] = 2;
        stack.pop_back();
      }
    }
  }
  return {};
}

// Cut the longest member of the cycle along the dominant view axis at another
// member's face; the near half keeps the slot, the far half is appended
// (:445-563).
bool split_cycle_member(const std::vector<int>& cycle, std::vector<OrderBox>& boxes,
                        const OrderMatrices& m) {
  if (cycle.size() < 2) return false;
  const float* dir = m.dir;
  int axis = 0;
  float strongest = std::fabs(dir[0]);
  for (int a = 1; a < 3; ++a)
    if (std::fabs(dir[a]) > strongest) {
      strongest = std::fabs(dir[a]);
      axis = a;
    }
  if (strongest <= 1e-6f) {
    float widest = -1.0f;
    for (int a = 0; a < 3; ++a)
      for (int id : cycle) {
        const float len = boxes[id].hi[a] - boxes[id].lo[a];
        if (len > widest) {
          widest = len;
          axis = a;
        }
      }
  }
  const float along = dir[axis];
  if (std::fabs(along) <= 1e-6f) return false;
  int target = cycle.front();
  float target_len = -1.0f;
  for (int id : cycle) {
    const float len = boxes[id].hi[axis] - boxes[id].lo[axis];
    if (len > target_len && len > 1e-6f) {
      target_len = len;
      target = id;
    }
  }
  if (target_len <= 1e-6f) return false;
  const OrderBox whole = boxes[target];
  const float lo = whole.lo[axis], hi = whole.hi[axis];
  const float margin = fmax2(1e-5f * (hi - lo), 1e-6f);
  bool any = false;
  float cut_hi = 0.0f, cut_lo = 0.0f;
  for (int id : cycle) {
    if (id == target) continue;
    for (float face : {boxes[id].lo[axis], boxes[id].hi[axis]}) {
      if (!(face > lo + margin && face < hi - margin)) continue;
      if (!any) {
        cut_hi = cut_lo = face;
        any = true;
      } else {
        if (cut_hi < face) cut_hi = face;
        if (face < cut_lo) cut_lo = face;
      }
    }
  }
  float cut = 0.5f * (lo + hi);
  if (any) cut = (along > 0.0f) ? cut_hi : cut_lo;
  if (cut <= lo + margin) cut = lo + margin;
  if (cut >= hi - margin) cut = hi - margin;
  if (!(cut > lo && cut < hi)) return false;
  OrderBox near_half = whole, far_half = whole;
  if (along > 0.0f) {
    near_half.hi[axis] = cut;
    far_half.lo[axis] = cut;
  } else {
    near_half.lo[axis] = cut;
    far_half.hi[axis] = cut;
  }
  order_box_depth_range(m, near_half);
  order_box_depth_range(m, far_half);
  boxes[target] = near_half;
  boxes.push_back(far_half);
  return true;
}

// exportGraph (:319-352): nodes labelled with owner and NDC depth interval
// (std::fixed, six decimals), then one line per edge, rows in node order.
std::string graph_dot(const std::vector<OrderBox>& boxes, const Graph& g) {
  std::string out = "digraph VisibilityGraph {\n  rankdir=LR;\n";
  char line[256];
  for (size_t i = 0; i < boxes.size(); ++i) {
    std::snprintf(line, sizeof(line),
                  "  box%zu [label=\"box %zu\\nrank %d\\nminDepth %.6f\\nmaxDepth %.6f\"];\n", i, i,
                  boxes[i].owner, static_cast<double>(boxes[i].min_depth),
                  static_cast<double>(boxes[i].max_depth));
    out += line;
  }
  for (int from = 0; from < g.n; ++from)
    for (int k = 0; k < g.degree(from); ++k) {
      std::snprintf(line, sizeof(line), "  box%d -> box%d;\n", from, g.target(from, k));
      out += line;
    }
  out += "}\n";
  return out;
}

}  // namespace

bool order_ranks(int nranks, int nglobal, const amrvr_box_record* recs,
                 const amrvr_camera& cam, float aspect, bool use_graph,
                 OrderBackend& backend, int* rank_order, std::vector<std::string>* graph_dots) {
  for (int r = 0; r < nranks; ++r) rank_order[r] = r;
  if (!use_graph || nglobal <= 0) return true;

  OrderMatrices m;
  make_order_matrices(cam, aspect, m);
  std::vector<OrderBox> boxes(static_cast<size_t>(nglobal));
  for (int i = 0; i < nglobal; ++i) {
    for (int d = 0; d < 3; ++d) {
      boxes[i].lo[d] = static_cast<float>(recs[i].min_corner[d]);
      boxes[i].hi[d] = static_cast<float>(recs[i].max_corner[d]);
    }
    boxes[i].owner = recs[i].owner;
  }
  order_box_depth_ranges(m, boxes.data(), nglobal);

  const int max_rounds = std::max(nglobal, 1) * 8 + 32;  // :565-566
  std::vector<OrderEdge> edges;
  std::vector<int32_t> priority;
  std::vector<int> order, residual;
  Graph g;
  for (int round = 0; round < max_rounds; ++round) {
    const int n = static_cast<int>(boxes.size());
    backend.pair_edges(boxes, m.dir, edges);
    build_graph(n, edges, g);
    if (graph_dots != nullptr) graph_dots->push_back(graph_dot(boxes, g));
    backend.rank_boxes(boxes, priority);
    if (topo_order(g, priority, order, residual)) {
      std::vector<char> seen(static_cast<size_t>(nranks), 0);
      int filled = 0;
      for (int id : order) {  // :576-593 owners in first-appearance order
        const int owner = boxes[id].owner;
        if (owner >= 0 && owner < nranks && !seen[owner]) {
          seen[owner] = 1;
          rank_order[filled++] = owner;
        }
      }
      for (int r = 0; r < nranks; ++r)
        if (!seen[r]) rank_order[filled++] = r;
      return true;
    }
    const std::vector<int> cycle = first_cycle(g, residual);
    if (cycle.empty()) break;
    if (!split_cycle_member(cycle, boxes, m)) break;
  }
  for (int r = 0; r < nranks; ++r) rank_order[r] = r;  // :615-623
  return false;
}

// ---------------------------------------------------------------------------
// Wireframe overlay geometry (VolumeRenderer.cpp:139-335, host part)
// ---------------------------------------------------------------------------
void make_overlay_edges(const double* tight_min, const double* tight_max,
                        const amrvr_camera& cam, int sqrt_aa, int width, int height,
                        OverlayEdges& out) {
  std::memset(&out, 0, sizeof(out));
  if (width <= 0 || height <= 0) return;
  const Frame3 fr = camera_frame(cam);
  const float aspect = static_cast<float>(width) / static_cast<float>(std::max(height, 1));
  const float tan_half = std::tan(cam.fov_y_degrees * 0.5f * kPiF / 180.0f);
  const float sx_scale = (width > 1) ? static_cast<float>(width - 1) : 0.0f;
  const float sy_scale = (height > 1) ? static_cast<float>(height - 1) : 0.0f;
  float sx[8], sy[8];
  bool ok[8];
  for (int c = 0; c < 8; ++c) {
    const double p[3] = {(c & 1) ? tight_max[0] : tight_min[0],
                         (c & 2) ? tight_max[1] : tight_min[1],
                         (c & 4) ? tight_max[2] : tight_min[2]};
    const V3 rel = sub(p, cam.eye);
    ok[c] = false;
    sx[c] = sy[c] = 0.0f;
    const float depth = static_cast<float>(dot(rel, fr.f));
    if (!(depth > 0.0f) || !std::isfinite(depth)) continue;
    const float xc = static_cast<float>(dot(rel, fr.r));
    const float yc = static_cast<float>(dot(rel, fr.u));
    const float nx = xc / (depth * tan_half * aspect);
    const float ny = yc / (depth * tan_half);
    if (!std::isfinite(nx) || !std::isfinite(ny)) continue;
    sx[c] = (nx * 0.5f + 0.5f) * sx_scale;
    sy[c] = (ny * 0.5f + 0.5f) * sy_scale;
    ok[c] = true;
  }
  static const int kEdge[12][2] = {{0, 1}, {1, 3}, {3, 2}, {2, 0}, {4, 5}, {5, 7},
                                   {7, 6}, {6, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};
  const float reach = 0.5f * static_cast<float>(std::max(sqrt_aa, 1)) + 0.5f;
  for (int e = 0; e < 12; ++e) {
    const int a = kEdge[e][0], b = kEdge[e][1];
    if (!ok[a] || !ok[b]) continue;
    const int k = out.count++;
    out.ax[k] = sx[a];
    out.ay[k] = sy[a];
    out.bx[k] = sx[b];
    out.by[k] = sy[b];
    const float ex = sx[b] - sx[a], ey = sy[b] - sy[a];
    out.degenerate[k] = !((ex * ex + ey * ey) > 0.0f);
    if (out.degenerate[k]) {
      out.x0[k] = out.x1[k] = static_cast<int32_t>(std::lround(sx[a]));
      out.y0[k] = out.y1[k] = static_cast<int32_t>(std::lround(sy[a]));
      continue;
    }
    out.x0[k] = std::max(0, static_cast<int>(std::floor(fmin2(sx[a], sx[b]) - reach)));
    out.x1[k] = std::min(width - 1, static_cast<int>(std::ceil(fmax2(sx[a], sx[b]) + reach)));
    out.y0[k] = std::max(0, static_cast<int>(std::floor(fmin2(sy[a], sy[b]) - reach)));
    out.y1[k] = std::min(height - 1, static_cast<int>(std::ceil(fmax2(sy[a], sy[b]) + reach)));
  }
}

}  // namespace amrvr

// ------------------------------------------------- visibility graph as DOT text
extern "C" int amrvr_order_graph_dot_host(int nranks, int nglobal, const amrvr_box_record* records,
                                          const amrvr_camera* camera, float aspect, int round,
                                          char* text, int64_t capacity, int64_t* needed,
                                          int* nrounds) {
  if (nranks < 1 || nglobal < 0 || (nglobal > 0 && records == nullptr) || camera == nullptr ||
      round < 0 || capacity < 0 || (needed == nullptr && nrounds == nullptr && text == nullptr))
    return amrvr::report_error(AMRVR_ERR_INVALID, "bad argument");
  std::vector<int> rank_order(static_cast<size_t>(nranks));
  std::vector<std::string> dots;
  amrvr::order_ranks(nranks, nglobal, records, *camera, aspect, true, *amrvr::host_order_backend(),
                     rank_order.data(), &dots);
  if (nrounds != nullptr) *nrounds = static_cast<int>(dots.size());
  if (round >= static_cast<int>(dots.size())) {
    if (needed != nullptr) *needed = 0;
    if (text == nullptr) return AMRVR_OK;  // size query of a round that does not exist
    return amrvr::report_error(AMRVR_ERR_INVALID, "round out of range");
  }
  const std::string& doc = dots[static_cast<size_t>(round)];
  if (needed != nullptr) *needed = static_cast<int64_t>(doc.size());
  if (text == nullptr) return AMRVR_OK;
  if (capacity < static_cast<int64_t>(doc.size()))
    return amrvr::report_error(AMRVR_ERR_INVALID, "text is too small");
  std::memcpy(text, doc.data(), doc.size());
  return AMRVR_OK;
}

// ------------------------------------------------------------ automatic camera
// reference: VolumeRenderer/VolumeRenderer.cpp:975-1030 (renderScene without an
// explicit camera).  The two angles come from std::mt19937(seed) through
// std::uniform_real_distribution<float> -- the same libstdc++ templates the
// reference instantiates -- azimuth first, then altitude.
#include <random>

namespace {

struct D3 {
  double v[3];
};

double length3(const D3& a) { return std::sqrt(a.v[0] * a.v[0] + a.v[1] * a.v[1] + a.v[2] * a.v[2]); }

D3 cross3(const D3& a, const D3& b) {
  return D3{{a.v[1] * b.v[2] - a.v[2] * b.v[1], a.v[2] * b.v[0] - a.v[0] * b.v[2],
             a.v[0] * b.v[1] - a.v[1] * b.v[0]}};
}

// safeNormalize of the reference: (0, 0, -1) when the length is unusable
D3 unit_or_minus_z(const D3& a) {
  const double len = length3(a);
  if (!(len > 0.0) || !std::isfinite(len)) return D3{{0.0, 0.0, -1.0}};
  return D3{{a.v[0] / len, a.v[1] / len, a.v[2] / len}};
}

}  // namespace

extern "C" int amrvr_auto_camera(const double* bounds_min, const double* bounds_max, unsigned seed,
                                 const double* custom_up, amrvr_camera* out) {
  if (bounds_min == nullptr || bounds_max == nullptr || out == nullptr)
    return amrvr::report_error(AMRVR_ERR_INVALID, "bad argument");
  constexpr float pi = 3.14159265358979323846f;
  D3 center, half_extent;
  for (int d = 0; d < 3; ++d) {
    center.v[d] = 0.5f * (bounds_min[d] + bounds_max[d]);
    half_extent.v[d] = 0.5f * (bounds_max[d] - bounds_min[d]);
  }
  float radius = static_cast<float>(length3(half_extent));
  if (radius <= 0.0f) radius = 1.0f;

  const float fov = pi * 0.25f, max_altitude = pi * 0.25f, half_fov = fov * 0.5f;
  const float fit = (half_fov > 0.0f) ? radius / static_cast<float>(std::tan(half_fov)) : radius;
  const float distance = fit + std::max(0.25f * radius, 0.5f);

  std::mt19937 rng(seed);
  std::uniform_real_distribution<float> pick_azimuth(0.0f, 2.0f * pi);
  std::uniform_real_distribution<float> pick_altitude(-max_altitude, max_altitude);
  const float azimuth = pick_azimuth(rng);
  const float altitude = pick_altitude(rng);
  const float cos_alt = std::cos(altitude);

  const D3 eye{{center.v[0] + distance * cos_alt * std::sin(azimuth),
                center.v[1] + distance * std::sin(altitude),
                center.v[2] + distance * cos_alt * std::cos(azimuth)}};
  const D3 view = unit_or_minus_z(D3{{center.v[0] - eye.v[0], center.v[1] - eye.v[1], center.v[2] - eye.v[2]}});
  D3 up = (custom_up != nullptr) ? D3{{custom_up[0], custom_up[1], custom_up[2]}} : D3{{0.0, 1.0, 0.0}};
  const D3 spare[2] = {D3{{0.0, 0.0, 1.0}}, D3{{1.0, 0.0, 0.0}}};
  for (int k = 0; k < 2 && length3(cross3(view, up)) <= 1e-4f; ++k) up = spare[k];
  up = unit_or_minus_z(up);

  for (int d = 0; d < 3; ++d) {
    out->eye[d] = eye.v[d];
    out->look_at[d] = center.v[d];
    out->up[d] = up.v[d];
  }
  out->fov_y_degrees = fov * 180.0f / pi;
  out->near_plane = 0.1f;
  out->far_plane = distance * 4.0f;
  return AMRVR_OK;
}
