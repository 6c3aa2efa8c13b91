/* oracle/restate/orc_color.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates the transfer-function table of Common/VolumePainter.cpp:107-516. */
#include "oracle.h"

#include <math.h>
#include <string.h>

#define ORC_MAX_NODES 1024

typedef struct {
  float value, r, g, b;
} orc_cnode; /* VolumePainter.cpp:46-51 */

typedef struct {
  float value, alpha, midpoint, sharpness;
} orc_onode; /* VolumePainter.cpp:53-58 */

typedef struct {
  int lab; /* ColorSpace::kLab, :60-63 */
  int ncolors, nopacity;
  orc_cnode colors[ORC_MAX_NODES];
  orc_onode opacity[ORC_MAX_NODES];
} orc_table;

static float clampf(float v, float lo, float hi) { /* std::clamp */
  return (v < lo) ? lo : ((hi < v) ? hi : v);
}
static float maxf(float a, float b) { return (a < b) ? b : a; } /* std::max */
static float minf(float a, float b) { return (b < a) ? b : a; } /* std::min */

/* VolumePainter.cpp:107-125 computeScaledAlpha */
static float scaled_alpha(float base, float alpha_scale, float norm) {
  const float scaled = clampf(base * alpha_scale, 0.0f, 1.0f);
  if (norm <= 0.0f || scaled <= 0.0f) return 0.0f;
  if (scaled >= 1.0f) return 1.0f;
  const double transmittance = pow(1.0 - (double)scaled, (double)norm);
  float a = (float)(1.0 - transmittance);
  if (!isfinite(a)) a = scaled;
  return clampf(a, 0.0f, 1.0f);
}

/* VolumePainter.cpp:127-138 insertColorNode: sorted insert, equal value replaces */
static void insert_color(orc_table* t, orc_cnode n) {
  int pos = 0;
  while (pos < t->ncolors && t->colors[pos].value < n.value) ++pos;
  if (pos < t->ncolors && t->colors[pos].value == n.value) {
    t->colors[pos] = n;
    return;
  }
  if (t->ncolors >= ORC_MAX_NODES) return;
  memmove(&t->colors[pos + 1], &t->colors[pos],
          sizeof(orc_cnode) * (size_t)(t->ncolors - pos));
  t->colors[pos] = n;
  ++t->ncolors;
}

/* VolumePainter.cpp:140-151 insertOpacityNode */
static void insert_opacity(orc_table* t, orc_onode n) {
  int pos = 0;
  while (pos < t->nopacity && t->opacity[pos].value < n.value) ++pos;
  if (pos < t->nopacity && t->opacity[pos].value == n.value) {
    t->opacity[pos] = n;
    return;
  }
  if (t->nopacity >= ORC_MAX_NODES) return;
  memmove(&t->opacity[pos + 1], &t->opacity[pos],
          sizeof(orc_onode) * (size_t)(t->nopacity - pos));
  t->opacity[pos] = n;
  ++t->nopacity;
}

/* VolumePainter.cpp:153-200 getTableRange + rescaleTableToRange */
static void rescale_table(orc_table* t, float range_min, float range_max) {
  int has = 0;
  float lo = 0.0f, hi = 0.0f;
  for (int i = 0; i < t->ncolors; ++i) {
    const float v = t->colors[i].value;
    if (!has) {
      lo = hi = v;
      has = 1;
    } else {
      lo = minf(lo, v);
      hi = maxf(hi, v);
    }
  }
  for (int i = 0; i < t->nopacity; ++i) {
    const float v = t->opacity[i].value;
    if (!has) {
      lo = hi = v;
      has = 1;
    } else {
      lo = minf(lo, v);
      hi = maxf(hi, v);
    }
  }
  if (!has) lo = hi = 0.0f;
  const float old_span = hi - lo;
  const float new_span = range_max - range_min;
  if (!(old_span > 0.0f) || !(new_span > 0.0f)) return;
  for (int i = 0; i < t->ncolors; ++i) {
    const float u = (t->colors[i].value - lo) / old_span;
    t->colors[i].value = range_min + new_span * u;
  }
  for (int i = 0; i < t->nopacity; ++i) {
    const float u = (t->opacity[i].value - lo) / old_span;
    t->opacity[i].value = range_min + new_span * u;
  }
}

static float srgb_to_linear(float c) { /* :206-220 */
  return (c > 0.04045f) ? powf((c + 0.055f) / 1.055f, 2.4f) : c / 12.92f;
}
static float lab_f(float v) { /* :234-248 */
  const float one_third = 1.0f / 3.0f;
  const float k16 = 16.0f / 116.0f;
  return (v > 0.008856f) ? powf(v, one_third) : (7.787f * v) + k16;
}
static float lab_finv(float v) { /* :263-277 */
  const float k16 = 16.0f / 116.0f;
  return (powf(v, 3.0f) > 0.008856f) ? powf(v, 3.0f) : (v - k16) / 7.787f;
}
static float linear_to_srgb(float c) { /* :291-305 */
  const float inv = 1.0f / 2.4f;
  return (c > 0.0031308f) ? 1.055f * powf(c, inv) - 0.055f : 12.92f * c;
}

/* VolumePainter.cpp:202-256 rgbToLab (alpha rides along in [3]) */
static void rgb_to_lab(const float* rgb, float* lab) {
  const float r = srgb_to_linear(rgb[0]);
  const float g = srgb_to_linear(rgb[1]);
  const float b = srgb_to_linear(rgb[2]);
  const float x = r * 0.4124f + g * 0.3576f + b * 0.1805f;
  const float y = r * 0.2126f + g * 0.7152f + b * 0.0722f;
  const float z = r * 0.0193f + g * 0.1192f + b * 0.9505f;
  const float vx = lab_f(x / 0.9505f);
  const float vy = lab_f(y / 1.0f);
  const float vz = lab_f(z / 1.089f);
  lab[0] = (116.0f * vy) - 16.0f;
  lab[1] = 500.0f * (vx - vy);
  lab[2] = 200.0f * (vy - vz);
  lab[3] = rgb[3];
}

/* VolumePainter.cpp:258-320 labToRgb */
static void lab_to_rgb(const float* lab, float* rgb) {
  float y = (lab[0] + 16.0f) / 116.0f;
  float x = lab[1] / 500.0f + y;
  float z = y - lab[2] / 200.0f;
  x = lab_finv(x);
  y = lab_finv(y);
  z = lab_finv(z);
  x *= 0.9505f;
  y *= 1.0f;
  z *= 1.089f;
  float r = x * 3.2406f + y * -1.5372f + z * -0.4986f;
  float g = x * -0.9689f + y * 1.8758f + z * 0.0415f;
  float b = x * 0.0557f + y * -0.2040f + z * 1.0570f;
  r = linear_to_srgb(r);
  g = linear_to_srgb(g);
  b = linear_to_srgb(b);
  const float m = maxf(r, maxf(g, b));
  if (m > 1.0f) {
    r /= m;
    g /= m;
    b /= m;
  }
  rgb[0] = maxf(r, 0.0f);
  rgb[1] = maxf(g, 0.0f);
  rgb[2] = maxf(b, 0.0f);
  rgb[3] = lab[3];
}

/* VolumePainter.cpp:331-379 mapColorValue (useClamping is always true, :448) */
static void map_color(const orc_table* t, float value, float* out) {
  if (!isfinite(value)) { /* nanColor; unreachable for table sampling */
    out[0] = t->lab ? 1.0f : 0.25f;
    out[1] = 0.0f;
    out[2] = 0.0f;
    out[3] = 1.0f;
    return;
  }
  if (t->ncolors == 0) { /* belowRange */
    out[0] = out[1] = out[2] = 0.0f;
    out[3] = 1.0f;
    return;
  }
  const orc_cnode* first = &t->colors[0];
  const orc_cnode* last = &t->colors[t->ncolors - 1];
  /* nodes are sorted, so "< first", "> last", "== first", "== last" in the
   * reference's order collapses to two tests */
  if (value < first->value || (!(value > last->value) && value == first->value)) {
    out[0] = first->r;
    out[1] = first->g;
    out[2] = first->b;
    out[3] = 1.0f;
    return;
  }
  if (value > last->value || value == last->value) {
    out[0] = last->r;
    out[1] = last->g;
    out[2] = last->b;
    out[3] = 1.0f;
    return;
  }
  for (int i = 1; i < t->ncolors; ++i) {
    const orc_cnode* right = &t->colors[i];
    if (right->value >= value) {
      const orc_cnode* left = &t->colors[i - 1];
      const float span = right->value - left->value;
      const float u = (span > 0.0f) ? (value - left->value) / span : 0.0f;
      float a[4] = {left->r, left->g, left->b, 1.0f};
      float b[4] = {right->r, right->g, right->b, 1.0f};
      if (t->lab) {
        float la[4], lb[4], mix[4];
        rgb_to_lab(a, la);
        rgb_to_lab(b, lb);
        for (int c = 0; c < 4; ++c) mix[c] = la[c] + (lb[c] - la[c]) * u;
        lab_to_rgb(mix, out);
        return;
      }
      for (int c = 0; c < 4; ++c) out[c] = a[c] + (b[c] - a[c]) * u;
      return;
    }
  }
  out[0] = last->r;
  out[1] = last->g;
  out[2] = last->b;
  out[3] = 1.0f;
}

/* VolumePainter.cpp:381-440 mapOpacityValue */
static float map_opacity(const orc_table* t, float value) {
  if (!isfinite(value)) return 1.0f;
  if (t->nopacity == 0) return 1.0f;
  const orc_onode* first = &t->opacity[0];
  const orc_onode* last = &t->opacity[t->nopacity - 1];
  if (value <= first->value) return first->alpha;
  if (value >= last->value) return last->alpha;
  for (int i = 1; i < t->nopacity; ++i) {
    const orc_onode* right = &t->opacity[i];
    if (right->value >= value) {
      const orc_onode* left = &t->opacity[i - 1];
      const float span = right->value - left->value;
      float w = (span > 0.0f) ? (value - left->value) / span : 0.0f;
      if (w < left->midpoint) {
        w = 0.5f * w / left->midpoint;
      } else {
        w = 0.5f + 0.5f * (w - left->midpoint) / (1.0f - left->midpoint);
      }
      if (left->sharpness == 1.0f) return (w < 0.5f) ? left->alpha : right->alpha;
      if (left->sharpness == 0.0f)
        return left->alpha + (right->alpha - left->alpha) * w;
      if (w < 0.5f) {
        w = 0.5f * powf(w * 2.0f, 1.0f + 10.0f * left->sharpness);
      } else if (w > 0.5f) {
        w = 1.0f - 0.5f * powf((1.0f - w) * 2.0f, 1.0f + 10.0f * left->sharpness);
      }
      const float ww = w * w;
      const float www = ww * w;
      const float h1 = 2.0f * www - 3.0f * ww + 1.0f;
      const float h2 = -2.0f * www + 3.0f * ww;
      const float h3 = www - 2.0f * ww + w;
      const float h4 = www - ww;
      const float slope = right->alpha - left->alpha;
      const float tt = (1.0f - left->sharpness) * slope;
      float result = h1 * left->alpha + h2 * right->alpha + h3 * tt + h4 * tt;
      result = maxf(result, minf(left->alpha, right->alpha));
      result = minf(result, maxf(left->alpha, right->alpha));
      return result;
    }
  }
  return last->alpha;
}

/* VolumePainter.cpp:442-516 buildColorTable */
void orc_color_table(float alpha_scale, float normalization_factor,
                     float range_min, float range_max,
                     const amrvr_colormap_pt* pts, int npts, float* table) {
  static _Thread_local orc_table t;
  t.ncolors = 0;
  t.nopacity = 0;
  if (pts != 0 && npts > 0) {
    t.lab = 1;
    for (int i = 0; i < npts; ++i) {
      orc_cnode c = {pts[i].value, clampf(pts[i].red, 0.0f, 1.0f),
                     clampf(pts[i].green, 0.0f, 1.0f),
                     clampf(pts[i].blue, 0.0f, 1.0f)};
      insert_color(&t, c);
      orc_onode o = {pts[i].value,
                     scaled_alpha(pts[i].alpha, alpha_scale,
                                  normalization_factor),
                     0.5f, 0.0f};
      insert_opacity(&t, o);
    }
  } else {
    t.lab = 0;
    static const orc_cnode jet[7] = {
        {0.0f, 0.0f, 0.0f, 0.5625f},      {0.111111f, 0.0f, 0.0f, 1.0f},
        {0.3650795f, 0.0f, 1.0f, 1.0f},   {0.4920635f, 0.5f, 1.0f, 0.5f},
        {0.6190475f, 1.0f, 1.0f, 0.0f},   {0.873016f, 1.0f, 0.0f, 0.0f},
        {1.0f, 0.5f, 0.0f, 0.0f}};
    for (int i = 0; i < 7; ++i) insert_color(&t, jet[i]);
    static const float positions[6] = {0.0f, 0.15f, 0.35f, 0.6f, 0.85f, 1.0f};
    static const float alphas[6] = {0.05f, 0.15f, 0.22f, 0.3f, 0.38f, 0.5f};
    const float span = range_max - range_min;
    for (int i = 0; i < 6; ++i) {
      orc_onode o = {positions[i] * span + range_min,
                     scaled_alpha(alphas[i], alpha_scale, normalization_factor),
                     0.5f, 0.0f};
      insert_opacity(&t, o);
    }
    rescale_table(&t, range_min, range_max);
  }
  const float span = range_max - range_min;
  for (int i = 0; i < AMRVR_COLOR_TABLE_SIZE; ++i) {
    const float u = (float)i / (float)(AMRVR_COLOR_TABLE_SIZE - 1);
    const float value = range_min + span * u;
    float rgba[4];
    map_color(&t, value, rgba);
    rgba[3] = map_opacity(&t, value);
    memcpy(table + 4 * i, rgba, sizeof(rgba));
  }
}
