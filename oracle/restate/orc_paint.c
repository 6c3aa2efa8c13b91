/* oracle/restate/orc_paint.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates VolumePainter::paint (Common/VolumePainter.cpp:548-961), the scalar
 * transform (Common/VolumeTypes.hpp:33-67) and the soft clip (:75-105). */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static const float kPi = 3.14159265358979323846f; /* VolumePainter.cpp:37 */
static const float kSoftClipTolerance = 1e-5f;     /* :36 */

static float clampf(float v, float lo, float hi) {
  return (v < lo) ? lo : ((hi < v) ? hi : v);
}
static float maxf(float a, float b) { return (a < b) ? b : a; }

/* VolumeTypes.hpp:33-67 applyScalarTransform (double in, float out) */
static float apply_transform(double raw, const amrvr_scalar_transform* t) {
  double v = isfinite(raw) ? raw : 0.0; /* sanitizeScalarSample */
  if (t->log_scale_input) {              /* toProcessedScalar */
    if (!(v > 0.0)) {
      v = t->positive_floor;
    } else if (v < t->positive_floor) {
      v = t->positive_floor;
    }
    v = log(v);
  }
  if (t->normalize_to_unit_range) {
    v = (v - t->normalization_min) * t->inverse_normalization_span;
    if (v < 0.0) {
      v = 0.0;
    } else if (v > 1.0) {
      v = 1.0;
    }
  }
  return (float)v;
}

/* VolumePainter.cpp:75-105 saturateSoftTail */
static float soft_tail(float value, float clip_start, float rolloff_end) {
  const float end = maxf(clip_start, rolloff_end);
  float c = value;
  if (c < 0.0f) {
    c = 0.0f;
  } else if (c > end) {
    c = end;
  }
  if (!(end > clip_start + kSoftClipTolerance)) return c;
  if (!(c > clip_start)) return c;
  if (!(c < end)) return end;
  const float n = (c - clip_start) / (end - clip_start);
  const float smooth = n + n * n - n * n * n;
  return clip_start + (end - clip_start) * smooth;
}

/* The per-sample value path of VolumePainter.cpp:870-883 for one raw cell. */
int orc_cell_index(double raw, const amrvr_scalar_transform* t,
                   float range_min, float range_max) {
  float inverse_range = 1.0f; /* :719-722 */
  if (range_max != range_min) inverse_range = 1.0f / (range_max - range_min);
  const float clip_start = clampf(range_max, 0.0f, 1.0f); /* :723 */
  const int apply_clip = 1.0f > clip_start + kSoftClipTolerance;
  float scalar = apply_transform(raw, t);
  if (apply_clip) scalar = soft_tail(scalar, clip_start, 1.0f);
  float normalized = (scalar - range_min) * inverse_range;
  normalized = (normalized < 0.0f) ? 0.0f : normalized;
  normalized = (normalized > 1.0f) ? 1.0f : normalized;
  int idx = (int)(normalized * (float)(AMRVR_COLOR_TABLE_SIZE - 1));
  idx = (idx < 0) ? 0 : idx;
  idx = (idx > AMRVR_COLOR_TABLE_SIZE - 1) ? (AMRVR_COLOR_TABLE_SIZE - 1) : idx;
  return idx;
}

static double box_value(const amrvr_box* b, int i, int j, int k) {
  /* amrex::Array4::operator() with the box's strides */
  return b->data[(long long)(b->lo[0] + i - b->begin[0]) +
                 (long long)(b->lo[1] + j - b->begin[1]) * b->jstride +
                 (long long)(b->lo[2] + k - b->begin[2]) * b->kstride +
                 (long long)b->component * b->nstride];
}

void orc_quantize_box(const amrvr_box* box, const amrvr_scalar_transform* t,
                      float range_min, float range_max, uint8_t* out) {
  size_t n = 0;
  for (int k = 0; k < box->dims[2]; ++k)
    for (int j = 0; j < box->dims[1]; ++j)
      for (int i = 0; i < box->dims[0]; ++i)
        out[n++] = (uint8_t)orc_cell_index(box_value(box, i, j, k), t,
                                           range_min, range_max);
}

/* VolumePainter.cpp:571-610 */
void orc_box_sampling(const amrvr_box* box, const double* bounds_min,
                      const double* bounds_max, float reference_sample_distance,
                      float* sample_distance, float* normalization_factor) {
  double spacing[3] = {0.0, 0.0, 0.0};
  for (int d = 0; d < 3; ++d)
    if (box->dims[d] > 0)
      spacing[d] = (box->max_corner[d] - box->min_corner[d]) / (double)box->dims[d];
  float min_spacing = FLT_MAX;
  for (int d = 0; d < 3; ++d) {
    const float v = (float)spacing[d];
    if (v > 0.0f && v < min_spacing && isfinite(v)) min_spacing = v;
  }
  if (!(min_spacing > 0.0f && isfinite(min_spacing))) {
    double m = bounds_max[0] - bounds_min[0];
    for (int d = 1; d < 3; ++d) {
      const double s = bounds_max[d] - bounds_min[d];
      if (s < m) m = s;
    }
    min_spacing = maxf(1e-4f, (float)m * 0.01f);
  }
  const float step = maxf(min_spacing * 0.5f, 1e-5f);
  float ref = reference_sample_distance;
  if (!(ref > 0.0f && isfinite(ref))) ref = step;
  float norm = step / ref;
  if (!isfinite(norm)) norm = 1.0f;
  norm = maxf(norm, 0.0f);
  *sample_distance = step;
  *normalization_factor = norm;
}

/* Camera basis of VolumePainter.cpp:631-656 (double math, then float). */
static void camera_basis(const amrvr_camera* c, float* fwd, float* right,
                         float* up) {
  double f[3], r[3], u[3];
  for (int d = 0; d < 3; ++d) f[d] = c->look_at[d] - c->eye[d];
  const double flen = sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
  if (flen > 0.0 && isfinite(flen)) { /* CameraUtils.hpp:16-22 safeNormalize */
    for (int d = 0; d < 3; ++d) f[d] = f[d] / flen;
  } else {
    f[0] = 0.0;
    f[1] = 0.0;
    f[2] = -1.0;
  }
  r[0] = f[1] * c->up[2] - f[2] * c->up[1];
  r[1] = f[2] * c->up[0] - f[0] * c->up[2];
  r[2] = f[0] * c->up[1] - f[1] * c->up[0];
  const double rlen = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
  if (rlen > 0.0 && isfinite(rlen)) {
    for (int d = 0; d < 3; ++d) r[d] /= rlen;
  } else {
    r[0] = 1.0;
    r[1] = 0.0;
    r[2] = 0.0;
  }
  u[0] = r[1] * f[2] - r[2] * f[1];
  u[1] = r[2] * f[0] - r[0] * f[2];
  u[2] = r[0] * f[1] - r[1] * f[0];
  for (int d = 0; d < 3; ++d) {
    fwd[d] = (float)f[d];
    right[d] = (float)r[d];
    up[d] = (float)u[d];
  }
}

typedef struct {
  float tmin, tmax;
} slab_t;

/* VolumePainter.cpp:775-796 updateBounds */
static void slab_update(slab_t* s, float origin, float direction, float lo,
                        float hi) {
  if (fabsf(direction) < 1e-8f) {
    if (origin < lo || origin > hi) {
      s->tmin = INFINITY;
      s->tmax = -INFINITY;
    }
    return;
  }
  const float inv = 1.0f / direction;
  float t1 = (lo - origin) * inv;
  float t2 = (hi - origin) * inv;
  if (t1 > t2) {
    const float tmp = t1;
    t1 = t2;
    t2 = tmp;
  }
  s->tmin = (s->tmin > t1) ? s->tmin : t1;
  s->tmax = (s->tmax < t2) ? s->tmax : t2;
}

void orc_paint(const amrvr_box* box, const amrvr_scalar_transform* t,
               float range_min, float range_max, float box_transparency,
               float reference_sample_distance, int width, int height,
               const amrvr_camera* camera, const amrvr_colormap_pt* pts,
               int npts, float* out5, long long* samples) {
  long long count = 0;
  if (samples) *samples = 0;
  if (width <= 0 || height <= 0) return;

  float step, norm;
  orc_box_sampling(box, box->min_corner, box->max_corner,
                   reference_sample_distance, &step, &norm);
  const float alpha_scale = clampf(1.0f - box_transparency, 0.0f, 1.0f);
  float table[AMRVR_COLOR_TABLE_SIZE * 4];
  orc_color_table(alpha_scale, norm, range_min, range_max, pts, npts, table);

  const float aspect = (float)width / (float)((height > 1) ? height : 1);
  float fwd[3], right[3], up[3];
  camera_basis(camera, fwd, right, up);
  const float eye[3] = {(float)camera->eye[0], (float)camera->eye[1],
                        (float)camera->eye[2]};
  const float lo[3] = {(float)box->min_corner[0], (float)box->min_corner[1],
                       (float)box->min_corner[2]};
  const float hi[3] = {(float)box->max_corner[0], (float)box->max_corner[1],
                       (float)box->max_corner[2]};
  const int nx = box->dims[0], ny = box->dims[1], nz = box->dims[2];
  const size_t npix = (size_t)width * (size_t)height;
  if (nx <= 0 || ny <= 0 || nz <= 0) { /* :670-673 clear() */
    for (size_t p = 0; p < npix; ++p) {
      out5[5 * p + 0] = out5[5 * p + 1] = out5[5 * p + 2] = out5[5 * p + 3] = 0.0f;
      out5[5 * p + 4] = INFINITY;
    }
    return;
  }
  const float dx = (hi[0] - lo[0]) / (float)nx;
  const float dy = (hi[1] - lo[1]) / (float)ny;
  const float dz = (hi[2] - lo[2]) / (float)nz;
  const float extent =
      sqrtf((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1]) +
            (hi[2] - lo[2]) * (hi[2] - lo[2]));
  const float eps = extent * 0.0001f;

  /* K5-equivalent hoist is NOT applied here: the oracle evaluates the value
   * path per sample exactly like the reference (:870-883). */
  const float tan_half = tanf(camera->fov_y_degrees * 0.5f * kPi / 180.0f);
  const float inv_w = 1.0f / (float)width;
  const float inv_h = 1.0f / (float)height;

  for (int index = 0; index < (int)npix; ++index) {
    const int px = index % width;
    const int py = index / width;
    const float ndc_x = ((float)px + 0.5f) * inv_w * 2.0f - 1.0f;
    const float ndc_y = ((float)py + 0.5f) * inv_h * 2.0f - 1.0f;
    const float plane_x = ndc_x * tan_half * aspect;
    const float plane_y = ndc_y * tan_half;
    float dir[3];
    for (int d = 0; d < 3; ++d)
      dir[d] = fwd[d] + plane_x * right[d] + plane_y * up[d];
    const float len_sq = dir[0] * dir[0] + dir[1] * dir[1] + dir[2] * dir[2];
    /* amrex::Math::rsqrt on the host is 1./sqrt(x) (double divide, innocuous
     * double rounding) */
    const float len =
        (len_sq > 0.0f) ? (1.0f / (float)(1.0 / (double)sqrtf(len_sq))) : 0.0f;
    if (len > 0.0f) {
      const float inv = 1.0f / len;
      dir[0] *= inv;
      dir[1] *= inv;
      dir[2] *= inv;
    }
    slab_t s = {-INFINITY, INFINITY};
    for (int d = 0; d < 3; ++d) slab_update(&s, eye[d], dir[d], lo[d], hi[d]);

    float* px_out = out5 + 5 * (size_t)index;
    if (!(s.tmax >= s.tmin)) {
      px_out[0] = px_out[1] = px_out[2] = px_out[3] = 0.0f;
      px_out[4] = INFINITY;
      continue;
    }
    float distance = s.tmin + eps;
    if (distance < 0.0f) distance = eps;
    float acc[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    float pos[3];
    for (int d = 0; d < 3; ++d) pos[d] = eye[d] + dir[d] * distance;
#define ORC_INSIDE()                                                       \
  (!(pos[0] < lo[0] || pos[0] > hi[0] || pos[1] < lo[1] || pos[1] > hi[1] || \
     pos[2] < lo[2] || pos[2] > hi[2]))
    while (distance < s.tmax && !ORC_INSIDE()) { /* :830-835 */
      distance += step;
      for (int d = 0; d < 3; ++d) pos[d] = eye[d] + dir[d] * distance;
    }
    while (distance < s.tmax && acc[3] < 1.0f) { /* :837-900 */
      if (ORC_INSIDE()) {
        const float fx = (pos[0] - lo[0]) / dx;
        const float fy = (pos[1] - lo[1]) / dy;
        const float fz = (pos[2] - lo[2]) / dz;
        int i = (int)floorf(fx);
        int j = (int)floorf(fy);
        int k = (int)floorf(fz);
        i = (i < 0) ? 0 : ((i >= nx) ? nx - 1 : i);
        j = (j < 0) ? 0 : ((j >= ny) ? ny - 1 : j);
        k = (k < 0) ? 0 : ((k >= nz) ? nz - 1 : k);
        const int idx =
            orc_cell_index(box_value(box, i, j, k), t, range_min, range_max);
        ++count;
        const float* e = table + 4 * idx;
        const float a = e[3] * (1.0f - acc[3]);
        acc[0] += e[0] * a;
        acc[1] += e[1] * a;
        acc[2] += e[2] * a;
        acc[3] += a;
      }
      distance += step;
      for (int d = 0; d < 3; ++d) pos[d] = eye[d] + dir[d] * distance;
    }
#undef ORC_INSIDE
    for (int c = 0; c < 4; ++c) acc[c] = (acc[c] > 1.0f) ? 1.0f : acc[c];
    float depth = INFINITY;
    if (acc[3] > 0.0f) { /* :912-921 entry depth along the view axis */
      const float ex = eye[0] + dir[0] * s.tmin;
      const float ey = eye[1] + dir[1] * s.tmin;
      const float ez = eye[2] + dir[2] * s.tmin;
      depth = (ex - eye[0]) * fwd[0] + (ey - eye[1]) * fwd[1] +
              (ez - eye[2]) * fwd[2];
    }
    /* host epilogue :939-955 */
    for (int c = 0; c < 4; ++c) px_out[c] = clampf(acc[c], 0.0f, 1.0f);
    if (!isfinite(depth) || px_out[3] <= 0.0f) depth = INFINITY;
    px_out[4] = depth;
  }
  if (samples) *samples = count;
}
