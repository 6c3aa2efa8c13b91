/* oracle/restate/orc_post.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates the tail of VolumeRenderer::renderSingleTrial
 * (VolumeRenderer/VolumeRenderer.cpp:1103-1339): downsampleImage (:479-528),
 * renderBoundingBoxLayer (:139-335), the byte rule of Common/Color.hpp:66-91
 * with SavePPM's row order (Common/SavePPM.cpp:25-32), the automatic camera
 * (:974-1023), and the frame driver itself. */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static const float kPi = 3.14159265358979323846f;

static float clampf(float v, float lo, float hi) {
  return (v < lo) ? lo : ((hi < v) ? hi : v);
}
static float maxf(float a, float b) { return (a < b) ? b : a; }
static float minf(float a, float b) { return (b < a) ? b : a; }

void orc_downsample(const float* src5, int width, int height, int sqrt_aa,
                    float* dst5) {
  const int block = (sqrt_aa > 1) ? sqrt_aa : 1;
  const int src_w = width * block;
  const float inv = 1.0f / (float)(block * block);
  for (int y = 0; y < height; ++y)
    for (int x = 0; x < width; ++x) {
      float sum[4] = {0.0f, 0.0f, 0.0f, 0.0f};
      for (int dy = 0; dy < block; ++dy)
        for (int dx = 0; dx < block; ++dx) {
          const float* s =
              src5 + 5 * ((size_t)(y * block + dy) * src_w + (x * block + dx));
          for (int c = 0; c < 4; ++c) sum[c] += s[c];
        }
      float* d = dst5 + 5 * ((size_t)y * width + x);
      for (int c = 0; c < 4; ++c) d[c] = sum[c] * inv;
      d[4] = INFINITY;
    }
}

static void camera_frame(const amrvr_camera* c, double* f, double* r, double* u) {
  const double d[3] = {c->look_at[0] - c->eye[0], c->look_at[1] - c->eye[1],
                       c->look_at[2] - c->eye[2]};
  const double len = sqrt(d[0] * d[0] + d[1] * d[1] + d[2] * d[2]);
  if (len > 0.0 && isfinite(len)) {
    f[0] = d[0] / len;
    f[1] = d[1] / len;
    f[2] = d[2] / len;
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
    r[0] /= rlen;
    r[1] /= rlen;
    r[2] /= rlen;
  } else {
    r[0] = 1.0;
    r[1] = 0.0;
    r[2] = 0.0;
  }
  u[0] = r[1] * f[2] - r[2] * f[1];
  u[1] = r[2] * f[0] - r[0] * f[2];
  u[2] = r[0] * f[1] - r[1] * f[0];
}

static double dot3(const double* a, const double* b) {
  return a[0] * b[0] + a[1] * b[1] + a[2] * b[2];
}

/* :228-252 blendSample */
static void overlay_sample(float* image5, int width, int height, int px, int py,
                           float coverage) {
  if (px < 0 || px >= width || py < 0 || py >= height) return;
  const float a = clampf(coverage, 0.0f, 1.0f);
  if (a <= 0.0f) return;
  float* b = image5 + 5 * ((size_t)py * width + px);
  const float src = 1.0f * a; /* white line colour */
  b[0] = src + b[0] * (1.0f - a);
  b[1] = src + b[1] * (1.0f - a);
  b[2] = src + b[2] * (1.0f - a);
  b[3] = a + b[3] * (1.0f - a);
  b[4] = -FLT_MAX; /* numeric_limits<float>::lowest() */
}

void orc_overlay(const double* tight_min, const double* tight_max,
                 const amrvr_camera* camera, int sqrt_aa, int width, int height,
                 float* image5) {
  if (width <= 0 || height <= 0) return;
  const float aspect = (float)width / (float)((height > 1) ? height : 1);
  double f[3], r[3], u[3];
  camera_frame(camera, f, r, u);
  const float tan_half = tanf(camera->fov_y_degrees * 0.5f * kPi / 180.0f);
  const float w_scale = (width > 1) ? (float)(width - 1) : 0.0f;
  const float h_scale = (height > 1) ? (float)(height - 1) : 0.0f;

  float sx[8], sy[8];
  int valid[8];
  for (int i = 0; i < 8; ++i) { /* :179-209 */
    const double corner[3] = {(i & 1) ? tight_max[0] : tight_min[0],
                              (i & 2) ? tight_max[1] : tight_min[1],
                              (i & 4) ? tight_max[2] : tight_min[2]};
    const double rel[3] = {corner[0] - camera->eye[0], corner[1] - camera->eye[1],
                           corner[2] - camera->eye[2]};
    valid[i] = 0;
    sx[i] = sy[i] = 0.0f;
    const float depth = (float)dot3(rel, f);
    if (!(depth > 0.0f) || !isfinite(depth)) continue;
    const float xc = (float)dot3(rel, r);
    const float yc = (float)dot3(rel, u);
    const float nx = xc / (depth * tan_half * aspect);
    const float ny = yc / (depth * tan_half);
    if (!isfinite(nx) || !isfinite(ny)) continue;
    sx[i] = (nx * 0.5f + 0.5f) * w_scale;
    sy[i] = (ny * 0.5f + 0.5f) * h_scale;
    valid[i] = 1;
  }
  static const int edges[12][2] = {{0, 1}, {1, 3}, {3, 2}, {2, 0}, {4, 5}, {5, 7},
                                   {7, 6}, {6, 4}, {0, 4}, {1, 5}, {2, 6}, {3, 7}};
  const float pixel_radius = 0.5f * (float)((sqrt_aa > 1) ? sqrt_aa : 1);
  const float influence = pixel_radius + 0.5f;
  const float coverage_scale = 0.6f;
  for (int e = 0; e < 12; ++e) {
    const int a = edges[e][0], b = edges[e][1];
    if (!valid[a] || !valid[b]) continue;
    const float min_x = minf(sx[a], sx[b]) - influence;
    const float max_x = maxf(sx[a], sx[b]) + influence;
    const float min_y = minf(sy[a], sy[b]) - influence;
    const float max_y = maxf(sy[a], sy[b]) + influence;
    int x0 = (int)floorf(min_x), x1 = (int)ceilf(max_x);
    int y0 = (int)floorf(min_y), y1 = (int)ceilf(max_y);
    if (x0 < 0) x0 = 0;
    if (x1 > width - 1) x1 = width - 1;
    if (y0 < 0) y0 = 0;
    if (y1 > height - 1) y1 = height - 1;
    const float ex = sx[b] - sx[a], ey = sy[b] - sy[a];
    const float len_sq = ex * ex + ey * ey;
    if (!(len_sq > 0.0f)) {
      overlay_sample(image5, width, height, (int)lroundf(sx[a]),
                     (int)lroundf(sy[a]), 1.0f);
      continue;
    }
    for (int py = y0; py <= y1; ++py) {
      const float cy = (float)py + 0.5f;
      for (int px = x0; px <= x1; ++px) {
        const float cx = (float)px + 0.5f;
        const float apx = cx - sx[a], apy = cy - sy[a];
        float t = (apx * ex + apy * ey) / len_sq;
        t = clampf(t, 0.0f, 1.0f);
        const float qx = sx[a] + (sx[b] - sx[a]) * t;
        const float qy = sy[a] + (sy[b] - sy[a]) * t;
        const float dx = cx - qx, dy = cy - qy;
        const float dist = sqrtf(dx * dx + dy * dy);
        const float coverage =
            clampf((pixel_radius + 0.5f - dist) * coverage_scale, 0.0f, 1.0f);
        if (coverage <= 0.0f) continue;
        overlay_sample(image5, width, height, px, py, coverage);
      }
    }
  }
}

void orc_quantize_rgb8(const float* image5, int width, int height,
                       uint8_t* rgb8) {
  size_t o = 0;
  for (int y = height - 1; y >= 0; --y)
    for (int x = 0; x < width; ++x) {
      const float* p = image5 + 5 * ((size_t)y * width + x);
      for (int c = 0; c < 3; ++c) {
        const int v = (int)(p[c] * 256.f);
        rgb8[o++] = (uint8_t)((v < 0) ? 0 : ((v > 255) ? 255 : v));
      }
    }
}

void orc_auto_camera(const double* bounds_min, const double* bounds_max,
                     float azimuth, float altitude, const double* up_or_null,
                     amrvr_camera* out) {
  double center[3], half[3];
  for (int d = 0; d < 3; ++d) {
    center[d] = 0.5f * (bounds_min[d] + bounds_max[d]);
    half[d] = 0.5f * (bounds_max[d] - bounds_min[d]);
  }
  float radius = (float)sqrt(half[0] * half[0] + half[1] * half[1] + half[2] * half[2]);
  if (radius <= 0.0f) radius = 1.0f;
  const float fov_y = kPi * 0.25f;
  const float half_fov = fov_y * 0.5f;
  const float min_distance =
      (half_fov > 0.0f) ? radius / (float)tanf(half_fov) : radius;
  const float margin = maxf(0.25f * radius, 0.5f);
  const float distance = min_distance + margin;
  const float cos_alt = cosf(altitude);
  /* Vec3 eye(center + float products): float * float, promoted on the add */
  out->eye[0] = center[0] + distance * cos_alt * sinf(azimuth);
  out->eye[1] = center[1] + distance * sinf(altitude);
  out->eye[2] = center[2] + distance * cos_alt * cosf(azimuth);
  double up[3] = {0.0, 1.0, 0.0};
  if (up_or_null) memcpy(up, up_or_null, sizeof(up));
  const double raw[3] = {center[0] - out->eye[0], center[1] - out->eye[1],
                         center[2] - out->eye[2]};
  double vd[3];
  {
    const double len = sqrt(raw[0] * raw[0] + raw[1] * raw[1] + raw[2] * raw[2]);
    if (len > 0.0 && isfinite(len)) {
      for (int d = 0; d < 3; ++d) vd[d] = raw[d] / len;
    } else {
      vd[0] = 0.0;
      vd[1] = 0.0;
      vd[2] = -1.0;
    }
  }
  for (int attempt = 0; attempt < 2; ++attempt) { /* :1008-1013 */
    const double c[3] = {vd[1] * up[2] - vd[2] * up[1], vd[2] * up[0] - vd[0] * up[2],
                         vd[0] * up[1] - vd[1] * up[0]};
    if (sqrt(c[0] * c[0] + c[1] * c[1] + c[2] * c[2]) <= 1e-4f) {
      if (attempt == 0) {
        up[0] = 0.0;
        up[1] = 0.0;
        up[2] = 1.0;
      } else {
        up[0] = 1.0;
        up[1] = 0.0;
        up[2] = 0.0;
      }
    } else {
      break;
    }
  }
  {
    const double len = sqrt(up[0] * up[0] + up[1] * up[1] + up[2] * up[2]);
    if (len > 0.0 && isfinite(len)) {
      for (int d = 0; d < 3; ++d) up[d] = up[d] / len;
    } else {
      up[0] = 0.0;
      up[1] = 0.0;
      up[2] = -1.0;
    }
  }
  for (int d = 0; d < 3; ++d) {
    out->look_at[d] = center[d];
    out->up[d] = up[d];
  }
  out->fov_y_degrees = fov_y * 180.0f / kPi;
  out->near_plane = 0.1f;
  out->far_plane = distance * 4.0f;
}

int orc_render_frame(int nranks, int nboxes, const amrvr_box* boxes,
                     const int32_t* owners, const amrvr_scalar_transform* t,
                     float range_min, float range_max, const double* bounds_min,
                     const double* bounds_max, const amrvr_camera* camera,
                     const amrvr_frame_params* params,
                     const amrvr_colormap_pt* pts, int npts, float* final5,
                     uint8_t* rgb8, int* rank_order_out, long long* samples) {
  const int sqrt_aa = (int)lround(sqrt((double)params->antialiasing));
  const int mult = (sqrt_aa > 1) ? sqrt_aa : 1;
  const int rw = params->width * mult, rh = params->height * mult;
  const float aspect = (float)params->width / (float)params->height;
  const size_t layer_floats = (size_t)rw * rh * 5;

  /* global table in owner-major order (= MPI_Allgatherv order) */
  amrvr_box_record* recs =
      (amrvr_box_record*)calloc((size_t)nboxes + 1, sizeof(amrvr_box_record));
  int* src = (int*)malloc(sizeof(int) * (size_t)(nboxes + 1));
  int n = 0;
  for (int rank = 0; rank < nranks; ++rank) {
    int local = 0;
    for (int i = 0; i < nboxes; ++i) {
      if (owners[i] != rank) continue;
      for (int d = 0; d < 3; ++d) {
        recs[n].min_corner[d] = boxes[i].min_corner[d];
        recs[n].max_corner[d] = boxes[i].max_corner[d];
        recs[n].dims[d] = boxes[i].dims[d];
      }
      recs[n].owner = rank;
      recs[n].local_index = local++;
      src[n] = i;
      ++n;
    }
  }
  const float ref_distance =
      orc_reference_sample_distance(n, recs, bounds_min, bounds_max);

  float** layers = (float**)calloc((size_t)n + 1, sizeof(float*));
  int* layer_owner = (int*)malloc(sizeof(int) * (size_t)(n + 1));
  float* hints = (float*)malloc(sizeof(float) * (size_t)(n + 1));
  long long total = 0;
  for (int g = 0; g < n; ++g) { /* :1201-1219 */
    layers[g] = (float*)malloc(sizeof(float) * layer_floats);
    long long count = 0;
    orc_paint(&boxes[src[g]], t, range_min, range_max, params->box_transparency,
              ref_distance, rw, rh, camera, pts, npts, layers[g], &count);
    total += count;
    layer_owner[g] = recs[g].owner;
    hints[g] = orc_depth_hint(&recs[g], camera);
  }
  int* rank_order = (int*)malloc(sizeof(int) * (size_t)nranks);
  orc_order(nranks, n, recs, camera, aspect, params->use_visibility_graph,
            rank_order);
  float* full = (float*)malloc(sizeof(float) * layer_floats);
  orc_compose(nranks, rw, rh, n, layer_owner, (const float* const*)layers, hints,
              rank_order, full, 0);
  for (int g = 0; g < n; ++g) free(layers[g]);

  const size_t out_floats = (size_t)params->width * params->height * 5;
  float* image = (float*)malloc(sizeof(float) * out_floats);
  if (sqrt_aa > 1) {
    orc_downsample(full, params->width, params->height, sqrt_aa, image);
  } else {
    memcpy(image, full, sizeof(float) * out_floats);
  }
  if (params->draw_bounding_box) { /* computeTightBounds :791-849 + :1311-1314 */
    double tmin[3], tmax[3];
    if (n > 0) {
      for (int d = 0; d < 3; ++d) {
        double lo = recs[0].min_corner[d], hi = recs[0].max_corner[d];
        for (int g = 1; g < n; ++g) {
          if (recs[g].min_corner[d] < lo) lo = recs[g].min_corner[d];
          if (recs[g].max_corner[d] > hi) hi = recs[g].max_corner[d];
        }
        tmin[d] = (double)(float)lo;
        tmax[d] = (double)(float)hi;
      }
    } else {
      memcpy(tmin, bounds_min, sizeof(tmin));
      memcpy(tmax, bounds_max, sizeof(tmax));
    }
    orc_overlay(tmin, tmax, camera, 1, params->width, params->height, image);
  }
  if (final5) memcpy(final5, image, sizeof(float) * out_floats);
  if (rgb8) orc_quantize_rgb8(image, params->width, params->height, rgb8);
  if (rank_order_out) memcpy(rank_order_out, rank_order, sizeof(int) * (size_t)nranks);
  if (samples) *samples = total;
  free(image);
  free(full);
  free(rank_order);
  free(hints);
  free(layer_owner);
  free(layers);
  free(src);
  free(recs);
  return 0;
}
