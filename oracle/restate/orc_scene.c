/* oracle/restate/orc_scene.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates the scene ingest either side of the hot path:
 *   detail::BuildSceneGeometry   VolumeRenderer/SceneBuilder.cpp:113-425
 *   reduceLocalScalarStats       VolumeRenderer/SceneBuilder.cpp:53-94
 *   SetSceneNormalizationRange   VolumeRenderer/SceneBuilder.cpp:427-443
 *   detail::ComputeSceneHistogram VolumeRenderer/SceneBuilder.cpp:445-577
 * for ALREADY-CONVEX levels (amrex::convexify is un-vendored AMReX).  The levels
 * passed in hold the FABs of ALL ranks: every reduction of the reference
 * (MPI_MIN / MPI_MAX / MPI_SUM) is exact and order-independent, so the
 * rank-independent outputs do not depend on the box->rank map. */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <string.h>

static double fab_value(const amrvr_box* b, int i, int j, int k) {
  /* amrex::Array4::operator()(i,j,k,n) */
  return b->data[(i - b->begin[0]) + (int64_t)(j - b->begin[1]) * b->jstride +
                 (int64_t)(k - b->begin[2]) * b->kstride +
                 (int64_t)b->component * b->nstride];
}

/* Common/VolumeTypes.hpp:33-67 applyScalarTransform */
static float apply_transform(double raw, const amrvr_scalar_transform* t) {
  double v = isfinite(raw) ? raw : 0.0;
  if (t->log_scale_input) {
    if (!(v > 0.0)) v = t->positive_floor;
    else if (v < t->positive_floor) v = t->positive_floor;
    v = log(v);
  }
  if (t->normalize_to_unit_range) {
    v = (v - t->normalization_min) * t->inverse_normalization_span;
    if (v < 0.0) v = 0.0;
    else if (v > 1.0) v = 1.0;
  }
  return (float)v;
}

static void make_range(double lo, double hi, float* out) { /* :103-109 */
  if (lo == hi) hi = lo + 1.0;
  out[0] = (float)lo;
  out[1] = (float)hi;
}

/* status: 0 ok, 1 "no data", 2 "invalid scalar range", 3 no positive value for
 * the log scale, 4 normalisation span not finite, 5 boxes_out too small. */
int orc_scene_build(int nlevels, const amrvr_level* levels, int min_level,
                    int max_level, int component, int log_scale_input,
                    int normalize_to_data_range, amrvr_box* boxes_out,
                    int capacity, amrvr_scene_summary* summary) {
  memset(summary, 0, sizeof(*summary));
  double gmin[3] = {DBL_MAX, DBL_MAX, DBL_MAX};
  double gmax[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  double smin = INFINITY, smax = -INFINITY, spos = INFINITY;
  long long finite = 0;
  int n = 0;
  for (int level = min_level; level <= max_level; ++level) {
    if (level < 0 || level >= nlevels) continue;
    const amrvr_level* L = &levels[level];
    for (int f = 0; f < L->nfabs; ++f) {
      const amrvr_fab* F = &L->fabs[f];
      const int nx = F->hi[0] - F->lo[0] + 1, ny = F->hi[1] - F->lo[1] + 1,
                nz = F->hi[2] - F->lo[2] + 1;
      if (nx <= 0 || ny <= 0 || nz <= 0) continue;
      if (n >= capacity) return 5;
      amrvr_box* B = &boxes_out[n++];
      memset(B, 0, sizeof(*B));
      for (int d = 0; d < 3; ++d) { /* :166-184 */
        B->min_corner[d] = L->prob_lo[d] + (double)F->lo[d] * L->cell_size[d];
        B->max_corner[d] = L->prob_lo[d] + (double)(F->hi[d] + 1) * L->cell_size[d];
        B->lo[d] = F->lo[d];
        B->begin[d] = F->begin[d];
        if (B->min_corner[d] < gmin[d]) gmin[d] = B->min_corner[d];
        if (gmax[d] < B->max_corner[d]) gmax[d] = B->max_corner[d];
      }
      B->dims[0] = nx;
      B->dims[1] = ny;
      B->dims[2] = nz;
      B->data = F->data;
      B->jstride = F->jstride;
      B->kstride = F->kstride;
      B->nstride = F->nstride;
      B->component = component;
      for (int k = 0; k < nz; ++k) /* :53-94, non-finite cells skipped */
        for (int j = 0; j < ny; ++j)
          for (int i = 0; i < nx; ++i) {
            const double raw = fab_value(B, F->lo[0] + i, F->lo[1] + j, F->lo[2] + k);
            if (!isfinite(raw)) continue;
            if (raw < smin) smin = raw;
            if (smax < raw) smax = raw;
            if (raw > 0.0 && raw < spos) spos = raw;
            ++finite;
          }
    }
  }
  /* :223-257 rescale so that the shortest side of the global extent is 1 */
  double min_extent = DBL_MAX;
  for (int d = 0; d < 3; ++d) {
    const double len = fabs(gmax[d] - gmin[d]);
    if (len > 0.0 && isfinite(len) && len < min_extent) min_extent = len;
  }
  double scale = 1.0;
  if (min_extent > 0.0 && isfinite(min_extent)) scale = 1.0 / min_extent;
  if (!isfinite(scale) || !(scale > 0.0)) scale = 1.0;
  if (scale != 1.0) {
    for (int b = 0; b < n; ++b)
      for (int d = 0; d < 3; ++d) {
        boxes_out[b].min_corner[d] *= scale;
        boxes_out[b].max_corner[d] *= scale;
      }
  }
  /* :256-292 bounds of the scaled boxes */
  double bmin[3] = {DBL_MAX, DBL_MAX, DBL_MAX}, bmax[3] = {-DBL_MAX, -DBL_MAX, -DBL_MAX};
  for (int b = 0; b < n; ++b)
    for (int d = 0; d < 3; ++d) {
      if (boxes_out[b].min_corner[d] < bmin[d]) bmin[d] = boxes_out[b].min_corner[d];
      if (bmax[d] < boxes_out[b].max_corner[d]) bmax[d] = boxes_out[b].max_corner[d];
    }
  if (bmin[0] > bmax[0] || bmin[1] > bmax[1] || bmin[2] > bmax[2]) return 1;
  const double e0 = bmax[0] - bmin[0], e1 = bmax[1] - bmin[1], e2 = bmax[2] - bmin[2];
  const double e12 = (e1 < e2) ? e2 : e1;
  const double max_extent = (e0 < e12) ? e12 : e0;
  const double pad = (max_extent > 0.0) ? max_extent * 0.05 : 1.0;
  for (int d = 0; d < 3; ++d) {
    summary->bounds_min[d] = bmin[d] - pad;
    summary->bounds_max[d] = bmax[d] + pad;
  }
  summary->global_scale = scale;
  summary->nboxes = n;

  const double gsmin = (finite > 0) ? smin : INFINITY; /* :312-341 */
  const double gsmax = (finite > 0) ? smax : -INFINITY;
  if (finite > 0 && isfinite(gsmin) && isfinite(gsmax)) {
    make_range(gsmin, gsmax, summary->original_range);
    summary->has_original_range = 1;
  }
  double pmin = gsmin, pmax = gsmax;
  amrvr_scalar_transform* T = &summary->transform;
  T->log_scale_input = log_scale_input ? 1 : 0;
  T->normalize_to_unit_range = 0;
  T->positive_floor = 0.0;
  T->processed_max = 1.0; /* defaults of volume::ScalarTransform */
  T->inverse_processed_span = 1.0;
  T->normalization_max = 1.0;
  T->inverse_normalization_span = 1.0;
  if (log_scale_input) { /* :357-396 */
    const double gpos = (spos > 0.0 && isfinite(spos)) ? spos : INFINITY;
    if (!(gpos < INFINITY) || !isfinite(gpos) || !(gpos > 0.0)) return 3;
    T->positive_floor = gpos;
    pmin = log(gpos);
    pmax = log((gsmax < gpos) ? gpos : gsmax);
  }
  if (!isfinite(pmin) || !isfinite(pmax)) return 2;
  if (pmin == pmax) pmax = pmin + 1.0;
  make_range(pmin, pmax, summary->processed_range);
  T->processed_min = pmin;
  T->processed_max = pmax;
  T->inverse_processed_span = 1.0 / (pmax - pmin);
  T->normalization_min = pmin;
  T->normalization_max = pmax;
  T->inverse_normalization_span = T->inverse_processed_span;
  summary->scalar_range[0] = summary->processed_range[0];
  summary->scalar_range[1] = summary->processed_range[1];
  if (normalize_to_data_range) { /* :427-443 */
    const double span = pmax - pmin;
    if (!(span > 0.0) || !isfinite(span)) return 4;
    T->normalize_to_unit_range = 1;
    T->normalization_min = pmin;
    T->normalization_max = pmax;
    T->inverse_normalization_span = 1.0 / span;
    summary->scalar_range[0] = 0.0f;
    summary->scalar_range[1] = 1.0f;
  }
  return 0;
}

/* SceneBuilder.cpp:445-577 (per-cell body :503-539); counts of all boxes. */
void orc_scene_histogram(int nboxes, const amrvr_box* boxes,
                         const amrvr_scalar_transform* t, float range_min,
                         float range_max, int bin_count, uint64_t* counts,
                         uint64_t* sample_count) {
  for (int b = 0; b < bin_count; ++b) counts[b] = 0;
  if (sample_count) *sample_count = 0;
  const float width = range_max - range_min;
  if (!(width > 0.0f) || !isfinite(width)) return;
  const float inverse_width = 1.0f / width;
  uint64_t total = 0;
  for (int bx = 0; bx < nboxes; ++bx) {
    const amrvr_box* B = &boxes[bx];
    const int nx = B->dims[0], ny = B->dims[1], nz = B->dims[2];
    if (nx <= 0 || ny <= 0 || nz <= 0) continue;
    for (int k = 0; k < nz; ++k)
      for (int j = 0; j < ny; ++j)
        for (int i = 0; i < nx; ++i) {
          float value = apply_transform(fab_value(B, B->lo[0] + i, B->lo[1] + j, B->lo[2] + k), t);
          if (value < range_min) value = range_min;
          else if (value > range_max) value = range_max;
          float normalized = (value - range_min) * inverse_width;
          if (normalized < 0.0f) normalized = 0.0f;
          else if (normalized > 1.0f) normalized = 1.0f;
          int index = (int)(normalized * (float)bin_count);
          if (index >= bin_count) index = bin_count - 1;
          else if (index < 0) index = 0;
          ++counts[index];
          ++total;
        }
  }
  if (sample_count) *sample_count = total;
}
