/* oracle/restate/orc_order.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * Restates BuildVisibilityOrderedGroup (Common/VisibilityOrdering.cpp:63-632),
 * makeViewMatrix (Common/CameraUtils.hpp:24-61), computeBoxDepthHint
 * (VolumeRenderer/VolumeRenderer.cpp:541-553), the layer order of
 * DirectSendBase.cpp:363-388 and referenceSampleDistance
 * (VolumeRenderer.cpp:1138-1190). */
#include "oracle.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float maxf(float a, float b) { return (a < b) ? b : a; }
static float minf(float a, float b) { return (b < a) ? b : a; }

/* CameraUtils.hpp:16-22 */
static void safe_normalize(const double* in, double* out) {
  const double len = sqrt(in[0] * in[0] + in[1] * in[1] + in[2] * in[2]);
  if (len > 0.0 && isfinite(len)) {
    out[0] = in[0] / len;
    out[1] = in[1] / len;
    out[2] = in[2] / len;
  } else {
    out[0] = 0.0;
    out[1] = 0.0;
    out[2] = -1.0;
  }
}

static void view_dir(const amrvr_camera* c, double* out) {
  const double d[3] = {c->look_at[0] - c->eye[0], c->look_at[1] - c->eye[1],
                       c->look_at[2] - c->eye[2]};
  safe_normalize(d, out);
}

/* m[r][c]; 4x4 float */
typedef struct {
  float m[4][4];
} mat4;

static mat4 identity4(void) {
  mat4 a;
  memset(&a, 0, sizeof(a));
  for (int i = 0; i < 4; ++i) a.m[i][i] = 1.0f;
  return a;
}

/* CameraUtils.hpp:24-61 makeViewMatrix, element for element (translation
 * lands in row 3, exactly as the reference writes it). */
static mat4 view_matrix(const amrvr_camera* c) {
  double f[3], r[3], u[3];
  view_dir(c, f);
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
  const double* e = c->eye;
  mat4 v = identity4();
  v.m[0][0] = (float)r[0];
  v.m[1][0] = (float)r[1];
  v.m[2][0] = (float)r[2];
  v.m[3][0] = (float)(-(r[0] * e[0] + r[1] * e[1] + r[2] * e[2]));
  v.m[0][1] = (float)u[0];
  v.m[1][1] = (float)u[1];
  v.m[2][1] = (float)u[2];
  v.m[3][1] = (float)(-(u[0] * e[0] + u[1] * e[1] + u[2] * e[2]));
  v.m[0][2] = (float)(-f[0]);
  v.m[1][2] = (float)(-f[1]);
  v.m[2][2] = (float)(-f[2]);
  v.m[3][2] = (float)(f[0] * e[0] + f[1] * e[1] + f[2] * e[2]);
  v.m[0][3] = 0.0f;
  v.m[1][3] = 0.0f;
  v.m[2][3] = 0.0f;
  v.m[3][3] = 1.0f;
  return v;
}

/* VisibilityOrdering.cpp:34-59 makePerspectiveMatrix */
static mat4 perspective_matrix(float fov_y, float aspect, float near_plane,
                               float far_plane) {
  const float pi = 3.14159265358979323846f;
  mat4 p = identity4();
  const float tangent = tanf(fov_y * pi / 180.0f * 0.5f);
  const float size = near_plane * tangent;
  const float left = -size * aspect, right = size * aspect;
  const float bottom = -size, top = size;
  p.m[0][0] = 2.0f * near_plane / (right - left);
  p.m[1][1] = 2.0f * near_plane / (top - bottom);
  p.m[0][2] = (right + left) / (right - left);
  p.m[1][2] = (top + bottom) / (top - bottom);
  p.m[2][2] = -(far_plane + near_plane) / (far_plane - near_plane);
  p.m[3][2] = -1.0f;
  p.m[2][3] = -(2.0f * far_plane * near_plane) / (far_plane - near_plane);
  p.m[3][3] = 0.0f;
  return p;
}

/* SmallMatrix<float,4,4> * SmallMatrix<float,4,1>: k ascending from 0 */
static void mat_vec(const mat4* a, const float* v, float* out) {
  for (int i = 0; i < 4; ++i) {
    float acc = 0.0f;
    for (int k = 0; k < 4; ++k) acc += a->m[i][k] * v[k];
    out[i] = acc;
  }
}

typedef struct {
  float lo[3], hi[3]; /* float-truncated corners (:114-119) */
  int owner;
  float min_depth, max_depth;
} box_info;

/* VisibilityOrdering.cpp:164-190 computeDepthRange */
static void depth_range(const mat4* view, const mat4* proj, box_info* b) {
  float lo = INFINITY, hi = -INFINITY;
  for (int corner = 0; corner < 8; ++corner) {
    const float h[4] = {(corner & 1) ? b->hi[0] : b->lo[0],
                        (corner & 2) ? b->hi[1] : b->lo[1],
                        (corner & 4) ? b->hi[2] : b->lo[2], 1.0f};
    float vs[4], clip[4];
    mat_vec(view, h, vs);
    mat_vec(proj, vs, clip);
    if (clip[3] != 0.0f) {
      const float d = clip[2] / clip[3];
      lo = minf(lo, d);
      hi = maxf(hi, d);
    }
  }
  if (!isfinite(lo) || !isfinite(hi)) {
    lo = INFINITY;
    hi = INFINITY;
  }
  b->min_depth = lo;
  b->max_depth = hi;
}

/* :211-214 */
static int nearly_equal(float a, float b) {
  const float scale = maxf(maxf(1.0f, fabsf(a)), fabsf(b));
  return fabsf(a - b) <= 1e-5f * scale;
}

/* :216-228 */
static int overlaps(float a_min, float a_max, float b_min, float b_max) {
  const float o_min = maxf(a_min, b_min);
  const float o_max = minf(a_max, b_max);
  float scale = 1.0f;
  scale = maxf(scale, fabsf(a_min));
  scale = maxf(scale, fabsf(a_max));
  scale = maxf(scale, fabsf(b_min));
  scale = maxf(scale, fabsf(b_max));
  scale = maxf(scale, fabsf(o_min));
  scale = maxf(scale, fabsf(o_max));
  return (o_max - o_min) > 1e-5f * scale;
}

/* :238-257 compareBoxes: 1 when lhs sorts before rhs */
static int box_before(const box_info* boxes, int lhs, int rhs) {
  const box_info* a = &boxes[lhs];
  const box_info* b = &boxes[rhs];
  const int af = isfinite(a->min_depth), bf = isfinite(b->min_depth);
  if (af != bf) return af && !bf;
  if (a->min_depth == b->min_depth) {
    if (a->max_depth == b->max_depth) {
      if (a->owner == b->owner) return lhs < rhs;
      return a->owner < b->owner;
    }
    return a->max_depth < b->max_depth;
  }
  return a->min_depth < b->min_depth;
}

typedef struct {
  int* v;
  int n, cap;
} ivec;

static void ivec_push(ivec* a, int x) {
  if (a->n == a->cap) {
    a->cap = a->cap ? a->cap * 2 : 8;
    a->v = (int*)realloc(a->v, sizeof(int) * (size_t)a->cap);
  }
  a->v[a->n++] = x;
}

static void add_edge(ivec* adj, int* indegree, int from, int to) { /* :268-277 */
  if (from == to) return;
  for (int i = 0; i < adj[from].n; ++i)
    if (adj[from].v[i] == to) return;
  ivec_push(&adj[from], to);
  ++indegree[to];
}

/* :261-317 rebuildAdjacency */
static void build_adjacency(const box_info* boxes, int n, const float* dir,
                            ivec* adj, int* indegree) {
  const float tol = 1e-6f;
  for (int i = 0; i < n; ++i) {
    adj[i].n = 0;
    indegree[i] = 0;
  }
  for (int i = 0; i < n; ++i) {
    const box_info* a = &boxes[i];
    for (int j = i + 1; j < n; ++j) {
      const box_info* b = &boxes[j];
      for (int axis = 0; axis < 3; ++axis) {
        const int a1 = (axis + 1) % 3, a2 = (axis + 2) % 3;
        if (!overlaps(a->lo[a1], a->hi[a1], b->lo[a1], b->hi[a1]) ||
            !overlaps(a->lo[a2], a->hi[a2], b->lo[a2], b->hi[a2]))
          continue;
        const float dc = dir[axis];
        if (nearly_equal(a->hi[axis], b->lo[axis])) {
          if (dc > tol) {
            add_edge(adj, indegree, j, i);
          } else if (dc < -tol) {
            add_edge(adj, indegree, i, j);
          }
        } else if (nearly_equal(b->hi[axis], a->lo[axis])) {
          if (dc > tol) {
            add_edge(adj, indegree, i, j);
          } else if (dc < -tol) {
            add_edge(adj, indegree, j, i);
          }
        }
      }
    }
  }
}

/* :360-401 topoSortBoxes.  The reference re-sorts the ready list with
 * compareBoxes after every pop and takes the front: that is "take the minimum
 * under compareBoxes", done here by a linear scan. */
static int topo_sort(const box_info* boxes, int n, const ivec* adj,
                     const int* indegree0, int* order, int* residual) {
  memcpy(residual, indegree0, sizeof(int) * (size_t)n);
  char* ready = (char*)calloc((size_t)n, 1);
  int count = 0;
  for (int i = 0; i < n; ++i)
    if (residual[i] == 0) ready[i] = 1;
  for (;;) {
    int best = -1;
    for (int i = 0; i < n; ++i)
      if (ready[i] && (best < 0 || box_before(boxes, i, best))) best = i;
    if (best < 0) break;
    ready[best] = 0;
    order[count++] = best;
    for (int e = 0; e < adj[best].n; ++e) {
      const int next = adj[best].v[e];
      if (--residual[next] == 0) ready[next] = 1;
    }
  }
  free(ready);
  return count == n;
}

/* :403-443 findCycle (recursive DFS in adjacency order) */
typedef struct {
  const ivec* adj;
  int* state;
  int* parent;
  ivec cycle;
} dfs_ctx;

static int dfs_visit(dfs_ctx* c, int node) {
  c->state[node] = 1;
  for (int e = 0; e < c->adj[node].n; ++e) {
    const int next = c->adj[node].v[e];
    if (c->state[next] == 0) {
      c->parent[next] = node;
      if (dfs_visit(c, next)) return 1;
    } else if (c->state[next] == 1) {
      c->cycle.n = 0;
      ivec_push(&c->cycle, next);
      for (int cur = node; cur != next && cur != -1; cur = c->parent[cur])
        ivec_push(&c->cycle, cur);
      for (int a = 0, b = c->cycle.n - 1; a < b; ++a, --b) {
        const int tmp = c->cycle.v[a];
        c->cycle.v[a] = c->cycle.v[b];
        c->cycle.v[b] = tmp;
      }
      return 1;
    }
  }
  c->state[node] = 2;
  return 0;
}

/* :445-563 breakCycle; returns 1 and appends the far half to boxes[*n] */
static int break_cycle(const ivec* cycle, box_info* boxes, int* n,
                       const float* dir, const mat4* view, const mat4* proj) {
  const float tol = 1e-6f;
  if (cycle->n < 2) return 0;
  int axis = 0;
  float best = fabsf(dir[0]);
  for (int a = 1; a < 3; ++a) {
    const float al = fabsf(dir[a]);
    if (al > best) {
      best = al;
      axis = a;
    }
  }
  if (best <= tol) {
    float widest = -1.0f;
    for (int a = 0; a < 3; ++a)
      for (int i = 0; i < cycle->n; ++i) {
        const box_info* b = &boxes[cycle->v[i]];
        const float len = b->hi[a] - b->lo[a];
        if (len > widest) {
          widest = len;
          axis = a;
        }
      }
  }
  const float dc = dir[axis];
  if (fabsf(dc) <= tol) return 0;
  const float min_len_tol = 1e-6f;
  int target = cycle->v[0];
  float target_len = -1.0f;
  for (int i = 0; i < cycle->n; ++i) {
    const box_info* b = &boxes[cycle->v[i]];
    const float len = b->hi[axis] - b->lo[axis];
    if (len > target_len && len > min_len_tol) {
      target_len = len;
      target = cycle->v[i];
    }
  }
  if (target_len <= min_len_tol) return 0;
  const box_info tb = boxes[target];
  const float lo = tb.lo[axis], hi = tb.hi[axis];
  const float length = hi - lo;
  const float eps = maxf(1e-5f * length, 1e-6f);
  int have = 0;
  float cand_max = 0.0f, cand_min = 0.0f;
  for (int i = 0; i < cycle->n; ++i) {
    if (cycle->v[i] == target) continue;
    const box_info* o = &boxes[cycle->v[i]];
    const float c[2] = {o->lo[axis], o->hi[axis]};
    for (int s = 0; s < 2; ++s) {
      if (c[s] > lo + eps && c[s] < hi - eps) {
        if (!have) {
          cand_max = cand_min = c[s];
          have = 1;
        } else {
          if (cand_max < c[s]) cand_max = c[s]; /* std::max_element: first max */
          if (c[s] < cand_min) cand_min = c[s]; /* std::min_element: first min */
        }
      }
    }
  }
  float split = 0.5f * (lo + hi);
  if (have) split = (dc > 0.0f) ? cand_max : cand_min;
  if (split <= lo + eps) split = lo + eps;
  if (split >= hi - eps) split = hi - eps;
  if (!(split > lo && split < hi)) return 0;
  box_info near_box = tb, far_box = tb;
  if (dc > 0.0f) {
    near_box.hi[axis] = split;
    far_box.lo[axis] = split;
  } else {
    near_box.lo[axis] = split;
    far_box.hi[axis] = split;
  }
  depth_range(view, proj, &near_box);
  depth_range(view, proj, &far_box);
  boxes[target] = near_box;
  boxes[(*n)++] = far_box;
  return 1;
}

int orc_order(int nranks, int nglobal, const amrvr_box_record* recs,
              const amrvr_camera* camera, float aspect, int use_graph,
              int* rank_order) {
  for (int i = 0; i < nranks; ++i) rank_order[i] = i; /* defaultOrder */
  if (!use_graph) return 1;
  if (nglobal <= 0) return 1; /* :98-100 */

  const mat4 view = view_matrix(camera);
  const mat4 proj = perspective_matrix(camera->fov_y_degrees, aspect,
                                       camera->near_plane, camera->far_plane);
  double vd[3];
  view_dir(camera, vd);
  const float dir[3] = {(float)vd[0], (float)vd[1], (float)vd[2]};

  const int max_iter = ((nglobal > 1) ? nglobal : 1) * 8 + 32; /* :565-566 */
  const int cap = nglobal + max_iter + 1;
  box_info* boxes = (box_info*)calloc((size_t)cap, sizeof(box_info));
  ivec* adj = (ivec*)calloc((size_t)cap, sizeof(ivec));
  int* indegree = (int*)calloc((size_t)cap, sizeof(int));
  int* residual = (int*)calloc((size_t)cap, sizeof(int));
  int* order = (int*)calloc((size_t)cap, sizeof(int));
  int* state = (int*)calloc((size_t)cap, sizeof(int));
  int* parent = (int*)calloc((size_t)cap, sizeof(int));
  int n = nglobal;
  for (int i = 0; i < n; ++i) {
    for (int d = 0; d < 3; ++d) {
      boxes[i].lo[d] = (float)recs[i].min_corner[d];
      boxes[i].hi[d] = (float)recs[i].max_corner[d];
    }
    boxes[i].owner = recs[i].owner;
    depth_range(&view, &proj, &boxes[i]);
  }

  int success = 0;
  dfs_ctx dfs;
  memset(&dfs, 0, sizeof(dfs));
  for (int iter = 0; iter < max_iter; ++iter) {
    build_adjacency(boxes, n, dir, adj, indegree);
    if (topo_sort(boxes, n, adj, indegree, order, residual)) {
      char* seen = (char*)calloc((size_t)nranks, 1);
      int m = 0;
      for (int i = 0; i < n; ++i) { /* :576-593 */
        const int owner = boxes[order[i]].owner;
        if (owner >= 0 && owner < nranks && !seen[owner]) {
          seen[owner] = 1;
          rank_order[m++] = owner;
        }
      }
      for (int r = 0; r < nranks; ++r)
        if (!seen[r]) rank_order[m++] = r;
      free(seen);
      success = 1;
      break;
    }
    for (int i = 0; i < n; ++i) {
      state[i] = 0;
      parent[i] = -1;
    }
    dfs.adj = adj;
    dfs.state = state;
    dfs.parent = parent;
    dfs.cycle.n = 0;
    for (int node = 0; node < n; ++node)
      if (residual[node] > 0 && state[node] == 0)
        if (dfs_visit(&dfs, node)) break;
    if (dfs.cycle.n == 0) break;
    if (!break_cycle(&dfs.cycle, boxes, &n, dir, &view, &proj)) break;
  }
  for (int i = 0; i < cap; ++i) free(adj[i].v);
  free(dfs.cycle.v);
  free(adj);
  free(boxes);
  free(indegree);
  free(residual);
  free(order);
  free(state);
  free(parent);
  if (!success)
    for (int i = 0; i < nranks; ++i) rank_order[i] = i; /* :615-623 fallback */
  return success;
}

/* VolumeRenderer.cpp:541-553 computeBoxDepthHint */
float orc_depth_hint(const amrvr_box_record* rec, const amrvr_camera* camera) {
  double vd[3];
  view_dir(camera, vd);
  float best = INFINITY;
  for (int corner = 0; corner < 8; ++corner) {
    const double c[3] = {
        (corner & 1) ? rec->max_corner[0] : rec->min_corner[0],
        (corner & 2) ? rec->max_corner[1] : rec->min_corner[1],
        (corner & 4) ? rec->max_corner[2] : rec->min_corner[2]};
    const double d0 = c[0] - camera->eye[0];
    const double d1 = c[1] - camera->eye[1];
    const double d2 = c[2] - camera->eye[2];
    const float depth = (float)(d0 * vd[0] + d1 * vd[1] + d2 * vd[2]);
    best = minf(best, depth);
  }
  return best;
}

typedef struct {
  float depth;
  int owner, local, global;
} layer_entry;

static int layer_cmp(const void* pa, const void* pb) { /* DirectSendBase.cpp:380-388 */
  const layer_entry* a = (const layer_entry*)pa;
  const layer_entry* b = (const layer_entry*)pb;
  if (a->depth == b->depth) {
    if (a->owner == b->owner) return (a->local > b->local) - (a->local < b->local);
    return (a->owner > b->owner) - (a->owner < b->owner);
  }
  return (a->depth < b->depth) ? -1 : 1;
}

void orc_layer_order(int nglobal, const amrvr_box_record* recs,
                     const amrvr_camera* camera, float* hints, int* order) {
  layer_entry* e = (layer_entry*)malloc(sizeof(layer_entry) * (size_t)(nglobal + 1));
  for (int i = 0; i < nglobal; ++i) {
    e[i].depth = orc_depth_hint(&recs[i], camera);
    e[i].owner = recs[i].owner;
    e[i].local = recs[i].local_index;
    e[i].global = i;
    if (hints) hints[i] = e[i].depth;
  }
  qsort(e, (size_t)nglobal, sizeof(layer_entry), layer_cmp);
  for (int i = 0; i < nglobal; ++i) order[i] = e[i].global;
  free(e);
}

/* VolumeRenderer.cpp:1138-1190 */
float orc_reference_sample_distance(int nglobal, const amrvr_box_record* recs,
                                    const double* bounds_min,
                                    const double* bounds_max) {
  float coarsest = 0.0f;
  for (int i = 0; i < nglobal; ++i) {
    double spacing[3] = {0.0, 0.0, 0.0};
    for (int d = 0; d < 3; ++d)
      if (recs[i].dims[d] > 0)
        spacing[d] = (recs[i].max_corner[d] - recs[i].min_corner[d]) /
                     (float)recs[i].dims[d];
    float min_spacing = FLT_MAX;
    for (int d = 0; d < 3; ++d)
      if (spacing[d] > 0.0f && spacing[d] < min_spacing && isfinite(spacing[d]))
        min_spacing = (float)spacing[d];
    if (min_spacing > 0.0f && isfinite(min_spacing))
      coarsest = maxf(coarsest, min_spacing);
  }
  if (!(coarsest > 0.0f && isfinite(coarsest))) {
    float fallback = FLT_MAX;
    for (int d = 0; d < 3; ++d) {
      const float len = (float)(bounds_max[d] - bounds_min[d]);
      if (len > 0.0f && isfinite(len)) fallback = minf(fallback, len);
    }
    if (!(fallback > 0.0f && isfinite(fallback))) fallback = 1.0f;
    coarsest = maxf(1e-4f, fallback * 0.01f);
  }
  return maxf(coarsest * 0.5f, 1e-5f);
}
