// internal.hpp -- declarations shared by the host logic, the CUDA kernels and the
// C ABI of libamrvr_b200.so.  Not installed; include/amrvr.h is the public face.
#pragma once

#include <cstddef>
#include <cstdint>
#include <string>
#include <vector>

#include "../../include/amrvr.h"

namespace amrvr {

// ----------------------------------------------------------------- host logic
// (host_logic.cpp; pure CPU, callable without a GPU)

struct CameraBasis {
  // double-precision frame, then the float copies the march uses
  // (reference: Common/VolumePainter.cpp:631-656).
  double fwd_d[3], right_d[3], up_d[3];
  float eye[3], fwd[3], right[3], up[3];
};
void make_camera_basis(const amrvr_camera& cam, CameraBasis& out);

// Transfer-function table (reference: Common/VolumePainter.cpp:107-516).
void build_color_table(float alpha_scale, float norm_factor, float range_min,
                       float range_max, const amrvr_colormap_pt* pts, int npts,
                       float* rgba256);

// Step length and opacity-normalisation factor of one box
// (reference: Common/VolumePainter.cpp:571-610).
void box_sampling(const double* min_corner, const double* max_corner,
                  const int32_t* dims, float reference_sample_distance,
                  float* step, float* norm_factor);

// reference: VolumeRenderer/VolumeRenderer.cpp:541-553
float box_depth_hint(const amrvr_box_record& rec, const amrvr_camera& cam);
// reference: DirectSend/Base/DirectSendBase.cpp:363-388
void layer_order(int nglobal, const amrvr_box_record* recs,
                 const amrvr_camera& cam, float* hints, int* order);
// reference: DirectSend/Base/DirectSendBase.cpp:59-74
void piece_range(int num_pixels, int piece, int num_pieces, int* begin, int* end);
// reference: VolumeRenderer/VolumeRenderer.cpp:1138-1190
float reference_sample_distance(int nglobal, const amrvr_box_record* recs,
                                const double* bounds_min, const double* bounds_max);

// Box as the visibility graph sees it: float-truncated corners, owner and NDC
// depth interval (reference: Common/VisibilityOrdering.cpp:103-207).
struct OrderBox {
  float lo[3], hi[3];
  int32_t owner;
  float min_depth, max_depth;
};
struct OrderMatrices {
  float view[16];  // row-major 4x4
  float proj[16];
  float dir[3];    // float view direction
};
void make_order_matrices(const amrvr_camera& cam, float aspect, OrderMatrices& m);
void order_box_depth_range(const OrderMatrices& m, OrderBox& b);
void order_box_depth_ranges(const OrderMatrices& m, OrderBox* boxes, int n);  // same values, whole table

// One directed "in front of" relation found by the pair test, tagged with the
// moment the reference would have inserted it (pair (i,j), i<j, axis).
struct OrderEdge {
  int32_t i, j;     // the pair, i < j
  int32_t axis;     // 0..2
  int32_t forward;  // 1: edge i->j, 0: edge j->i
};
// Host pair test (used when no GPU context is given and as the definition the
// device kernel must match).  Appends edges in (i, j, axis) order.
void order_pair_edges_host(const std::vector<OrderBox>& boxes, const float* dir,
                           std::vector<OrderEdge>& edges);

// Callbacks let the same sequential driver run with device or host back ends.
struct OrderBackend {
  virtual ~OrderBackend() = default;
  // priority[i] = position of box i in the total order of
  // VisibilityOrdering.cpp:238-257 (compareBoxes).
  virtual void rank_boxes(const std::vector<OrderBox>& boxes,
                          std::vector<int32_t>& priority) = 0;
  virtual void pair_edges(const std::vector<OrderBox>& boxes, const float* dir,
                          std::vector<OrderEdge>& edges) = 0;
};
OrderBackend* host_order_backend();
// Full ordering (reference: Common/VisibilityOrdering.cpp:63-632).  Returns
// true when the graph could be ordered; rank_order is identity otherwise.
bool order_ranks(int nranks, int nglobal, const amrvr_box_record* recs,
                 const amrvr_camera& cam, float aspect, bool use_graph,
                 OrderBackend& backend, int* rank_order,
                 std::vector<std::string>* graph_dots = nullptr);
// graph_dots (optional): receives the DOT document of every round's graph, the
// text exportGraph writes to visibility_graph_<n>.dot (:319-352).

// Screen-space footprint of a box: a conservative pixel rectangle that
// contains every pixel whose centre ray can hit the box.  x0 and w are
// multiples of 4 (w may overhang the frame on the right).  Empty boxes give
// w = h = 0.
struct Rect {
  int32_t x0, y0, w, h;
};
Rect box_footprint(const double* min_corner, const double* max_corner,
                   const int32_t* dims, const CameraBasis& basis,
                   const amrvr_camera& cam, int width, int height);
void box_footprints(int n, const amrvr_box_record* recs, const CameraBasis& basis,
                    const amrvr_camera& cam, int width, int height, Rect* out);

// Wireframe overlay geometry (reference: VolumeRenderer.cpp:139-335): projected
// corners of the tight bounds; the per-pixel coverage runs in a kernel.
struct OverlayEdges {
  int32_t count;
  float ax[12], ay[12], bx[12], by[12];
  int32_t x0[12], x1[12], y0[12], y1[12];
  int32_t degenerate[12];
};
void make_overlay_edges(const double* tight_min, const double* tight_max,
                        const amrvr_camera& cam, int sqrt_aa, int width,
                        int height, OverlayEdges& out);

// ------------------------------------------------------------ device structs

struct QuantJob {  // one brick of a batched quantise launch
  const double* src;
  int64_t jstride, kstride;
  uint8_t* dst;
  int32_t nx, ny, nz, pad;
};

struct MarchBox {
  float lo[3], hi[3];
  float cell[3];       // dx, dy, dz (float, as the reference derives them)
  int32_t n[3];
  float step, eps;
  int32_t table;       // index of the 256-entry table this box uses
  int32_t rx0, ry0, rw, rh;
  int32_t tile_base;   // first tile id of this box in the launch
  int32_t tiles_x;
  int32_t exact_recip; // 1: cell sizes are powers of two -> multiply by inverse is exact
  float inv_cell_tiny[3];  // 2^-149 / cell (exact when cell is a power of two)
  float margin;        // > 0: shrink distance of the provably-inside stretch; 0: always check
  uint64_t brick_off;  // byte offset into the brick pool
  uint64_t layer_off;  // float offset into the fragment pool of destination `dest`
  int32_t dest;        // which pool the fragments go to (MarchQueues::pool index)
  int32_t ry_first;    // frame row of the item's first tile row (== ry0 unless the frame is striped)
};

// A frame striped over P GPUs (MarchFrame::strip_stride == 16 * P): GPU q folds the
// 16-row strips k = q, q + P, ... of the frame, so march queue q holds, of every box,
// the tiles of the strips k % P == q that its footprint rows [ry0, ry0 + rh) touch --
// every GPU gets an even share of every footprint.  Per (queue, box) only these three
// numbers differ; boxes[] then holds one MarchBox per box (its tile_base / layer_off /
// dest / ry_first are filled in shared memory from queue_items[q * nboxes + box]).
// A layer of queue q stores its strips one after the other, 16 rows each.
struct QueueItem {
  int32_t tile_base;   // first tile id of (queue, box); equal to the next one's when it has no tile
  int32_t ry_first;    // first frame row of the first strip: 16 * (first k >= ry0 / 16 with k % P == q)
  uint64_t layer_off;  // float offset into pool[q]
};

#if defined(__CUDACC__)
#define AMRVR_HD __host__ __device__
#else
#define AMRVR_HD
#endif

// ---- striped frame arithmetic, shared by the host plan, the fold kernel and the CPU tests.
// Frame rows come in 16-row strips; strip k is folded by GPU k % P.
// First strip >= the one holding `row0` that GPU g folds.
AMRVR_HD inline int stripe_first_strip(int p, int g, int row0) {
  const int k0 = row0 >> 4;
  return k0 + (((g - k0) % p) + p) % p;
}
// Number of strips of GPU g that touch rows [row0, row1).
AMRVR_HD inline int stripe_count(int p, int g, int row0, int row1) {
  if (row1 <= row0) return 0;
  const int kf = stripe_first_strip(p, g, row0), k1 = (row1 - 1) >> 4;
  return (kf > k1) ? 0 : (k1 - kf) / p + 1;
}
// Row of GPU g's layer (its strips of a footprint starting at row0, one after the other,
// 16 rows each) that holds frame row y; y must lie in one of g's strips.
AMRVR_HD inline long long stripe_layer_row(int p, int g, int row0, int y) {
  return static_cast<long long>(((y >> 4) - stripe_first_strip(p, g, row0)) / p) * 16 + (y & 15);
}
// Plan of a striped frame: for every (queue q, box b) the first tile id, the first frame
// row of the queue's first strip of the box and the float offset of the box's strips in
// pool q; tile_begin[q] / pool_floats[q] per queue.  Boxes: footprint rows [ry0, ry0 + rh),
// padded width rw.  qitems: [p * nboxes].
void stripe_plan(int p, int nboxes, const int32_t* ry0, const int32_t* rh, const int32_t* rw,
                 QueueItem* qitems, int32_t* tile_begin /*[p + 1]*/, uint64_t* pool_floats /*[p]*/);

struct MarchFrame {
  float eye[3], fwd[3], right[3], up[3];
  float tan_half, aspect, inv_w, inv_h;
  int32_t width, height;
  int32_t nboxes, ntiles;
  int32_t strip_stride;  // rows from one tile row of an item to the next: 16, or 16 * P when the frame is striped
  int32_t pad;
};

// Peer-memory exchange (one box, NVLink/NVSwitch): pointers to the same buffer on
// every rank of the box, mapped into this process (CUDA IPC, or plain peer
// access when the ranks are threads of one process).
constexpr int kMaxPeers = 8;
constexpr int kFlagRows = 4;  // flag rows on every rank: [kFlagRows][kMaxPeers] u32
// Layout of the peer-mapped control block of a rank (u32 words): flag rows, then
// the barrier time-out word, the tile counter of the rank's march queue (its own
// 128 B line: peers hit it with atomics), the "queue q is drained" hints and the
// count of finished CTAs of the running march / fold launch.
constexpr int kCtlError = kFlagRows * kMaxPeers;  // word 32
constexpr int kCtlTileCounter = 64;
constexpr int kCtlHints = 96;                     // [kMaxPeers]
constexpr int kCtlDoneMarch = 128;
constexpr int kCtlDoneFold = 160;
constexpr int kCtlWords = 192;
struct PeerPtrs {
  void* p[kMaxPeers];
};

// The tiles of one march launch, as up to kMaxPeers queues.  A single-GPU march
// has one queue.  A frame sharded over the GPUs of a box has one queue per
// GPU: queue q holds every (box, tile) of the frame strips GPU q folds, its
// fragments go to pool[q] and its tile counter lives on that GPU.  A GPU drains `mine` first and then helps with the other
// queues in `steal_order`, writing the fragments of those tiles across NVLink.
struct MarchQueues {
  int32_t nqueues, mine;
  int32_t me, nranks;              // nranks > 0: the last CTA raises flag row `signal_row` on every rank
  int32_t tile_begin[kMaxPeers + 1];
  int32_t steal_order[kMaxPeers];
  const QueueItem* queue_items;    // striped frame: [nqueues][nboxes]; nullptr: boxes[] are the items of one queue
  unsigned* counter[kMaxPeers];    // next tile of queue q (relative to tile_begin[q])
  float* pool[kMaxPeers];
  unsigned* ctl[kMaxPeers];        // control block of rank r (hints, flags); [me] is the local one
  unsigned long long* stats;       // [0] samples  [1] bytes written to other GPUs  [2] tiles taken from other queues
  uint32_t epoch;
  int32_t signal_row;
};

// One layer as the fold kernel sees it, already in global blend order.
struct FoldLayer {
  const float* base;   // row r of the available slice at base + r*w*5
  int32_t x0, w;       // footprint rectangle origin (x) and padded width
  int32_t row0, row1;  // rows [row0,row1) of the frame held at `base`
  int32_t run;         // id of the same-owner run this layer belongs to
  int32_t interleaved; // 1: pixel (x,y) at base + ((y-row0)*w + x-x0)*5 (the reference's layout);
                       //    base may then point before the slice (only in-range pixels are read)
};

// A byte range to copy when packing slices for the exchange.
struct CopyJob {
  uint64_t src_off, dst_off;  // in floats
  uint32_t count;             // floats, multiple of 4
  uint32_t pad;               // push_slices_kernel: destination rank
};

// ------------------------------------------------------------ kernel launchers
// (kernels.cu; every launcher returns the cudaError_t of the launch as int and
// bumps *launches by the number of kernels it started)

int launch_brick_stats(const double* data, int64_t jstride, int64_t kstride,
                       int nx, int ny, int nz, double* partials /*[4*blocks]*/,
                       int blocks, void* stream, int64_t* launches);
int launch_brick_stats_final(const double* partials, int nparts, double* out4,
                             void* stream, int64_t* launches);
int launch_brick_histogram(const double* data, int64_t jstride, int64_t kstride,
                           int nx, int ny, int nz, const amrvr_scalar_transform& t,
                           float range_min, float range_max, int bin_count,
                           unsigned long long* counts_dev, void* stream, int64_t* launches);
int launch_brick_quantise_batch(const QuantJob* jobs_dev, int njobs, long long max_cells,
                                const amrvr_scalar_transform& t, float range_min, float range_max,
                                void* stream, int64_t* launches);
int launch_raymarch(const MarchFrame& frame, const MarchQueues& queues, const MarchBox* boxes_dev,
                    const uint8_t* bricks, const float* tables, bool count_samples, void* stream,
                    int64_t* launches);
// Rows a rank folds in a frame striped over `p` GPUs: the 16-row strips g, g + p, ... of a
// frame `height` rows high (see QueueItem).  nullptr / p <= 1: the pixel range the caller gives.
struct FoldStrips {
  int p, g, height;
};
// scratch: fold_scratch_bytes() of device memory (per-band layer lists)
size_t fold_scratch_bytes(int nlayers, int width, int pixel_begin, int pixel_end,
                          const FoldStrips* strips = nullptr);
// Byte epilogue of the fold: overlay edges in device memory (nullptr: no
// overlay), frame height, RGB8 frame (top row first; may be the root's frame on
// another GPU -- every rank writes the bytes of its own piece).
struct FoldBytes {
  const OverlayEdges* edges;
  int height;
  uint8_t* rgb8;
};
// Cross-GPU hand-shake fused into the fold of a sharded frame: before the first
// fragment is read every CTA waits until flag row `wait_row` of the own control
// block holds `epoch` from all nranks ranks (= every march that may have written
// into this GPU's pool has finished); the march queue counter is then reset for
// the next frame; the last CTA to finish raises row `signal_row` on every rank.
struct FoldSync {
  PeerPtrs ctl;
  int32_t me, nranks;
  int32_t wait_row, signal_row;
  uint32_t epoch;
  unsigned long long timeout_ns;
};
// Pass 1 (fold_bin): per 8-row band the ordered list of layers touching it.  It
// depends on the frame plan only, never on pixel data, so a frame launches it
// BEFORE the march and the march hides it.
int launch_fold_bin(const FoldLayer* layers_dev, int nlayers, int width, int pixel_begin,
                    int pixel_end, void* scratch, void* stream, int64_t* launches,
                    const FoldStrips* strips = nullptr);
// Pass 2: the fold proper over the lists pass 1 left in `scratch`.
int launch_fold(const FoldLayer* layers_dev, int nlayers, int width,
                int pixel_begin, int pixel_end, float* piece5, void* scratch,
                void* stream, int64_t* launches,
                const FoldBytes* bytes = nullptr, const FoldSync* sync = nullptr,
                const FoldStrips* strips = nullptr);
// Wait (on the stream) until flag row `row` of the own control block holds
// `epoch` from every rank; raises the time-out word instead of hanging.
int launch_peer_wait(const PeerPtrs& ctl, int me, int nranks, int row, uint32_t epoch,
                     unsigned long long timeout_ns, void* stream, int64_t* launches);
// NVLink probe: copy `bytes` between two device buffers (either may be a peer
// mapping) with plain 16 B loads and stores from `blocks` CTAs.
int launch_peer_copy(const void* src, void* dst, size_t bytes, int blocks, void* stream,
                     int64_t* launches);
int launch_pack(const CopyJob* jobs_dev, int njobs, const float* src, float* dst,
                void* stream, int64_t* launches);
// Cross-GPU barrier on the stream: store `epoch` into slot `me` of every peer's
// flag row `which`, then wait until all nranks slots of the own row hold >= epoch.
// Gives up after timeout_ns and raises *error_flag (own device memory) instead of
// hanging the GPU.
int launch_push(const CopyJob* jobs_dev, int njobs, const float* src, const PeerPtrs& recv, void* stream,
                int64_t* launches);
int launch_peer_barrier(const PeerPtrs& flags, int me, int nranks, int which, uint32_t epoch,
                        unsigned long long timeout_ns, uint32_t* error_flag, void* stream,
                        int64_t* launches);
// Overlay + byte quantisation of one composited piece, written straight into the
// root's RGB8 frame (a peer pointer when the root is another GPU).
int launch_post_piece(const float* piece5, int pixel_begin, int pixel_end, int width, int height,
                      const OverlayEdges* edges /*nullptr: no overlay*/, uint8_t* rgb8_root,
                      void* stream, int64_t* launches);
int launch_downsample(const float* src5, int width, int height, int sqrt_aa,
                      float* dst5, void* stream, int64_t* launches);
int launch_overlay(const OverlayEdges& edges, int sqrt_aa, int width, int height,
                   float* image5, void* stream, int64_t* launches);
int launch_quantize_rgb8(const float* image5, int width, int height,
                         uint8_t* rgb8, void* stream, int64_t* launches);
int launch_order_rank(const OrderBox* boxes_dev, int n, void* scratch_a /*16 B per box*/,
                      void* scratch_b, int owner_bits, int32_t* priority_dev,
                      void* stream, int64_t* launches);
// priority[i] = number of boxes that sort before box i (all-pairs count over the
// same keys; any n, many CTAs, no passes).  priority_dev must be zeroed.
int launch_order_count(const OrderBox* boxes_dev, int n, int32_t* priority_dev, void* stream,
                       int64_t* launches);
int launch_order_pairs(const OrderBox* boxes_dev, int n, float dx, float dy,
                       float dz, OrderEdge* edges_dev, int capacity,
                       int32_t* count_dev, void* stream, int64_t* launches);

// Records msg as the calling thread's last error (amrvr_last_error(NULL)) and
// returns status; for the host-only entry points that have no ctx (amr_io.cpp).
int report_error(int status, const std::string& msg);

}  // namespace amrvr
