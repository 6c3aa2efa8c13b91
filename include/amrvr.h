/* amrvr.h -- C ABI of libamrvr_b200.so, the B200-native (sm_100a) replacement for
 * amrVolumeRenderer's per-ray hot path and sort-last DirectSend composite.
 *
 * The reference has no FFI layer; its seams are C++ classes.  Each entry point
 * below names the reference interface it replaces (paths relative to the
 * reference tree).  amrvolumerenderer_b200/adapters/ holds thin C++ classes with
 * the reference's own names and signatures that forward here.
 *
 * Conventions: plain pointers and sizes only; every function returns an int
 * status (AMRVR_OK = 0); no exception crosses the ABI; amrvr_last_error() gives
 * the message the C++ adapters rethrow (std::invalid_argument for
 * AMRVR_ERR_INVALID, std::runtime_error otherwise).  One amrvr_ctx per GPU, used
 * from one thread at a time (the reference path is single-threaded per rank).
 * There is NO CPU fallback: every compute entry point fails with
 * AMRVR_ERR_CUDA when no sm_100 device is usable.  Entry points marked
 * [host-only] never touch the GPU and work on a machine without one.
 */
#ifndef AMRVR_H
#define AMRVR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AMRVR_ABI_VERSION 2

enum {
  AMRVR_OK = 0,
  AMRVR_ERR_INVALID = 1, /* bad argument (reference: std::invalid_argument)  */
  AMRVR_ERR_RUNTIME = 2, /* data problem (reference: std::runtime_error)     */
  AMRVR_ERR_CUDA = 3,    /* CUDA / device failure, or no usable GPU          */
  AMRVR_ERR_NCCL = 4     /* NCCL failure or NCCL not loadable                */
};

#define AMRVR_COLOR_TABLE_SIZE 256 /* Common/VolumePainter.cpp:35 */
#define AMRVR_NCCL_ID_BYTES 128

typedef struct amrvr_ctx amrvr_ctx;
typedef struct amrvr_scene amrvr_scene;
typedef struct amrvr_layers amrvr_layers;

/* volume::AmrBox (Common/VolumeTypes.hpp:69-76) flattened together with the
 * amrex::Array4<Real const> view it carries.  `data` may be a host pointer
 * (pageable or pinned) or a device pointer; element (i,j,k,n) lives at
 *   data[(i-begin[0]) + (j-begin[1])*jstride + (k-begin[2])*kstride + n*nstride]
 * exactly as amrex::Array4::operator().  Real is double (reference default). */
typedef struct amrvr_box {
  double min_corner[3];
  double max_corner[3];
  int32_t dims[3];  /* cellDimensions */
  int32_t lo[3];    /* validBox.smallEnd() */
  const double* data;
  int64_t jstride, kstride, nstride;
  int32_t begin[3]; /* Array4::begin */
  int32_t component;
} amrvr_box;

/* volume::ScalarTransform (Common/VolumeTypes.hpp:21-31). */
typedef struct amrvr_scalar_transform {
  int32_t log_scale_input;
  int32_t normalize_to_unit_range;
  double positive_floor;
  double processed_min;
  double processed_max;
  double inverse_processed_span;
  double normalization_min;
  double normalization_max;
  double inverse_normalization_span;
} amrvr_scalar_transform;

/* volume::CameraParameters (Common/VolumeTypes.hpp:83-90). */
typedef struct amrvr_camera {
  double eye[3];
  double look_at[3];
  double up[3];
  float fov_y_degrees;
  float near_plane;
  float far_plane;
} amrvr_camera;

/* volume::ColorMapControlPoint (Common/VolumeTypes.hpp:92-98). */
typedef struct amrvr_colormap_pt {
  float value, red, green, blue, alpha;
} amrvr_colormap_pt;

/* One global box record, identical on every rank after amrvr_scene_upload():
 * what VisibilityOrdering.cpp:85-152 allgathers (bounds, owner) plus what
 * DirectSendBase.cpp:329-376 needs (owner-local layer index). */
typedef struct amrvr_box_record {
  double min_corner[3];
  double max_corner[3];
  int32_t dims[3];
  int32_t owner;       /* rank that paints this box */
  int32_t local_index; /* index within the owner's box list (= layer index) */
  int32_t reserved;
} amrvr_box_record;

/* Per-frame statistics of one amrvr_raymarch() call. */
typedef struct amrvr_march_stats {
  int64_t samples;          /* executions of VolumePainter.cpp:869-894 (0 unless counting on) */
  int64_t rays;             /* (pixel, box) pairs inside a footprint rectangle */
  int64_t fragment_pixels;  /* pixels stored in all footprint layers */
  float kernel_ms;          /* CUDA-event time of the march kernel */
} amrvr_march_stats;

/* ------------------------------------------------------------------ context */

/* [host-only] Version of this ABI. */
int amrvr_abi_version(void);

/* Number of CUDA devices visible to this process (0 when there is none). */
int amrvr_device_count(int* count);

/* [host-only] Create an NCCL unique id (rank 0 calls it and ships the
 * AMRVR_NCCL_ID_BYTES to the other ranks with whatever the host already has:
 * MPI_Bcast in the reference's world, torch.distributed here). */
int amrvr_nccl_unique_id(void* id_out);

/* Replaces the implicit per-rank state of the reference (MPI_COMM_WORLD rank /
 * size read in VolumeRenderer::VolumeRenderer(), VolumeRenderer.cpp:557-560).
 * nccl_unique_id may be NULL iff nranks == 1. */
int amrvr_ctx_create(int cuda_device, int rank, int nranks,
                     const void* nccl_unique_id, amrvr_ctx** out);
void amrvr_ctx_destroy(amrvr_ctx* ctx);
/* Message of the last failure on this ctx (or of the last failed call that had
 * no ctx when ctx == NULL).  Never NULL. */
const char* amrvr_last_error(const amrvr_ctx* ctx);
/* The CUDA stream every kernel of this ctx is launched on (cudaStream_t). */
void* amrvr_ctx_stream(amrvr_ctx* ctx);
/* Number of kernels this ctx has launched so far. */
int64_t amrvr_ctx_kernel_launches(const amrvr_ctx* ctx);
/* Device pointer of the RGB8 image (width*height*3, top row first) the last
 * amrvr_render_frame built on this ctx (root only); valid until the next frame. */
const uint8_t* amrvr_ctx_frame_rgb8_device(const amrvr_ctx* ctx);
/* The visibility order amrvr_render_frame computed for its last frame
 * (BuildVisibilityOrderedGroup): rank_order[i] = rank the reference lets own
 * image piece i.  rank_order has nranks entries. */
int amrvr_ctx_last_rank_order(const amrvr_ctx* ctx, int* rank_order, int* graph_failed);
/* amrvr_order_boxes / amrvr_order_records run the device radix sort and the
 * device pair test from this many global boxes upwards (default 1024); below
 * it the same sequential driver uses its host back end, which is faster than a
 * kernel launch plus readback for a few hundred boxes.  0 = always device. */
int amrvr_ctx_set_order_device_min_boxes(amrvr_ctx* ctx, int min_boxes);

/* ------------------------------------------------------------------- scene */

/* K2.  Replaces reduceLocalScalarStats (VolumeRenderer/SceneBuilder.cpp:53-94):
 * min, max, min-positive and finite count over all cells of the given boxes. */
int amrvr_brick_stats(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                      double* min_value, double* max_value,
                      double* min_positive, int64_t* finite_count);

/* No counterpart in the reference.  amrvr_brick_stats for callers that cannot allocate device
 * memory: it needs no amrvr_ctx and calls neither cudaMalloc nor cudaFree nor anything else that
 * allocates.  Meant for in-situ hosts (an AMReX run whose arena already holds almost all of the
 * GPU's memory, or any code that bans allocation inside its time loop): there a cudaMalloc made by
 * a library can fail outright, and the matching cudaFree makes the whole device wait.
 * The per-CTA partial records of the reduction live in one fixed array in device memory that every
 * GPU loading this library holds from load to exit (about 2 MiB).  Consequences:
 *   - at most AMRVR_BRICK_STATS_STATIC_MAX_BOXES boxes per call (AMRVR_ERR_INVALID beyond that);
 *   - every box's `data` must be device (or managed) memory -- a host FAB would need a staging
 *     buffer, so it is refused with AMRVR_ERR_INVALID instead;
 *   - all calls on one GPU share that array: calls that overlap on the GPU (two streams, two
 *     threads) race and return garbage.  The call waits for `stream` before returning, so calls
 *     issued one after another are fine; nothing has to be cleared between them;
 *   - the first call on each GPU looks up the array's device address and keeps it; make that
 *     first call outside any stream capture.
 * cuda_device is made current.  stream is a cudaStream_t (NULL = the default stream).  Results are
 * host pointers (any may be NULL) and equal amrvr_brick_stats' bit for bit.  Errors are read with
 * amrvr_last_error(NULL). */
#define AMRVR_BRICK_STATS_STATIC_MAX_BOXES 4096
int amrvr_brick_stats_static(int cuda_device, void* stream, int nboxes,
                             const amrvr_box* boxes, double* min_value, double* max_value,
                             double* min_positive, int64_t* finite_count);

/* amrvr_scene_upload without the final wait (one rank; with several ranks it is
 * amrvr_scene_upload): returns while the FAB copies are still on PCIe.  The
 * caller keeps the FABs alive and unchanged until amrvr_scene_wait() or the
 * first amrvr_render_frame() of the scene has returned.  That first frame
 * marches the boxes of a copy group as soon as the group is quantised, so the
 * march hides behind the upload (the in-situ case of a known scalar range:
 * SceneBuilder.cpp:160-187 hands the data over, VolumeRenderer.cpp:1103-1339
 * renders it once).  Every other entry point waits for the upload first. */
int amrvr_scene_upload_async(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                             const amrvr_scalar_transform* transform,
                             float range_min, float range_max, amrvr_scene** out);
int amrvr_scene_wait(amrvr_ctx* ctx, amrvr_scene* scene);

/* amrvr_scene_upload for the in-situ call sequence (VolumeRendererApi.cpp:80-124, 257-395:
 * BuildSceneGeometry, then the painter reads the same FABs): amrvr_scene_build /
 * amrvr_brick_stats already moved the host FABs into the context's staging buffer to
 * reduce their statistics.  When `boxes` are the views that call staged last (same
 * pointers, extents, strides; nothing else used the staging buffer since) and the caller
 * asserts that their contents have not changed, the bricks are quantised from the staged
 * doubles and nothing crosses PCIe a second time; otherwise this is amrvr_scene_upload.
 * amrvr_scene_upload_was_staged tells which happened. */
int amrvr_scene_upload_staged(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                              const amrvr_scalar_transform* transform,
                              float range_min, float range_max, amrvr_scene** out);
int amrvr_scene_upload_was_staged(const amrvr_scene* scene);

/* K5.  Replaces the zero-copy Array4 views BuildSceneGeometry hands to the
 * painter (SceneBuilder.cpp:160-187): each FAB becomes a device-resident brick
 * of one byte per cell holding the colour-table index the reference computes
 * per sample (VolumeTypes.hpp:33-67 applyScalarTransform, VolumePainter.cpp:
 * 75-105 saturateSoftTail, :873-883 normalise/quantise).  Also allgathers the
 * box records of all ranks (collective when nranks > 1). */
int amrvr_scene_upload(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                       const amrvr_scalar_transform* transform,
                       float range_min, float range_max, amrvr_scene** out);
int amrvr_scene_free(amrvr_ctx* ctx, amrvr_scene* scene);
/* 1 when every rank holds the bricks of all ranks (NVLink box: allgathered at
 * upload), so that amrvr_render_frame deals march tiles, not boxes, to GPUs. */
int amrvr_scene_is_replicated(const amrvr_scene* scene);
/* Device time of the upload's FAB copies + quantise launches (CUDA events on the
 * ctx stream), FAB bytes and cells this rank brought.  Any pointer may be NULL. */
int amrvr_scene_upload_stats(const amrvr_scene* scene, float* device_ms,
                             int64_t* fab_bytes, int64_t* cells);
/* Global box table (all ranks, owner-major, owner-local order). */
int amrvr_scene_global_boxes(const amrvr_scene* scene, int* nglobal,
                             const amrvr_box_record** records);
/* Copy brick `local_box` (dims[0]*dims[1]*dims[2] bytes, i fastest) to host. */
int amrvr_scene_read_brick(amrvr_ctx* ctx, const amrvr_scene* scene,
                           int local_box, uint8_t* dst);

/* ------------------------------------------------------ scene construction */

/* One FAB of a MultiFab as MFIter hands it out (VolumeRenderer/SceneBuilder.cpp:
 * 148-187): valid box + the amrex::Array4<Real const> view.  `data` may be a
 * host or a device pointer (same addressing rule as amrvr_box). */
typedef struct amrvr_fab {
  int32_t lo[3];    /* validbox().smallEnd() */
  int32_t hi[3];    /* validbox().bigEnd()   */
  const double* data;
  int64_t jstride, kstride, nstride;
  int32_t begin[3]; /* Array4::begin (valid box grown by the ghost cells) */
  int32_t reserved;
} amrvr_fab;

/* One AMR level of api::AmrData (VolumeRenderer/VolumeRendererApi.hpp:22-26):
 * this rank's FABs in MFIter order + LevelGridGeometry (SceneBuilder.hpp). */
typedef struct amrvr_level {
  double prob_lo[3];
  double cell_size[3];
  int32_t nfabs;
  int32_t reserved;
  const amrvr_fab* fabs;
} amrvr_level;

/* What BuildSceneGeometry leaves in VolumeRenderer::SceneGeometry besides the
 * boxes (VolumeRenderer/VolumeRenderer.hpp:46-62). */
typedef struct amrvr_scene_summary {
  amrvr_scalar_transform transform;
  float scalar_range[2];
  float processed_range[2];
  float original_range[2];
  int32_t has_original_range;
  int32_t nboxes;          /* boxes written to boxes_out */
  double bounds_min[3];    /* explicitBounds (padded by 5 % of the longest side) */
  double bounds_max[3];
  double global_scale;     /* 1 / shortest side of the global extent */
} amrvr_scene_summary;

/* f1.  Replaces detail::BuildSceneGeometry (VolumeRenderer/SceneBuilder.cpp:
 * 113-425) for already-convex levels: box corners from the level geometry,
 * global rescale to a unit shortest side, scalar statistics on the GPU (K2),
 * the ScalarTransform and the three scalar ranges.  Levels outside
 * [min_level, max_level] are skipped as in the reference.  Collective when
 * nranks > 1 (one min- and one sum-allreduce instead of the reference's eleven
 * MPI_Allreduce calls).  boxes_out (capacity entries) receives this rank's
 * boxes in (level, MFIter) order, ready for amrvr_scene_upload.
 * Errors: AMRVR_ERR_RUNTIME with the reference's messages ("no data", "invalid
 * scalar range", "Log scaling requested but no positive scalar values were
 * found."); AMRVR_ERR_INVALID when capacity is too small. */
int amrvr_scene_build(amrvr_ctx* ctx, int nlevels, const amrvr_level* levels,
                      int min_level, int max_level, int component,
                      int log_scale_input, int normalize_to_data_range,
                      amrvr_box* boxes_out, int capacity,
                      amrvr_scene_summary* summary);

/* f4.  Replaces detail::ComputeSceneHistogram (SceneBuilder.cpp:445-577):
 * bin_count bins over [range_min, range_max] of the transformed scalar of
 * every cell of the given boxes, summed over all ranks (collective when
 * nranks > 1).  counts_host[bin_count]; *sample_count = total cells. */
int amrvr_scene_histogram(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                          const amrvr_scalar_transform* transform,
                          float range_min, float range_max, int bin_count,
                          uint64_t* counts_host, uint64_t* sample_count);

/* -------------------------------------------- plotfiles and convexify (f2, f3) */

/* f3.  Replaces the use the reference makes of amrex::PlotFileData
 * (VolumeRenderer/VolumeRenderer.cpp:599-679; AMReX 26.04, un-vendored): the
 * plotfile's `Header`, every level's `Level_n/Cell_H` (VisMF header: BoxArray +
 * FabOnDisk table) and the FAB records in `Cell_D_*` (text descriptor, then
 * the reals, component-major, Fortran order; 8- or 4-byte IEEE, either byte
 * order -- values are widened to double as amrex::Real = double does).
 * [host-only]; errors go to amrvr_last_error(NULL). */
typedef struct amrvr_plotfile amrvr_plotfile;

typedef struct amrvr_plotfile_info {
  int32_t space_dim;    /* PlotFileData::spaceDim()    */
  int32_t ncomp;        /* varNames().size()           */
  int32_t finest_level; /* finestLevel()               */
  int32_t reserved;
  double time;
  double prob_lo[3];    /* probLo()                    */
  double prob_hi[3];
} amrvr_plotfile_info;

typedef struct amrvr_plotfile_level {
  double cell_size[3];   /* cellSize(level)                               */
  int32_t ref_ratio[3];  /* refRatio(level) towards level+1; 0 on the finest */
  int32_t domain_lo[3];  /* probDomain(level)                             */
  int32_t domain_hi[3];
  int32_t nboxes;        /* boxArray(level).size()                        */
  int32_t real_bytes;    /* 8 or 4: precision of the FABs on disk         */
} amrvr_plotfile_level;

int amrvr_plotfile_open(const char* path, amrvr_plotfile** out);
void amrvr_plotfile_close(amrvr_plotfile* pf);
int amrvr_plotfile_get_info(const amrvr_plotfile* pf, amrvr_plotfile_info* info);
/* varNames()[component]; NULL when out of range.  Owned by pf. */
const char* amrvr_plotfile_var_name(const amrvr_plotfile* pf, int component);
/* Index of a variable name, -1 when absent (VolumeRenderer.cpp:614-624). */
int amrvr_plotfile_var_index(const amrvr_plotfile* pf, const char* name);
int amrvr_plotfile_get_level(const amrvr_plotfile* pf, int level,
                             amrvr_plotfile_level* out);
/* boxArray(level): lo_hi[6*b .. 6*b+5] = lo[3], hi[3] of box b. */
int amrvr_plotfile_level_boxes(const amrvr_plotfile* pf, int level,
                               int32_t* lo_hi, int capacity);
/* One component of one FAB into dst[nx*ny*nz] (i fastest), = what
 * PlotFileData::get(level, name) leaves in that FAB.  Thread-safe per call
 * (pread), so a caller may read several FABs concurrently. */
int amrvr_plotfile_read_fab(const amrvr_plotfile* pf, int level, int box,
                            int component, double* dst, int64_t capacity);
/* All FABs of `level` dealt to `rank` (box b belongs to rank b % nranks when
 * rule = 0, to contiguous chunks of ceil(n/nranks) when rule = 1) read with
 * `threads` reader threads into dst (dense, one after another in box order).
 * box_index_out[k] / offset_out[k] (may be NULL) name box k of the rank and
 * its first element in dst.  Returns the count through nread. */
int amrvr_plotfile_read_level(const amrvr_plotfile* pf, int level, int component,
                              int rank, int nranks, int rule, int threads,
                              double* dst, int64_t capacity,
                              int32_t* box_index_out, int64_t* offset_out,
                              int box_capacity, int32_t* nread);

/* f2.  Replaces amrex::convexify (AMReX 26.04 MultiFabUtil, un-vendored; call
 * sites VolumeRenderer.cpp:669-670, VolumeRendererApi.cpp:97-98) for one level
 * transition: every coarse box minus the union of the fine boxes coarsened by
 * `ratio` (floor division, as amrex::coarsen), cut into disjoint boxes with
 * AMReX's BoxList boxDiff order (highest dimension first, low side before high
 * side).  Piece k = out_lo_hi[6k..6k+5], lying wholly inside coarse box
 * out_source[k]; pieces come in coarse-box order.  No cell data moves: the
 * caller re-points an amrvr_fab at the source FAB with the piece as valid box
 * (amrvr_convexify_levels does that), and amrvr_scene_upload's strided
 * quantise pass is the masked copy.  [host-only]
 * Returns AMRVR_ERR_INVALID when capacity is too small (*npieces = needed). */
int amrvr_convexify_boxes(int ncoarse, const int32_t* coarse_lo_hi, int nfine,
                          const int32_t* fine_lo_hi, const int32_t ratio[3],
                          int32_t* out_lo_hi, int32_t* out_source, int capacity,
                          int32_t* npieces);

/* Whole hierarchy: levels[l] holds this rank's FABs; fine_boxes[l] /
 * nfine_boxes[l] the GLOBAL BoxArray of level l (what MultiFab::boxArray()
 * returns on every rank; NULL = this rank's FABs are the whole level, i.e. one
 * rank); ref_ratio[l] the ratio between level l and l+1.  fabs_out receives,
 * level after level, views into the callers' FABs restricted to the cells no
 * finer level covers (finest level unchanged); nfabs_out[l] counts them.
 * Pieces stay on the rank that owns the source FAB -- the reference lets AMReX
 * build a fresh DistributionMapping and ParallelCopy; here nothing is
 * exchanged.  [host-only] */
int amrvr_convexify_levels(int nlevels, const amrvr_level* levels,
                           const int32_t* const* fine_boxes,
                           const int32_t* nfine_boxes,
                           const int32_t (*ref_ratio)[3], amrvr_fab* fabs_out,
                           int capacity, int32_t* nfabs_out);

/* ---------------------------------------------------------------- ordering */

/* a10.  Replaces BuildVisibilityOrderedGroup (Common/VisibilityOrdering.hpp:
 * 14-23, .cpp:63-632): rank_order[i] = rank that owns image piece i.  The
 * MPI_Group of the reference is this int array.  Depth keys are sorted by the
 * device radix sort; returns identity order when use_graph == 0 or the graph
 * cannot be ordered (same fallback as :610-627, *graph_failed set to 1). */
int amrvr_order_boxes(amrvr_ctx* ctx, const amrvr_scene* scene,
                      const amrvr_camera* camera, float aspect, int use_graph,
                      int* rank_order /*[nranks]*/, int* graph_failed);

/* The same on an explicit global table (any nranks, independent of the ctx's
 * own rank count): what amrvr_order_boxes runs after the allgather. */
int amrvr_order_records(amrvr_ctx* ctx, int nranks, int nglobal,
                        const amrvr_box_record* records,
                        const amrvr_camera* camera, float aspect, int use_graph,
                        int* rank_order, int* graph_failed);

/* [host-only] Same computation on the host from an explicit global table;
 * used by the CPU tests and by amrvr_order_boxes for the sequential part. */
int amrvr_order_boxes_host(int nranks, int nglobal,
                           const amrvr_box_record* records,
                           const amrvr_camera* camera, float aspect,
                           int use_graph, int* rank_order, int* graph_failed);

/* [host-only] a10, the debugging aid behind writeVisibilityGraph.  The DOT
 * documents exportGraph writes to visibility_graph_<n>.dot (Common/
 * VisibilityOrdering.cpp:319-352): one per round of the cycle-breaking loop
 * (:567-569), byte for byte.  *nrounds = number of rounds the ordering took;
 * *needed = bytes of round `round` (no terminator).  With text == NULL this is
 * a size query; otherwise capacity must be >= *needed.  The caller decides
 * where the text goes (the reference: the working directory of rank 0). */
int amrvr_order_graph_dot_host(int nranks, int nglobal,
                               const amrvr_box_record* records,
                               const amrvr_camera* camera, float aspect,
                               int round, char* text, int64_t capacity,
                               int64_t* needed, int* nrounds);

/* [host-only] a6 + a11.  computeBoxDepthHint (VolumeRenderer.cpp:541-553) for
 * every global box and the layer order of DirectSendBase.cpp:363-388:
 * order[k] = index into the global table of the k-th layer to blend. */
int amrvr_layer_order_host(int nglobal, const amrvr_box_record* records,
                           const amrvr_camera* camera, float* depth_hints,
                           int* order);

/* [host-only] The plan of a frame striped over the GPUs of a box (one rank per GPU):
 * GPU q folds the 16-row strips q, q + nranks, ... of the frame and its march queue
 * holds, of every box, the tiles of those strips.  rects: x0, y0, w (padded to 4), h of
 * every visible footprint.  Out, indexed [q * nboxes + b]: first tile id, first frame row
 * of the queue's first strip of the box, float offset of the box's strips in pool q (a
 * layer stores them one after the other, 16 rows each); queue_tile_begin[nranks + 1],
 * pool_floats[nranks].  Where the reference deals a CONTIGUOUS piece to every rank
 * (DirectSendBase.cpp:59-74) this deals strips: which pixels a rank folds does not change
 * a pixel of the gathered frame. */
int amrvr_stripe_plan_host(int nranks, int nboxes, const int32_t* rects,
                           int32_t* tile_base, int32_t* ry_first, uint64_t* layer_off,
                           int32_t* queue_tile_begin, uint64_t* pool_floats);
/* [host-only] row of rank's layer of a footprint starting at row0 that holds frame row y
 * (-1: not a row of this rank) -- the address arithmetic of the fold kernel. */
int64_t amrvr_stripe_layer_row_host(int nranks, int rank, int row0, int y);

/* [host-only] a12.  getPieceRange (DirectSendBase.cpp:59-74). */
int amrvr_piece_range(int num_pixels, int piece, int num_pieces, int* begin,
                      int* end);

/* [host-only] a7.  referenceSampleDistance (VolumeRenderer.cpp:1138-1190). */
int amrvr_reference_sample_distance(int nglobal,
                                    const amrvr_box_record* records,
                                    const double bounds_min[3],
                                    const double bounds_max[3], float* out);

/* [host-only] a2.  buildColorTable (VolumePainter.cpp:442-516): 256 RGBA
 * entries for one normalisation factor. */
int amrvr_color_table(float alpha_scale, float normalization_factor,
                      float range_min, float range_max,
                      const amrvr_colormap_pt* pts, int npts,
                      float* table_rgba /*[256*4]*/);

/* --------------------------------------------------------------- ray-march */

/* a1-a5.  Replaces the loop of VolumePainter::paint calls in
 * renderSingleTrial (VolumeRenderer.cpp:1201-1219, Common/VolumePainter.cpp:
 * 548-961): all local boxes in one launch, each into a layer that stores only
 * the box's screen-space footprint rectangle.  count_samples != 0 turns the
 * per-sample counter on (slightly slower kernel). */
int amrvr_raymarch(amrvr_ctx* ctx, const amrvr_scene* scene,
                   const amrvr_camera* camera, int width, int height,
                   float box_transparency, float reference_sample_distance,
                   const amrvr_colormap_pt* pts, int npts, int count_samples,
                   amrvr_layers** out);
int amrvr_layers_free(amrvr_ctx* ctx, amrvr_layers* layers);
int amrvr_layers_count(const amrvr_layers* layers);
int amrvr_layers_stats(const amrvr_layers* layers, amrvr_march_stats* stats);
/* Footprint rectangle of a layer: pixels [x0,x0+w) x [y0,y0+h). */
int amrvr_layers_rect(const amrvr_layers* layers, int layer, int* x0, int* y0,
                      int* w, int* h);
/* Dense materialisation of one layer as the reference's
 * ImageRGBAFloatColorDepthSort buffer: width*height*5 floats
 * [r,g,b,a,depth], premultiplied, +inf depth where empty
 * (Common/ImageRGBAFloatColorDepthSort.hpp:9-44).  dst5 is a HOST pointer. */
int amrvr_layers_read(amrvr_ctx* ctx, const amrvr_layers* layers, int layer,
                      float* dst5);

/* --------------------------------------------------------------- composite */

/* a11-a13.  Replaces Compositor::compose / DirectSendBase::composeLayered
 * (Common/Compositor.hpp:35-37, DirectSend/Base/DirectSendBase.cpp:316-458):
 * this rank's fully blended piece, pixels [*region_begin, *region_end) of the
 * row-major frame, 5 floats per pixel, left on the device
 * (*piece5_device, owned by ctx, valid until the next composite). Collective.
 * The exchange of DirectSendBase.cpp:76-177, 400-446 (MPI_Isend/Irecv of layer pieces) is one
 * kernel here when the GPUs of the box can map each other's memory: it reads the row slices
 * another rank's piece needs out of the layer pool and stores them straight into that rank's
 * receive pool over NVLink, between two flag barriers; without peer access (or with
 * AMRVR_EXCHANGE=nccl) the slices are packed and travel with grouped ncclSend/ncclRecv. */
int amrvr_composite(amrvr_ctx* ctx, const amrvr_scene* scene,
                    const amrvr_layers* layers, const amrvr_camera* camera,
                    const int* rank_order, int* region_begin, int* region_end,
                    float** piece5_device);
/* Where the device time of the last amrvr_composite of this ctx went when it stored its slices
 * into the peers' receive pools (nranks > 1 with peer access; AMRVR_ERR_INVALID otherwise):
 * ms5 = { wait until every peer's pool is free, push_slices_kernel (this rank's stores over
 * NVLink), wait until every peer's slices have arrived, fold table copy + host gap, fold_bin +
 * fold }.  Waits for the ctx stream.  No reference counterpart (the reference times nothing
 * inside DirectSendBase::compose). */
int amrvr_ctx_last_composite_ms(amrvr_ctx* ctx, float* ms5);
/* [host-only] The exchange amrvr_composite performs for `rank`, without doing
 * it: floats this rank sends to / receives from every other rank (row slices
 * of footprint layers that intersect the destination piece) and its own pixel
 * range.  Every rank derives the same plan from the global table, so no count
 * exchange is needed: send_floats[q] on rank r == recv_floats[r] on rank q. */
int amrvr_exchange_plan_host(int nranks, int rank, int nglobal,
                             const amrvr_box_record* records,
                             const amrvr_camera* camera, int width, int height,
                             const int* rank_order, int64_t* send_floats,
                             int64_t* recv_floats, int* region_begin,
                             int* region_end);
/* The same composite in the reference's own calling convention, for the C++
 * adapter of DirectSendBase::compose: layer l of this rank is a dense
 * width*height*5-float HOST image (ImageRGBAFloatColorDepthSort buffer) with
 * depth hint depth_hints[l] (LayeredVolumeImage, Common/LayeredVolumeImage.hpp:
 * 18-57).  piece5_host (may be NULL) receives this rank's piece, pixels
 * [*region_begin, *region_end).  Collective; blocking. */
int amrvr_composite_dense(amrvr_ctx* ctx, int width, int height, int nlayers,
                          const float* const* layers5_host,
                          const float* depth_hints, const int* rank_order,
                          int* region_begin, int* region_end,
                          float* piece5_host);
/* Single-GPU emulation of the P-rank composite (no reference counterpart; for
 * hosts with fewer GPUs than ranks and for parity tests): the layers of this
 * ctx are treated as if box l were painted by virtual rank box_owner[l]
 * (owner-local order = increasing l), blended with the run structure P ranks
 * would produce, and the whole gathered frame (width*height*5 floats) is
 * written to full5_device.  piece_owner_host (may be NULL, width*height ints)
 * receives the virtual rank that would own each pixel. */
int amrvr_composite_virtual(amrvr_ctx* ctx, const amrvr_scene* scene,
                            const amrvr_layers* layers,
                            const amrvr_camera* camera, int nvirtual,
                            const int* box_owner, const int* rank_order,
                            float* full5_device, int* piece_owner_host);
/* a14.  Replaces ImageColorOnly::Gather (Common/ImageColorOnly.hpp:220-270).
 * full5_device (root only) receives width*height*5 floats. Collective. */
int amrvr_gather(amrvr_ctx* ctx, const float* piece5_device, int region_begin,
                 int region_end, int num_pixels, int root,
                 float* full5_device);
/* a15.  Replaces downsampleImage (VolumeRenderer.cpp:479-528). */
int amrvr_downsample(amrvr_ctx* ctx, const float* full5_device, int width,
                     int height, int sqrt_aa, float* out5_device);
/* a16.  Color::GetComponentAsByte (Common/Color.hpp:66-91) + the bottom-up row
 * order of SavePPM/SavePNG (Common/SavePPM.cpp:25-32): RGB8, top row first. */
int amrvr_quantize_rgb8(amrvr_ctx* ctx, const float* full5_device, int width,
                        int height, uint8_t* rgb8_device);

/* Device <-> host helpers so hosts without a CUDA runtime of their own (ctypes,
 * plain C++) can move results; all on the ctx stream, blocking. */
int amrvr_device_alloc(amrvr_ctx* ctx, size_t bytes, void** out);
int amrvr_device_free(amrvr_ctx* ctx, void* ptr);
int amrvr_copy_to_host(amrvr_ctx* ctx, void* dst_host, const void* src_device,
                       size_t bytes);
int amrvr_copy_to_device(amrvr_ctx* ctx, void* dst_device, const void* src_host,
                         size_t bytes);
int amrvr_host_alloc_pinned(size_t bytes, void** out);
int amrvr_host_free_pinned(void* ptr);
int amrvr_synchronize(amrvr_ctx* ctx);

/* The camera VolumeRenderer::renderScene builds when none is given
 * (VolumeRenderer/VolumeRenderer.cpp:975-1030): orbit at a random azimuth and
 * altitude (std::mt19937(seed), RenderParameters::cameraSeed = 91021 by
 * default) around the centre of the scene bounds, far enough to see all of it
 * at 45 degrees.  custom_up = RenderParameters::cameraUp or NULL.  [host-only] */
int amrvr_auto_camera(const double* bounds_min, const double* bounds_max,
                      unsigned seed, const double* custom_up, amrvr_camera* out);

/* ------------------------------------------------------------ whole frame */

typedef struct amrvr_frame_params {
  int32_t width, height;       /* RenderParameters (VolumeRenderer.hpp:33-44) */
  float box_transparency;
  int32_t antialiasing;        /* perfect square */
  int32_t use_visibility_graph;
  int32_t count_samples;
  int32_t draw_bounding_box;   /* renderBoundingBoxLayer, VolumeRenderer.cpp:1311-1314 */
  int32_t rgb8_on_device;      /* != 0: build the RGB8 image on the root's GPU even when
                                  rgb8_host is NULL (read it with amrvr_ctx_frame_rgb8_device) */
} amrvr_frame_params;

typedef struct amrvr_frame_stats {
  float march_ms, order_ms, composite_ms, gather_ms, post_ms;
  float fold_ms;           /* the fold kernel alone (inside composite_ms); in a sharded frame it
                              starts by waiting for the slowest GPU's march */
  int64_t samples;         /* samples THIS rank marched (any box: a sharded frame deals tiles, not boxes) */
  int64_t exchanged_bytes; /* fragment bytes this rank sent to other ranks: stored into their pools
                              over NVLink by the march (sharded frame) or sent with NCCL */
  int64_t stolen_tiles;    /* march tiles this rank took from other GPUs' queues (sharded frame) */
  int64_t tiles;           /* march tiles of the whole frame */
  int64_t fragment_pixels; /* (pixel, layer) fragments stored for this rank's piece */
  /* time line of this rank's march launch on the device clock (globaltimer) */
  int64_t march_span_ns;     /* first CTA start -> last CTA end */
  int64_t march_cta_busy_ns; /* sum over CTAs of (end - start) */
  int64_t march_table_ns;    /* sum over CTAs of the time spent staging transfer tables */
  int64_t march_ctas;        /* CTAs of the launch */
  float pre_ms;              /* entry of amrvr_render_frame -> the march's parameters are on the device
                                (host planning + one small copy; the GPU idles meanwhile) */
  float pad_;
} amrvr_frame_stats;

/* a1-a17 in one call = VolumeRenderer::renderSingleTrial
 * (VolumeRenderer.cpp:1103-1339) without the file write: on `root`,
 * rgba5_host (may be NULL) receives width*height*5 floats after downsample and
 * overlay, rgb8_host (may be NULL) the width*height*3 bytes SavePPM would
 * write.  Collective. */
int amrvr_render_frame(amrvr_ctx* ctx, const amrvr_scene* scene,
                       const amrvr_camera* camera,
                       const amrvr_frame_params* params,
                       const amrvr_colormap_pt* pts, int npts, int root,
                       float* rgba5_host, uint8_t* rgb8_host,
                       amrvr_frame_stats* stats);

/* NVLink probe (collective, nranks > 1 with peer access): every rank copies
 * `bytes` between its own memory and that of rank (rank + peer_offset) mod
 * nranks with a plain load/store kernel, `iters` times, all ranks at once.
 * write != 0: stores into the peer (what the march does with the tiles it takes
 * from another GPU's queue); write == 0: loads from the peer.  *gbytes_per_s is
 * this rank's rate: the measured per-GPU, per-direction link figure. */
int amrvr_peer_bandwidth(amrvr_ctx* ctx, size_t bytes, int peer_offset, int write,
                         int iters, float* gbytes_per_s);

#ifdef __cplusplus
}
#endif
#endif /* AMRVR_H */
