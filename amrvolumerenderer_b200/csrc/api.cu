// api.cu -- the C ABI of libamrvr_b200.so (include/amrvr.h): contexts, scene
// upload, ray-march, ordering, sort-last composite, post-processing.  No CPU
// fallback: compute entry points need an sm_100 device.
#include <cuda_runtime.h>

#include <algorithm>
#include <chrono>
#include <climits>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>
#include <limits>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <string>
#include <type_traits>
#include <unistd.h>
#include <vector>

#include "internal.hpp"
#include "nccl_dyn.hpp"

using namespace amrvr;

// ---------------------------------------------------------------------------
// objects
// ---------------------------------------------------------------------------

struct DeviceBuffer {  // grow-only device allocation
  void* ptr = nullptr;
  size_t bytes = 0;
  cudaError_t reserve(size_t want) {
    if (want <= bytes) return cudaSuccess;
    if (ptr != nullptr) {
      cudaError_t e = cudaFree(ptr);
      ptr = nullptr;
      bytes = 0;
      if (e != cudaSuccess) return e;
    }
    const size_t rounded = (want + (1u << 20) - 1) & ~static_cast<size_t>((1u << 20) - 1);
    cudaError_t e = cudaMalloc(&ptr, rounded);
    if (e == cudaSuccess) bytes = rounded;
    return e;
  }
  void release() {
    if (ptr != nullptr) cudaFree(ptr);
    ptr = nullptr;
    bytes = 0;
  }
  template <class T>
  T* as() const {
    return static_cast<T*>(ptr);
  }
};

// A buffer that exists with the same capacity on every rank of the box and is
// mapped into every process: peer[q] is rank q's copy as seen from here.
struct SymBuffer {
  void* local = nullptr;
  size_t cap = 0;
  std::vector<void*> peer;   // [nranks]; peer[me] == local
  std::vector<char> opened;  // peer[q] came from cudaIpcOpenMemHandle
};

// One helper thread per context, started on first use and parked on a condition
// variable between frames: the visibility ordering of a few hundred boxes is
// pure host work, so it runs here while the calling thread enqueues the march.
// (A thread created per frame costs about as much as the ordering itself.)
class HostWorker {
 public:
  ~HostWorker() {
    {
      std::lock_guard<std::mutex> lock(m_);
      quit_ = true;
    }
    cv_.notify_all();
    if (thread_.joinable()) thread_.join();
  }
  void submit(std::function<void()> job) {
    if (!thread_.joinable()) thread_ = std::thread([this] { loop(); });
    {
      std::lock_guard<std::mutex> lock(m_);
      job_ = std::move(job);
      pending_ = true;
      done_ = false;
    }
    cv_.notify_all();
  }
  void wait() {
    std::unique_lock<std::mutex> lock(m_);
    cv_.wait(lock, [this] { return done_; });
  }

 private:
  void loop() {
    std::unique_lock<std::mutex> lock(m_);
    for (;;) {
      cv_.wait(lock, [this] { return pending_ || quit_; });
      if (quit_) return;
      pending_ = false;
      std::function<void()> job = std::move(job_);
      lock.unlock();
      job();
      lock.lock();
      done_ = true;
      cv_.notify_all();
    }
  }
  std::thread thread_;
  std::mutex m_;
  std::condition_variable cv_;
  std::function<void()> job_;
  bool pending_ = false, done_ = true, quit_ = false;
};

struct amrvr_ctx {
  int device = 0, rank = 0, nranks = 1;
  cudaStream_t stream = nullptr;
  cudaStream_t aux = nullptr;  // visibility ordering: runs beside the march, never waits for it
  cudaStream_t plan = nullptr; // fold table copy + fold_bin of a frame: beside the march, and apart from `aux`,
                               // which the helper thread may be driving (ordering of >= 1024 boxes) at the same time
  cudaStream_t copy = nullptr;  // host->device FAB copies of amrvr_scene_upload, ahead of the quantise
  cudaStream_t quant = nullptr; // quantise launches of an asynchronous upload (amrvr_scene_upload_async)
  cudaEvent_t copy_ev[4] = {};  // copy-group fences (round-robin)
  cudaEvent_t copy_start = nullptr;
  int64_t launches = 0;
  int64_t order_launches = 0;   // kernels of the visibility ordering (it may run on the helper thread)
  int order_device_min = 1024;  // global box count from which the ordering runs its device kernels
  std::string error;
  HostWorker worker;
  ncclComm_t comm = nullptr;
  cudaEvent_t ev[8] = {};
  // per-frame parameter blocks and scratch
  DeviceBuffer stage;        // FAB doubles waiting for quantisation
  // what stage_boxes left in `stage` (one 64-bit word per descriptor field and box); valid
  // until anything else writes the staging buffer -- amrvr_scene_upload_staged quantises
  // straight from it instead of copying the same FABs a second time
  std::vector<uint64_t> staged_key;
  bool staged_valid = false;
  bool async_upload_inflight = false;  // an amrvr_scene_upload_async may still read `stage` / `quant_jobs`
  const amrvr_scene* async_upload_owner = nullptr;  // the scene of that upload (cleared with the flag)
  DeviceBuffer quant_jobs;   // QuantJob[] of the upload in flight
  DeviceBuffer march_boxes;  // MarchBox[]
  DeviceBuffer tables;       // float[ntables*1024]
  DeviceBuffer counter;      // sample counter + misc small results
  DeviceBuffer fold_layers;  // FoldLayer[]
  DeviceBuffer fold_bins;    // per-band layer lists of the fold (fold_scratch_bytes)
  DeviceBuffer copy_jobs;    // CopyJob[]
  DeviceBuffer send_buf, recv_buf;
  DeviceBuffer piece;        // composited piece (5 floats / pixel)
  DeviceBuffer full, small, rgb8;  // gathered frame, downsampled frame, bytes
  DeviceBuffer order_boxes, order_scratch_a, order_scratch_b, order_priority, order_edges;
  DeviceBuffer stats_partials;
  std::vector<DeviceBuffer> layer_pools;  // recycled layer pools
  std::vector<DeviceBuffer> brick_pools;  // recycled brick pools (amrvr_scene_free -> next upload)
  DeviceBuffer gather_counts, gather_table;  // scratch of the box-record allgather
  std::vector<cudaEvent_t> event_pool;    // recycled timing events
  // peer-memory exchange (nranks > 1 on one NVLink box; off -> NCCL alltoallv)
  bool peer_ok = false;
  SymBuffer sym_flags, sym_pool, sym_piece, sym_rgb8;
  SymBuffer sym_recv;  // amrvr_composite with peer access: every rank's receive pool, written by its peers
  cudaEvent_t seam_ev[4] = {};  // that composite on the stream: entry, pools free, slices stored, slices arrived
  bool seam_timed = false;
  uint32_t barrier_epoch = 0;
  uint32_t* peer_error_host = nullptr;  // pinned copy of the barrier time-out flag
  unsigned char* param_host = nullptr;  // pinned staging of the per-frame march parameter block
  size_t param_host_bytes = 0;
  cudaEvent_t param_copied = nullptr;   // the last copy out of param_host has completed
  cudaEvent_t bin_done = nullptr;       // frame_fused: fold table copied and binned on the plan stream
  uint8_t* last_rgb8 = nullptr;         // RGB8 frame of the last amrvr_render_frame (root)
  DeviceBuffer frame_pool;              // fragment pool of single-rank fused frames
  std::vector<int> last_rank_order;     // visibility order of the last frame (BuildVisibilityOrderedGroup)
  int last_graph_failed = 0;
  int table_slot_boxes = 256;           // box records per rank in the upload's allgather (grows, same on all ranks)
  bool poisoned = false;                // a frame barrier timed out: peers are out of step
  unsigned long long* frame_stats_host = nullptr;  // pinned: [samples, pushed bytes, stolen tiles, error word]
  // transfer tables already built, keyed by every input of buildColorTable
  struct TableEntry {
    std::vector<unsigned char> key;
    std::vector<float> rgba;
  };
  std::vector<TableEntry> table_cache;
};

struct LocalBox {
  double min_corner[3], max_corner[3];
  int32_t dims[3];
  uint64_t brick_off;
};

struct amrvr_scene {
  std::vector<amrvr_box_record> global;
  std::vector<LocalBox> local;
  int first_global = 0;  // global index of local box 0
  DeviceBuffer bricks;
  // replicated: `bricks` holds the bricks of EVERY rank (allgathered once at
  // upload, 1 B per cell), global box g at global_brick_off[g] -- any GPU can
  // then march any tile of the frame.  Otherwise only the local bricks.
  bool replicated = false;
  std::vector<uint64_t> global_brick_off;
  std::vector<MarchBox> march_const_global;  // like march_const, for every global box
  cudaEvent_t up0 = nullptr, up1 = nullptr;  // around the FAB copies + quantise launches of the upload
  // asynchronous upload (amrvr_scene_upload_async, one rank): the bricks arrive group by
  // group; group_ready[g] fires when the bricks of copy group g are quantised, box_group[b]
  // is the group of local box b.  A frame marches the boxes of a group as soon as it is there.
  bool staged_reuse = false;  // amrvr_scene_upload_staged found the FABs in the staging buffer: no copy
  bool pending = false;
  std::vector<cudaEvent_t> group_ready;
  std::vector<int> box_group;
  int64_t fab_bytes = 0, fab_cells = 0;
  float range_min = 0.f, range_max = 1.f;
  // MarchBox fields that depend on the box alone (corners, cell sizes, exactness
  // of the reciprocal, epsilon, brick offset): built once, on the first frame
  std::vector<MarchBox> march_const;
};

struct amrvr_layers {
  int width = 0, height = 0;
  amrvr_camera camera{};
  std::vector<Rect> rects;
  std::vector<Rect> global_rects;  // footprints of ALL boxes of the scene (the composite's plan)
  std::vector<uint64_t> offs;  // float offsets into pool
  DeviceBuffer pool;
  bool pool_shared = false;  // pool is the context's peer-mapped pool (not owned)
  uint64_t pool_floats = 0;
  bool counted = false;
  cudaEvent_t t0 = nullptr, t1 = nullptr;
  amrvr_march_stats stats{};
  bool stats_ready = false;
};

namespace {

thread_local std::string g_error = "";

// AMRVR_TRACE=1: host wall-clock of every stage of amrvr_render_frame on stderr.
struct Trace {
  bool on;
  std::chrono::steady_clock::time_point t0, last;
  std::string line;
  Trace() : on(std::getenv("AMRVR_TRACE") != nullptr) { t0 = last = std::chrono::steady_clock::now(); }
  void mark(const char* what) {
    if (!on) return;
    const auto now = std::chrono::steady_clock::now();
    line += std::string(what) + "=" +
            std::to_string(std::chrono::duration<double, std::micro>(now - last).count()) + "us ";
    last = now;
  }
  ~Trace() {
    if (on) std::fprintf(stderr, "[amrvr trace] %s\n", line.c_str());
  }
};

int fail(amrvr_ctx* ctx, int status, const std::string& msg) {
  if (ctx != nullptr) ctx->error = msg;
  g_error = msg;
  return status;
}

}  // namespace
int amrvr::report_error(int status, const std::string& msg) { return fail(nullptr, status, msg); }
namespace {

#define CUDA_TRY(ctx, expr)                                                              \
  do {                                                                                   \
    const cudaError_t err__ = (expr);                                                    \
    if (err__ != cudaSuccess)                                                            \
      return fail(ctx, AMRVR_ERR_CUDA,                                                   \
                  std::string(#expr) + ": " + cudaGetErrorString(err__));                \
  } while (0)

#define LAUNCH_TRY(ctx, expr)                                                            \
  do {                                                                                   \
    const int err__ = (expr);                                                            \
    if (err__ != 0)                                                                      \
      return fail(ctx, AMRVR_ERR_CUDA,                                                   \
                  std::string(#expr) + ": " +                                            \
                      cudaGetErrorString(static_cast<cudaError_t>(err__)));              \
  } while (0)

#define NCCL_TRY(ctx, expr)                                                              \
  do {                                                                                   \
    const ncclResult_t err__ = (expr);                                                   \
    if (err__ != ncclSuccess)                                                            \
      return fail(ctx, AMRVR_ERR_NCCL,                                                   \
                  std::string(#expr) + ": " + nccl_api().GetErrorString(err__));         \
  } while (0)

#define REQUIRE(ctx, cond, msg) \
  do {                          \
    if (!(cond)) return fail(ctx, AMRVR_ERR_INVALID, msg); \
  } while (0)

bool is_pow2_float(float v) {
  if (!(v > 0.0f) || !std::isfinite(v)) return false;
  int e;
  return std::frexp(v, &e) == 0.5f;
}

// Distance by which the kernel shrinks a box before it trusts "this sample is
// inside and its cell index is in range" without testing.  The float position
// e + d*t carries at most 2^-24 * (|d*t| + |p|) of rounding error and the slab
// bounds of the shrunk box a few ulps more; 64 * 2^-24 * (that magnitude) is a
// 16-fold cover.  The floor() of (p - lo)/cell additionally needs the margin to
// be a visible fraction of a cell.  Returns 0 (always test) when the two cannot
// both hold.
float inside_margin(const MarchBox& M, const float* eye) {
  float coord = 0.0f, cell_min = INFINITY;
  for (int d = 0; d < 3; ++d) {
    coord = std::max(coord, std::max(std::fabs(M.lo[d]), std::fabs(M.hi[d])));
    coord = std::max(coord, std::fabs(eye[d]));
    if (!(M.cell[d] > 0.0f) || !std::isfinite(M.cell[d]) || M.n[d] < 1) return 0.0f;
    cell_min = std::min(cell_min, M.cell[d]);
  }
  const float magnitude = 4.0f * coord;  // |e| + |d t| + |p| <= 2 (|e| + |p|)
  const float rounding = 64.0f * 5.9604645e-8f * magnitude;
  const float margin = std::max(rounding, cell_min * 0.0625f);
  if (!std::isfinite(margin) || margin > 0.25f * cell_min) return 0.0f;
  return margin;
}

struct DeviceOrderBackend final : OrderBackend {
  amrvr_ctx* ctx;
  int status = 0;
  int nranks;
  explicit DeviceOrderBackend(amrvr_ctx* c) : ctx(c), nranks(c->nranks) {}

  int upload(const std::vector<OrderBox>& boxes) {
    const size_t n = boxes.size();
    if (ctx->order_boxes.reserve(n * sizeof(OrderBox)) != cudaSuccess) return AMRVR_ERR_CUDA;
    if (cudaMemcpyAsync(ctx->order_boxes.ptr, boxes.data(), n * sizeof(OrderBox),
                        cudaMemcpyHostToDevice, ctx->aux) != cudaSuccess)
      return AMRVR_ERR_CUDA;
    return 0;
  }
  void rank_boxes(const std::vector<OrderBox>& boxes, std::vector<int32_t>& priority) override {
    const int n = static_cast<int>(boxes.size());
    priority.assign(n, 0);
    if (status != 0 || n == 0) return;
    // the boxes were uploaded by pair_edges() of the same round
    if (ctx->order_scratch_a.reserve(static_cast<size_t>(n) * 16) != cudaSuccess ||
        ctx->order_scratch_b.reserve(static_cast<size_t>(n) * 16) != cudaSuccess ||
        ctx->order_priority.reserve(static_cast<size_t>(n) * 4) != cudaSuccess) {
      status = AMRVR_ERR_CUDA;
      return;
    }
    // up to 32 K boxes: rank by counting over all pairs (many CTAs, one launch);
    // beyond that, or with AMRVR_ORDER_SORT=radix, the single-CTA LSD radix sort
    const char* mode = std::getenv("AMRVR_ORDER_SORT");
    const bool radix = n > 32768 || (mode != nullptr && std::strcmp(mode, "radix") == 0);
    int launched;
    if (radix) {
      int max_owner = 0;
      for (const OrderBox& b : boxes) max_owner = std::max(max_owner, b.owner);
      int owner_bits = 0;
      while ((1 << owner_bits) <= max_owner) ++owner_bits;
      launched = launch_order_rank(ctx->order_boxes.as<OrderBox>(), n, ctx->order_scratch_a.ptr,
                                   ctx->order_scratch_b.ptr, owner_bits,
                                   ctx->order_priority.as<int32_t>(), ctx->aux, &ctx->order_launches);
    } else {
      launched = (cudaMemsetAsync(ctx->order_priority.ptr, 0, static_cast<size_t>(n) * 4, ctx->aux) ==
                  cudaSuccess)
                     ? launch_order_count(ctx->order_boxes.as<OrderBox>(), n,
                                          ctx->order_priority.as<int32_t>(), ctx->aux,
                                          &ctx->order_launches)
                     : 1;
    }
    if (launched != 0 ||
        cudaMemcpyAsync(priority.data(), ctx->order_priority.ptr, static_cast<size_t>(n) * 4,
                        cudaMemcpyDeviceToHost, ctx->aux) != cudaSuccess ||
        cudaStreamSynchronize(ctx->aux) != cudaSuccess)
      status = AMRVR_ERR_CUDA;
  }
  void pair_edges(const std::vector<OrderBox>& boxes, const float* dir,
                  std::vector<OrderEdge>& edges) override {
    edges.clear();
    const int n = static_cast<int>(boxes.size());
    if (status != 0 || n == 0) return;
    if ((status = upload(boxes)) != 0) return;
    int capacity = 16 * n + 64;
    for (int attempt = 0; attempt < 2; ++attempt) {
      if (ctx->order_edges.reserve(static_cast<size_t>(capacity) * sizeof(OrderEdge) + 16) !=
          cudaSuccess) {
        status = AMRVR_ERR_CUDA;
        return;
      }
      // count lives in the first 16 bytes
      int32_t* count_dev = ctx->order_edges.as<int32_t>();
      OrderEdge* edges_dev = reinterpret_cast<OrderEdge*>(ctx->order_edges.as<char>() + 16);
      int count = 0;
      if (cudaMemsetAsync(count_dev, 0, 16, ctx->aux) != cudaSuccess ||
          launch_order_pairs(ctx->order_boxes.as<OrderBox>(), n, dir[0], dir[1], dir[2], edges_dev,
                             capacity, count_dev, ctx->aux, &ctx->order_launches) != 0 ||
          cudaMemcpyAsync(&count, count_dev, 4, cudaMemcpyDeviceToHost, ctx->aux) !=
              cudaSuccess ||
          cudaStreamSynchronize(ctx->aux) != cudaSuccess) {
        status = AMRVR_ERR_CUDA;
        return;
      }
      if (count <= capacity) {
        edges.resize(static_cast<size_t>(count));
        if (count > 0 &&
            (cudaMemcpyAsync(edges.data(), edges_dev, static_cast<size_t>(count) * sizeof(OrderEdge),
                             cudaMemcpyDeviceToHost, ctx->aux) != cudaSuccess ||
             cudaStreamSynchronize(ctx->aux) != cudaSuccess))
          status = AMRVR_ERR_CUDA;
        return;
      }
      capacity = count;  // rare: dense contact graph; retry once with room
    }
    status = AMRVR_ERR_RUNTIME;
  }
};

// Everything the composite needs to know about the frame, identical on every
// rank: blend order, run ids, footprints, pieces.
struct FramePlan {
  std::vector<int> order;       // blend position -> global box
  std::vector<int> run_of;      // global box -> run id
  std::vector<Rect> rects;      // global box -> footprint
  std::vector<int> piece_of_rank, begin, end;  // per rank: piece index, pixel range
};

void make_frame_plan(const amrvr_scene& scene, const amrvr_camera& cam, int width, int height,
                     int nranks, const int* rank_order, FramePlan& plan,
                     const std::vector<Rect>* known_rects = nullptr) {
  const int n = static_cast<int>(scene.global.size());
  plan.order.assign(n, 0);
  layer_order(n, scene.global.data(), cam, nullptr, plan.order.data());
  plan.run_of.assign(n, 0);
  int run = -1, prev_owner = -1;
  for (int pos = 0; pos < n; ++pos) {  // DirectSendBase.cpp:400-407
    const int g = plan.order[pos];
    if (pos == 0 || scene.global[g].owner != prev_owner) ++run;
    prev_owner = scene.global[g].owner;
    plan.run_of[g] = run;
  }
  CameraBasis basis;
  make_camera_basis(cam, basis);
  if (known_rects != nullptr && static_cast<int>(known_rects->size()) == n) {
    plan.rects = *known_rects;  // computed by amrvr_raymarch for the same camera and size
  } else {
    plan.rects.resize(n);
    for (int g = 0; g < n; ++g)
      plan.rects[g] = box_footprint(scene.global[g].min_corner, scene.global[g].max_corner,
                                    scene.global[g].dims, basis, cam, width, height);
  }
  plan.piece_of_rank.assign(nranks, 0);
  plan.begin.assign(nranks, 0);
  plan.end.assign(nranks, 0);
  for (int piece = 0; piece < nranks; ++piece) {
    const int r = rank_order[piece];
    plan.piece_of_rank[r] = piece;
    piece_range(width * height, piece, nranks, &plan.begin[r], &plan.end[r]);
  }
}

// rows [s0,s1) of rect r needed by the pixel range [begin,end) of a W-wide frame
bool slice_rows(const Rect& r, int begin, int end, int width, int* s0, int* s1) {
  if (end <= begin || r.w <= 0 || r.h <= 0) return false;
  const int yb = begin / width, ye = (end - 1) / width;
  *s0 = std::max(r.y0, yb);
  *s1 = std::min(r.y0 + r.h, ye + 1);
  return *s1 > *s0;
}

}  // namespace

namespace {

struct SymRecord {  // what every rank tells the others about its new allocation
  cudaIpcMemHandle_t handle;
  unsigned long long pointer;
  long long pid;
  int device;
  int ok;
};

void sym_release(SymBuffer& b, int me) {
  for (size_t q = 0; q < b.peer.size(); ++q)
    if (static_cast<int>(q) != me && b.opened[q] && b.peer[q] != nullptr) cudaIpcCloseMemHandle(b.peer[q]);
  if (b.local != nullptr) cudaFree(b.local);
  b = SymBuffer{};
}

PeerPtrs peer_ptrs(const SymBuffer& b) {
  PeerPtrs out{};
  for (size_t q = 0; q < b.peer.size() && q < static_cast<size_t>(kMaxPeers); ++q) out.p[q] = b.peer[q];
  return out;
}

// Collective: afterwards every rank holds >= bytes and has every peer's copy
// mapped.  `bytes` must be the same on every rank (it is always derived from
// the global box table or the frame size), so all ranks grow in the same call.
// *all_ok reports whether EVERY rank could map every peer.
int sym_reserve(amrvr_ctx* ctx, SymBuffer& buf, size_t bytes, bool* all_ok) {
  if (all_ok) *all_ok = true;
  if (bytes <= buf.cap && buf.local != nullptr) return AMRVR_OK;
  const int P = ctx->nranks, me = ctx->rank;
  NcclApi& nc = nccl_api();
  const size_t want = ((bytes + bytes / 4) + (size_t(2) << 20) - 1) & ~((size_t(2) << 20) - 1);
  SymBuffer fresh;
  fresh.peer.assign(static_cast<size_t>(P), nullptr);
  fresh.opened.assign(static_cast<size_t>(P), 0);
  SymRecord mine{};
  mine.ok = 1;
  if (cudaMalloc(&fresh.local, want) != cudaSuccess) {
    (void)cudaGetLastError();
    fresh.local = nullptr;
    mine.ok = 0;
  } else {
    cudaMemsetAsync(fresh.local, 0, want, ctx->stream);
    if (cudaIpcGetMemHandle(&mine.handle, fresh.local) != cudaSuccess) {
      (void)cudaGetLastError();
      mine.ok = 0;
    }
  }
  mine.pointer = reinterpret_cast<unsigned long long>(fresh.local);
  mine.pid = static_cast<long long>(getpid());
  mine.device = ctx->device;
  // exchange the records (this allgather is also the barrier that makes it safe
  // to drop the old buffers afterwards: every rank has finished using them)
  DeviceBuffer tmp;
  CUDA_TRY(ctx, tmp.reserve(sizeof(SymRecord) * static_cast<size_t>(P + 1)));
  std::vector<SymRecord> all(static_cast<size_t>(P));
  auto allgather = [&](SymRecord& rec) -> int {
    CUDA_TRY(ctx, cudaMemcpyAsync(tmp.as<SymRecord>() + P, &rec, sizeof(SymRecord),
                                  cudaMemcpyHostToDevice, ctx->stream));
    NCCL_TRY(ctx, nc.AllGather(tmp.as<SymRecord>() + P, tmp.ptr, sizeof(SymRecord), ncclChar,
                               ctx->comm, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(all.data(), tmp.ptr, sizeof(SymRecord) * P, cudaMemcpyDeviceToHost,
                                  ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    return AMRVR_OK;
  };
  int st = allgather(mine);
  if (st != AMRVR_OK) return st;
  bool ok = true;
  for (int q = 0; q < P; ++q) ok = ok && all[q].ok != 0;
  if (ok) {
    fresh.peer[me] = fresh.local;
    for (int q = 0; q < P && ok; ++q) {
      if (q == me) continue;
      if (all[q].pid == mine.pid) {  // ranks are threads of one process: plain peer access
        if (all[q].device != ctx->device) {
          const cudaError_t e = cudaDeviceEnablePeerAccess(all[q].device, 0);
          if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) ok = false;
          (void)cudaGetLastError();
        }
        fresh.peer[q] = reinterpret_cast<void*>(all[q].pointer);
      } else {
        void* mapped = nullptr;
        if (cudaIpcOpenMemHandle(&mapped, all[q].handle, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) {
          (void)cudaGetLastError();
          ok = false;
        } else {
          fresh.peer[q] = mapped;
          fresh.opened[q] = 1;
        }
      }
    }
  }
  // second round: did everyone map everyone?
  mine.ok = ok ? 1 : 0;
  st = allgather(mine);
  if (st != AMRVR_OK) return st;
  for (int q = 0; q < P; ++q) ok = ok && all[q].ok != 0;
  tmp.release();
  if (!ok) {
    sym_release(fresh, me);
    if (all_ok) *all_ok = false;
    return AMRVR_OK;  // not an error: the caller falls back to the NCCL exchange
  }
  sym_release(buf, me);
  buf = fresh;
  buf.cap = want;
  return AMRVR_OK;
}

// Barrier kernel on the ctx stream (row `which` of the flag block).
int peer_barrier(amrvr_ctx* ctx, int which) {
  ++ctx->barrier_epoch;
  uint32_t* error_flag = static_cast<uint32_t*>(ctx->sym_flags.local) + kCtlError;
  LAUNCH_TRY(ctx, launch_peer_barrier(peer_ptrs(ctx->sym_flags), ctx->rank, ctx->nranks, which,
                                      ctx->barrier_epoch, 10ull * 1000 * 1000 * 1000, error_flag,
                                      ctx->stream, &ctx->launches));
  return AMRVR_OK;
}

}  // namespace

// ---------------------------------------------------------------------------
// context
// ---------------------------------------------------------------------------

extern "C" int amrvr_abi_version(void) { return AMRVR_ABI_VERSION; }

extern "C" int amrvr_device_count(int* count) {
  if (count == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "count is NULL");
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    (void)cudaGetLastError();
    n = 0;
  }
  *count = n;
  return AMRVR_OK;
}

extern "C" int amrvr_nccl_unique_id(void* id_out) {
  if (id_out == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "id_out is NULL");
  static_assert(sizeof(ncclUniqueId) == AMRVR_NCCL_ID_BYTES, "NCCL id size");
  const std::string why = nccl_api().load();
  if (!why.empty()) return fail(nullptr, AMRVR_ERR_NCCL, why);
  ncclUniqueId id;
  NCCL_TRY(nullptr, nccl_api().GetUniqueId(&id));
  std::memcpy(id_out, &id, sizeof(id));
  return AMRVR_OK;
}

extern "C" int amrvr_ctx_create(int cuda_device, int rank, int nranks, const void* nccl_unique_id,
                                amrvr_ctx** out) {
  if (out == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "out is NULL");
  *out = nullptr;
  if (nranks < 1 || rank < 0 || rank >= nranks)
    return fail(nullptr, AMRVR_ERR_INVALID, "rank/nranks out of range");
  if (nranks > 1 && nccl_unique_id == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "nccl_unique_id required when nranks > 1");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) {
    (void)cudaGetLastError();
    return fail(nullptr, AMRVR_ERR_CUDA,
                "no CUDA device visible: libamrvr_b200 has no CPU fallback");
  }
  if (cuda_device < 0 || cuda_device >= ndev)
    return fail(nullptr, AMRVR_ERR_INVALID, "cuda_device out of range");
  cudaDeviceProp prop;
  CUDA_TRY(nullptr, cudaGetDeviceProperties(&prop, cuda_device));
  if (prop.major != 10)
    return fail(nullptr, AMRVR_ERR_CUDA,
                std::string("device '") + prop.name + "' is sm_" + std::to_string(prop.major) +
                    std::to_string(prop.minor) + "; this library carries sm_100a code only");
  CUDA_TRY(nullptr, cudaSetDevice(cuda_device));
  // every early return below releases what has been created so far (streams, events, the
  // NCCL communicator, peer mappings): amrvr_ctx_destroy copes with a half-built context
  struct Destroy {
    void operator()(amrvr_ctx* c) const { amrvr_ctx_destroy(c); }
  };
  std::unique_ptr<amrvr_ctx, Destroy> ctx(new amrvr_ctx);
  ctx->device = cuda_device;
  ctx->rank = rank;
  ctx->nranks = nranks;
  CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
  CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->aux, cudaStreamNonBlocking));
  CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->plan, cudaStreamNonBlocking));
  CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->copy, cudaStreamNonBlocking));
  CUDA_TRY(nullptr, cudaStreamCreateWithFlags(&ctx->quant, cudaStreamNonBlocking));
  for (auto& e : ctx->copy_ev) CUDA_TRY(nullptr, cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
  CUDA_TRY(nullptr, cudaEventCreateWithFlags(&ctx->copy_start, cudaEventDisableTiming));
  for (auto& e : ctx->ev) CUDA_TRY(nullptr, cudaEventCreate(&e));
  CUDA_TRY(nullptr, ctx->counter.reserve(256));
  if (nranks > 1) {
    const std::string why = nccl_api().load();
    if (!why.empty()) return fail(nullptr, AMRVR_ERR_NCCL, why);
    ncclUniqueId id;
    std::memcpy(&id, nccl_unique_id, sizeof(id));
    NCCL_TRY(nullptr, nccl_api().CommInitRank(&ctx->comm, nranks, id, rank));
    // peer-memory exchange: every rank must be able to map every other rank's
    // buffers (one NVLink box).  AMRVR_EXCHANGE=nccl keeps the NCCL alltoallv.
    const char* mode = std::getenv("AMRVR_EXCHANGE");
    const bool want_peer = nranks <= kMaxPeers && !(mode != nullptr && std::strcmp(mode, "nccl") == 0);
    if (want_peer) {
      bool ok = false;
      amrvr_ctx* raw = ctx.get();
      const int st = sym_reserve(raw, raw->sym_flags, kCtlWords * sizeof(uint32_t), &ok);
      if (st != AMRVR_OK) {
        g_error = raw->error;
        return st;
      }
      raw->peer_ok = ok;
      if (ok) CUDA_TRY(nullptr, cudaMallocHost(reinterpret_cast<void**>(&raw->peer_error_host), 64));
      if (ok) *raw->peer_error_host = 0;
    }
  }
  *out = ctx.release();
  return AMRVR_OK;
}

extern "C" void amrvr_ctx_destroy(amrvr_ctx* ctx) {
  if (ctx == nullptr) return;
  cudaSetDevice(ctx->device);
  if (ctx->stream != nullptr) cudaStreamSynchronize(ctx->stream);
  for (SymBuffer* b : {&ctx->sym_flags, &ctx->sym_pool, &ctx->sym_piece, &ctx->sym_rgb8, &ctx->sym_recv})
    sym_release(*b, ctx->rank);
  if (ctx->peer_error_host != nullptr) cudaFreeHost(ctx->peer_error_host);
  if (ctx->param_host != nullptr) cudaFreeHost(ctx->param_host);
  if (ctx->frame_stats_host != nullptr) cudaFreeHost(ctx->frame_stats_host);
  ctx->frame_pool.release();
  if (ctx->param_copied != nullptr) cudaEventDestroy(ctx->param_copied);
  if (ctx->bin_done != nullptr) cudaEventDestroy(ctx->bin_done);
  if (ctx->comm != nullptr) nccl_api().CommDestroy(ctx->comm);
  for (DeviceBuffer* b :
       {&ctx->stage, &ctx->march_boxes, &ctx->tables, &ctx->counter, &ctx->fold_layers, &ctx->fold_bins,
        &ctx->copy_jobs, &ctx->send_buf, &ctx->recv_buf, &ctx->piece, &ctx->full, &ctx->small,
        &ctx->rgb8, &ctx->order_boxes, &ctx->order_scratch_a, &ctx->order_scratch_b,
        &ctx->order_priority, &ctx->order_edges, &ctx->stats_partials, &ctx->gather_counts,
        &ctx->gather_table})
    b->release();
  for (auto& p : ctx->layer_pools) p.release();
  for (auto& p : ctx->brick_pools) p.release();
  for (auto& e : ctx->ev)
    if (e != nullptr) cudaEventDestroy(e);
  for (cudaEvent_t e : ctx->event_pool) cudaEventDestroy(e);
  for (cudaStream_t st : {ctx->stream, ctx->aux, ctx->plan, ctx->quant, ctx->copy})
    if (st != nullptr) cudaStreamDestroy(st);
  for (auto& e : ctx->copy_ev)
    if (e != nullptr) cudaEventDestroy(e);
  if (ctx->copy_start != nullptr) cudaEventDestroy(ctx->copy_start);
  for (auto& e : ctx->seam_ev)
    if (e != nullptr) cudaEventDestroy(e);
  ctx->quant_jobs.release();
  delete ctx;
}

extern "C" const char* amrvr_last_error(const amrvr_ctx* ctx) {
  return (ctx != nullptr) ? ctx->error.c_str() : g_error.c_str();
}
extern "C" void* amrvr_ctx_stream(amrvr_ctx* ctx) { return ctx ? ctx->stream : nullptr; }
extern "C" int64_t amrvr_ctx_kernel_launches(const amrvr_ctx* ctx) {
  return ctx ? ctx->launches + ctx->order_launches : 0;
}
extern "C" int amrvr_ctx_last_rank_order(const amrvr_ctx* ctx, int* rank_order, int* graph_failed) {
  if (ctx == nullptr || rank_order == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "NULL argument");
  if (static_cast<int>(ctx->last_rank_order.size()) != ctx->nranks)
    return fail(nullptr, AMRVR_ERR_INVALID, "no frame has been rendered on this context");
  for (int i = 0; i < ctx->nranks; ++i) rank_order[i] = ctx->last_rank_order[i];
  if (graph_failed) *graph_failed = ctx->last_graph_failed;
  return AMRVR_OK;
}
extern "C" const uint8_t* amrvr_ctx_frame_rgb8_device(const amrvr_ctx* ctx) {
  return ctx ? ctx->last_rgb8 : nullptr;
}
extern "C" int amrvr_ctx_set_order_device_min_boxes(amrvr_ctx* ctx, int min_boxes) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  ctx->order_device_min = min_boxes < 0 ? 0 : min_boxes;
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// scene
// ---------------------------------------------------------------------------
namespace {

cudaError_t take_event(amrvr_ctx* ctx, cudaEvent_t* out) {
  if (!ctx->event_pool.empty()) {
    *out = ctx->event_pool.back();
    ctx->event_pool.pop_back();
    return cudaSuccess;
  }
  return cudaEventCreate(out);
}

struct StagedBox {
  const double* dev;  // device-visible pointer to the first valid cell
  int64_t jstride, kstride;
};

// Make the valid region of every box device-visible.  Host FABs are copied
// into ctx->stage (compacted, no ghost cells); device FABs are used in place.
// The staging buffer and the quantise job table are shared by every upload / statistics pass
// of a context: anything that is about to overwrite them first lets an asynchronous upload
// that may still be reading them finish.
int drain_async_upload(amrvr_ctx* ctx) {
  if (!ctx->async_upload_inflight) return AMRVR_OK;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->copy));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->quant));
  ctx->async_upload_inflight = false;
  ctx->async_upload_owner = nullptr;
  return AMRVR_OK;
}

// the asynchronous upload of `scene` is known to be complete: nothing reads the staging buffer any more
void async_upload_done(amrvr_ctx* ctx, const amrvr_scene* scene) {
  if (ctx->async_upload_owner != scene) return;  // a later upload owns the flag
  ctx->async_upload_inflight = false;
  ctx->async_upload_owner = nullptr;
}

// identity of a list of FAB views: pointers, extents, strides (not the corners: scene_build rescales them)
std::vector<uint64_t> staging_key(int nboxes, const amrvr_box* boxes) {
  std::vector<uint64_t> key;
  key.reserve(static_cast<size_t>(std::max(nboxes, 0)) * 14 + 1);
  key.push_back(static_cast<uint64_t>(nboxes));
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    key.push_back(reinterpret_cast<uintptr_t>(B.data));
    for (int d = 0; d < 3; ++d) {
      key.push_back(static_cast<uint64_t>(static_cast<uint32_t>(B.dims[d])));
      key.push_back(static_cast<uint64_t>(static_cast<uint32_t>(B.lo[d])));
      key.push_back(static_cast<uint64_t>(static_cast<uint32_t>(B.begin[d])));
    }
    key.push_back(static_cast<uint64_t>(B.jstride));
    key.push_back(static_cast<uint64_t>(B.kstride));
    key.push_back(static_cast<uint64_t>(B.nstride));
    key.push_back(static_cast<uint64_t>(static_cast<uint32_t>(B.component)));
  }
  return key;
}

int stage_boxes(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes, std::vector<StagedBox>& out) {
  if (drain_async_upload(ctx) != AMRVR_OK) return AMRVR_ERR_CUDA;
  ctx->staged_valid = false;
  out.resize(static_cast<size_t>(std::max(nboxes, 0)));
  size_t need = 0;
  std::vector<char> on_device(out.size(), 0);
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    const size_t cells = static_cast<size_t>(std::max(B.dims[0], 0)) * std::max(B.dims[1], 0) *
                         std::max(B.dims[2], 0);
    if (cells == 0) continue;
    REQUIRE(ctx, B.data != nullptr, "box data pointer is NULL");
    cudaPointerAttributes attr;
    const cudaError_t e = cudaPointerGetAttributes(&attr, B.data);
    if (e != cudaSuccess) (void)cudaGetLastError();
    on_device[b] =
        (e == cudaSuccess && (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged));
    if (!on_device[b]) need += (cells * sizeof(double) + 255) & ~static_cast<size_t>(255);
  }
  CUDA_TRY(ctx, ctx->stage.reserve(need));
  size_t cursor = 0;
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    const int nx = B.dims[0], ny = B.dims[1], nz = B.dims[2];
    const size_t cells = static_cast<size_t>(std::max(nx, 0)) * std::max(ny, 0) * std::max(nz, 0);
    if (cells == 0) {
      out[b] = StagedBox{nullptr, 0, 0};
      continue;
    }
    const double* first = B.data + (B.lo[0] - B.begin[0]) + (B.lo[1] - B.begin[1]) * B.jstride +
                          (B.lo[2] - B.begin[2]) * B.kstride + B.component * B.nstride;
    if (on_device[b]) {
      out[b] = StagedBox{first, B.jstride, B.kstride};
      continue;
    }
    double* dst = reinterpret_cast<double*>(ctx->stage.as<char>() + cursor);
    cursor += (cells * sizeof(double) + 255) & ~static_cast<size_t>(255);
    if (B.jstride == nx && B.kstride == static_cast<int64_t>(nx) * ny) {
      CUDA_TRY(ctx, cudaMemcpyAsync(dst, first, cells * sizeof(double), cudaMemcpyHostToDevice,
                                    ctx->stream));
    } else {
      for (int k = 0; k < nz; ++k)  // one pitched copy per k-slab
        CUDA_TRY(ctx, cudaMemcpy2DAsync(dst + static_cast<size_t>(k) * nx * ny, nx * sizeof(double),
                                        first + k * B.kstride, B.jstride * sizeof(double),
                                        nx * sizeof(double), ny, cudaMemcpyHostToDevice,
                                        ctx->stream));
    }
    out[b] = StagedBox{dst, nx, static_cast<int64_t>(nx) * ny};
  }
  ctx->staged_key = staging_key(nboxes, boxes);  // same offsets as upload_bricks lays the FABs out
  ctx->staged_valid = true;
  return AMRVR_OK;
}

// Host FABs -> bytes, pipelined: the copy stream moves FAB doubles into the
// staging buffer group by group (about 16 MB, a third of a millisecond of PCIe
// Gen5) while the main stream quantises the group before it with ONE batched
// launch, so the quantise and its launches hide behind the H2D copies and the
// upload costs the copy time plus one small kernel.  Device-resident FABs
// (in-situ) are quantised in place with a single launch.
int upload_bricks(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                  const amrvr_scalar_transform& transform, float range_min, float range_max,
                  amrvr_scene& scene, bool async_groups = false, bool reuse_staged = false) {
  // reuse_staged: the caller asserts that the FABs are the ones the last amrvr_brick_stats /
  // amrvr_scene_build of this context staged and have not changed since: when the views
  // match, their doubles are still in the staging buffer and nothing is copied
  const bool staged = reuse_staged && !async_groups && ctx->staged_valid &&
                      ctx->staged_key == staging_key(nboxes, boxes);
  ctx->staged_valid = false;  // an upload that copies overwrites the staging buffer
  // asynchronous: quantise launches go to their own stream and leave one event per group
  cudaStream_t qstream = async_groups ? ctx->quant : ctx->stream;
  if (async_groups) scene.box_group.assign(static_cast<size_t>(nboxes), 0);
  struct Item {
    int box;
    bool on_device;
    size_t cells, stage_off;
  };
  std::vector<Item> items;
  size_t need = 0;
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    const size_t cells = static_cast<size_t>(std::max(B.dims[0], 0)) * std::max(B.dims[1], 0) *
                         std::max(B.dims[2], 0);
    if (cells == 0) continue;
    REQUIRE(ctx, B.data != nullptr, "box data pointer is NULL");
    cudaPointerAttributes attr;
    const cudaError_t e = cudaPointerGetAttributes(&attr, B.data);
    if (e != cudaSuccess) (void)cudaGetLastError();
    const bool dev =
        (e == cudaSuccess && (attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged));
    items.push_back(Item{b, dev, cells, need});
    if (!dev) need += (cells * sizeof(double) + 255) & ~static_cast<size_t>(255);
  }
  if (items.empty()) return AMRVR_OK;
  CUDA_TRY(ctx, ctx->stage.reserve(need));
  CUDA_TRY(ctx, ctx->quant_jobs.reserve(items.size() * sizeof(QuantJob)));

  std::vector<QuantJob> jobs(items.size());
  for (size_t i = 0; i < items.size(); ++i) {
    const amrvr_box& B = boxes[items[i].box];
    const int nx = B.dims[0], ny = B.dims[1];
    QuantJob& J = jobs[i];
    if (items[i].on_device) {
      J.src = B.data + (B.lo[0] - B.begin[0]) + (B.lo[1] - B.begin[1]) * B.jstride +
              (B.lo[2] - B.begin[2]) * B.kstride + B.component * B.nstride;
      J.jstride = B.jstride;
      J.kstride = B.kstride;
    } else {
      J.src = reinterpret_cast<const double*>(ctx->stage.as<char>() + items[i].stage_off);
      J.jstride = nx;
      J.kstride = static_cast<int64_t>(nx) * ny;
    }
    J.dst = scene.bricks.as<uint8_t>() + scene.local[items[i].box].brick_off;
    J.nx = nx;
    J.ny = ny;
    J.nz = B.dims[2];
    J.pad = 0;
  }
  // the staging buffer and the job table may still be read by work queued on the main stream
  CUDA_TRY(ctx, cudaEventRecord(ctx->copy_start, ctx->stream));
  CUDA_TRY(ctx, cudaStreamWaitEvent(ctx->copy, ctx->copy_start, 0));
  if (async_groups) CUDA_TRY(ctx, cudaStreamWaitEvent(qstream, ctx->copy_start, 0));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->quant_jobs.ptr, jobs.data(), jobs.size() * sizeof(QuantJob),
                                cudaMemcpyHostToDevice, qstream));
  if (async_groups) CUDA_TRY(ctx, cudaStreamSynchronize(qstream));  // `jobs` is a local

  const size_t group_bytes = 16u << 20;
  size_t first = 0, in_group = 0, max_cells = 0;
  int fence = 0;
  auto flush = [&](size_t end, bool copied) -> int {
    if (end == first) return AMRVR_OK;
    if (copied) {
      cudaEvent_t ev = ctx->copy_ev[fence++ & 3];
      CUDA_TRY(ctx, cudaEventRecord(ev, ctx->copy));
      CUDA_TRY(ctx, cudaStreamWaitEvent(qstream, ev, 0));
    }
    LAUNCH_TRY(ctx, launch_brick_quantise_batch(ctx->quant_jobs.as<QuantJob>() + first,
                                                static_cast<int>(end - first),
                                                static_cast<long long>(max_cells), transform, range_min,
                                                range_max, qstream, &ctx->launches));
    if (async_groups) {
      cudaEvent_t ready = nullptr;
      CUDA_TRY(ctx, take_event(ctx, &ready));
      scene.group_ready.push_back(ready);  // owned by the scene from here on
      CUDA_TRY(ctx, cudaEventRecord(ready, qstream));
      const int g = static_cast<int>(scene.group_ready.size()) - 1;
      for (size_t i = first; i < end; ++i) scene.box_group[static_cast<size_t>(items[i].box)] = g;
    }
    first = end;
    in_group = 0;
    max_cells = 0;
    return AMRVR_OK;
  };
  bool group_copied = false;
  const char* run_src = nullptr;  // pending run of adjacent dense copies
  char* run_dst = nullptr;
  size_t run_bytes = 0;
  auto flush_run = [&]() -> int {
    if (run_bytes > 0)
      CUDA_TRY(ctx, cudaMemcpyAsync(run_dst, run_src, run_bytes, cudaMemcpyHostToDevice, ctx->copy));
    run_bytes = 0;
    return AMRVR_OK;
  };
  for (size_t i = 0; i < items.size(); ++i) {
    const Item& it = items[i];
    const amrvr_box& B = boxes[it.box];
    if (!it.on_device && !staged) {
      const int nx = B.dims[0], ny = B.dims[1], nz = B.dims[2];
      const double* src = B.data + (B.lo[0] - B.begin[0]) + (B.lo[1] - B.begin[1]) * B.jstride +
                          (B.lo[2] - B.begin[2]) * B.kstride + B.component * B.nstride;
      double* dst = reinterpret_cast<double*>(ctx->stage.as<char>() + it.stage_off);
      if (B.jstride == nx && B.kstride == static_cast<int64_t>(nx) * ny) {
        // dense FAB: merged with the copy before it when both ends are adjacent
        // (FABs carved from one arena block travel as one DMA transfer)
        const size_t bytes = it.cells * sizeof(double);
        if (run_bytes > 0 && reinterpret_cast<const char*>(src) == run_src + run_bytes &&
            reinterpret_cast<char*>(dst) == run_dst + run_bytes) {
          run_bytes += bytes;
        } else {
          const int st = flush_run();
          if (st != AMRVR_OK) return st;
          run_src = reinterpret_cast<const char*>(src);
          run_dst = reinterpret_cast<char*>(dst);
          run_bytes = bytes;
        }
      } else if (B.kstride % B.jstride == 0 && B.kstride / B.jstride >= ny) {
        // ghost cells or a sub-view: one 3-D pitched copy
        cudaMemcpy3DParms p{};
        p.srcPtr = make_cudaPitchedPtr(const_cast<double*>(src), B.jstride * sizeof(double),
                                       static_cast<size_t>(B.jstride), static_cast<size_t>(B.kstride / B.jstride));
        p.dstPtr = make_cudaPitchedPtr(dst, nx * sizeof(double), static_cast<size_t>(nx), static_cast<size_t>(ny));
        p.extent = make_cudaExtent(nx * sizeof(double), static_cast<size_t>(ny), static_cast<size_t>(nz));
        p.kind = cudaMemcpyHostToDevice;
        CUDA_TRY(ctx, cudaMemcpy3DAsync(&p, ctx->copy));
      } else {
        for (int k = 0; k < nz; ++k)  // irregular strides: one pitched copy per k-slab
          CUDA_TRY(ctx, cudaMemcpy2DAsync(dst + static_cast<size_t>(k) * nx * ny, nx * sizeof(double),
                                          src + k * B.kstride, B.jstride * sizeof(double),
                                          nx * sizeof(double), ny, cudaMemcpyHostToDevice, ctx->copy));
      }
      group_copied = true;
      in_group += it.cells * sizeof(double);
    }
    max_cells = std::max(max_cells, it.cells);
    if (in_group >= group_bytes) {
      int st = flush_run();
      if (st == AMRVR_OK) st = flush(i + 1, group_copied);
      if (st != AMRVR_OK) return st;
      group_copied = false;
    }
  }
  {
    const int st = flush_run();
    if (st != AMRVR_OK) return st;
  }
  if (staged) {  // the buffer still holds exactly these FABs
    ctx->staged_key = staging_key(nboxes, boxes);
    ctx->staged_valid = true;
    scene.staged_reuse = true;
  }
  return flush(items.size(), group_copied);
}

}  // namespace

extern "C" int amrvr_brick_stats(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                                 double* min_value, double* max_value, double* min_positive,
                                 int64_t* finite_count) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, nboxes >= 0 && (nboxes == 0 || boxes != nullptr), "bad box list");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  std::vector<StagedBox> staged;
  const int st = stage_boxes(ctx, nboxes, boxes, staged);
  if (st != AMRVR_OK) return st;
  const int blocks_per_box = 148 * 2;
  CUDA_TRY(ctx, ctx->stats_partials.reserve(
                    (static_cast<size_t>(std::max(nboxes, 1)) * blocks_per_box + 1) * 32));
  double* partials = ctx->stats_partials.as<double>();
  int nparts = 0;
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    const long long cells = 1LL * B.dims[0] * B.dims[1] * B.dims[2];
    if (staged[b].dev == nullptr || cells <= 0) continue;
    const int blocks = static_cast<int>(std::min<long long>(blocks_per_box, (cells + 255) / 256));
    LAUNCH_TRY(ctx, launch_brick_stats(staged[b].dev, staged[b].jstride, staged[b].kstride,
                                       B.dims[0], B.dims[1], B.dims[2], partials + 4 * nparts,
                                       blocks, ctx->stream, &ctx->launches));
    nparts += blocks;
  }
  double* out4 = partials + 4 * static_cast<size_t>(std::max(nboxes, 1)) * blocks_per_box;
  LAUNCH_TRY(ctx, launch_brick_stats_final(partials, nparts, out4, ctx->stream, &ctx->launches));
  double host4[4];
  CUDA_TRY(ctx, cudaMemcpyAsync(host4, out4, sizeof(host4), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (min_value) *min_value = host4[0];
  if (max_value) *max_value = host4[1];
  if (min_positive) *min_positive = host4[2];
  if (finite_count) std::memcpy(finite_count, &host4[3], 8);
  return AMRVR_OK;
}

namespace {

size_t padded_cells(const int32_t* dims) {
  const size_t cells = static_cast<size_t>(std::max(dims[0], 0)) * std::max(dims[1], 0) *
                       std::max(dims[2], 0);
  return (cells + 255) & ~static_cast<size_t>(255);
}

// Global box table: every rank's records, owner-major (the MPI_Allgather(v) of
// VisibilityOrdering.cpp:85-152 and DirectSendBase.cpp:329-361, once per scene).
// ONE NCCL allgather of fixed slots [count | records...]; the slot capacity is a
// per-context value every rank grows in step, so a second round happens only
// the first time a rank brings more boxes than any rank had before.
int gather_box_table(amrvr_ctx* ctx, const std::vector<amrvr_box_record>& mine,
                     std::vector<amrvr_box_record>& global, int* first_global) {
  const int P = ctx->nranks;
  NcclApi& nc = nccl_api();
  const int nboxes = static_cast<int>(mine.size());
  for (;;) {
    const int cap = ctx->table_slot_boxes;
    const size_t slot = 16 + static_cast<size_t>(cap) * sizeof(amrvr_box_record);
    DeviceBuffer& table = ctx->gather_table;
    CUDA_TRY(ctx, table.reserve(slot * (P + 1)));
    std::vector<char> send(slot, 0);
    std::memcpy(send.data(), &nboxes, sizeof(int));
    if (nboxes > 0)
      std::memcpy(send.data() + 16, mine.data(),
                  sizeof(amrvr_box_record) * static_cast<size_t>(std::min(nboxes, cap)));
    char* send_dev = table.as<char>() + slot * P;
    CUDA_TRY(ctx, cudaMemcpyAsync(send_dev, send.data(), slot, cudaMemcpyHostToDevice, ctx->stream));
    NCCL_TRY(ctx, nc.AllGather(send_dev, table.ptr, slot, ncclChar, ctx->comm, ctx->stream));
    std::vector<char> all(slot * P);
    CUDA_TRY(ctx, cudaMemcpyAsync(all.data(), table.ptr, all.size(), cudaMemcpyDeviceToHost,
                                  ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    int most = 0;
    for (int r = 0; r < P; ++r) {
      int c = 0;
      std::memcpy(&c, all.data() + slot * r, sizeof(int));
      most = std::max(most, c);
    }
    if (most > cap) {  // every rank sees the same counts and grows the same way
      int grown = cap;
      while (grown < most) grown *= 2;
      ctx->table_slot_boxes = grown;
      continue;
    }
    global.clear();
    for (int r = 0; r < P; ++r) {
      int c = 0;
      std::memcpy(&c, all.data() + slot * r, sizeof(int));
      if (r == ctx->rank) *first_global = static_cast<int>(global.size());
      const amrvr_box_record* recs = reinterpret_cast<const amrvr_box_record*>(all.data() + slot * r + 16);
      global.insert(global.end(), recs, recs + c);
    }
    return AMRVR_OK;
  }
}

}  // namespace

namespace {
// waits for an asynchronous upload (no-op otherwise); every entry point that reads
// the bricks calls it, except the single-rank frame, which marches group by group
int wait_upload(amrvr_ctx* ctx, const amrvr_scene* scene) {
  if (scene == nullptr || !scene->pending) return AMRVR_OK;
  amrvr_scene* sc = const_cast<amrvr_scene*>(scene);
  if (!sc->group_ready.empty()) CUDA_TRY(ctx, cudaEventSynchronize(sc->group_ready.back()));
  sc->pending = false;
  async_upload_done(ctx, scene);
  return AMRVR_OK;
}
int scene_upload_impl(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                      const amrvr_scalar_transform* transform, float range_min, float range_max,
                      amrvr_scene** out, bool want_async, bool reuse_staged = false);
}  // namespace

extern "C" int amrvr_scene_upload(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                                  const amrvr_scalar_transform* transform, float range_min,
                                  float range_max, amrvr_scene** out) {
  return scene_upload_impl(ctx, nboxes, boxes, transform, range_min, range_max, out, false);
}

extern "C" int amrvr_scene_upload_async(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                                        const amrvr_scalar_transform* transform, float range_min,
                                        float range_max, amrvr_scene** out) {
  return scene_upload_impl(ctx, nboxes, boxes, transform, range_min, range_max, out, true);
}

extern "C" int amrvr_scene_upload_staged(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                                         const amrvr_scalar_transform* transform, float range_min,
                                         float range_max, amrvr_scene** out) {
  return scene_upload_impl(ctx, nboxes, boxes, transform, range_min, range_max, out, false, true);
}

extern "C" int amrvr_scene_upload_was_staged(const amrvr_scene* scene) {
  return (scene != nullptr && scene->staged_reuse) ? 1 : 0;
}

extern "C" int amrvr_scene_wait(amrvr_ctx* ctx, amrvr_scene* scene) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene != nullptr, "scene is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return wait_upload(ctx, scene);
}

namespace {
int scene_upload_impl(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                      const amrvr_scalar_transform* transform, float range_min, float range_max,
                      amrvr_scene** out, bool want_async, bool reuse_staged) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, out != nullptr, "out is NULL");
  *out = nullptr;
  REQUIRE(ctx, nboxes >= 0 && (nboxes == 0 || boxes != nullptr), "bad box list");
  REQUIRE(ctx, transform != nullptr, "transform is NULL");
  for (int b = 0; b < nboxes; ++b)
    for (int d = 0; d < 3; ++d)
      REQUIRE(ctx, boxes[b].dims[d] <= (1 << 20), "box dimension exceeds 2^20 cells");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  // every failure return below hands the half-built scene to amrvr_scene_free: streams drained,
  // events back in the pool, the brick pool (possibly a recycled multi-GB one) back in brick_pools
  struct Abandon {
    amrvr_ctx* ctx;
    void operator()(amrvr_scene* s) const {
      if (s == nullptr) return;
      s->pending = true;  // copies of an asynchronous upload may be in flight
      amrvr_scene_free(ctx, s);
    }
  };
  std::unique_ptr<amrvr_scene, Abandon> scene(new amrvr_scene, Abandon{ctx});
  scene->range_min = range_min;
  scene->range_max = range_max;
  scene->local.resize(static_cast<size_t>(nboxes));
  Trace tr;

  // the box table travels first (nothing is queued on the stream yet, so its one
  // synchronisation does not wait for FAB copies)
  std::vector<amrvr_box_record> mine(static_cast<size_t>(nboxes));
  for (int b = 0; b < nboxes; ++b) {
    std::memset(&mine[b], 0, sizeof(amrvr_box_record));
    for (int d = 0; d < 3; ++d) {
      mine[b].min_corner[d] = boxes[b].min_corner[d];
      mine[b].max_corner[d] = boxes[b].max_corner[d];
      mine[b].dims[d] = boxes[b].dims[d];
    }
    mine[b].owner = ctx->rank;
    mine[b].local_index = b;
  }
  if (ctx->nranks == 1) {
    scene->global = mine;
    scene->first_global = 0;
  } else {
    const int st = gather_box_table(ctx, mine, scene->global, &scene->first_global);
    if (st != AMRVR_OK) return st;
  }
  tr.mark("upload_table");

  // Brick layout.  On an NVLink box every GPU keeps every brick (1 B per cell: a
  // 1024^3 volume is 1 GB of 180): rank r's bricks at r * slot, so that a frame
  // can hand any tile to any GPU.  AMRVR_REPLICATE_MAX_MB bounds the pool
  // (default 32 GB); above it, and without peer access, each rank keeps its own.
  const int P = ctx->nranks;
  std::vector<size_t> rank_bytes(static_cast<size_t>(P), 0);
  for (const amrvr_box_record& r : scene->global)
    if (r.owner >= 0 && r.owner < P) rank_bytes[r.owner] += padded_cells(r.dims);
  const size_t slot = std::max<size_t>(*std::max_element(rank_bytes.begin(), rank_bytes.end()), 256);
  size_t limit_mb = 32768;
  if (const char* env = std::getenv("AMRVR_REPLICATE_MAX_MB")) limit_mb = std::strtoull(env, nullptr, 10);
  scene->replicated = P > 1 && ctx->peer_ok && slot * P <= limit_mb * (size_t(1) << 20);
  const size_t base = scene->replicated ? slot * ctx->rank : 0;
  size_t brick_bytes = 0;
  for (int b = 0; b < nboxes; ++b) {
    LocalBox& L = scene->local[b];
    for (int d = 0; d < 3; ++d) {
      L.min_corner[d] = boxes[b].min_corner[d];
      L.max_corner[d] = boxes[b].max_corner[d];
      L.dims[d] = boxes[b].dims[d];
    }
    L.brick_off = base + brick_bytes;
    brick_bytes += padded_cells(L.dims);
  }
  if (scene->replicated || P == 1) {
    scene->global_brick_off.resize(scene->global.size());
    std::vector<size_t> cursor(static_cast<size_t>(P), 0);
    for (size_t g = 0; g < scene->global.size(); ++g) {
      const int owner = scene->global[g].owner;
      scene->global_brick_off[g] = (scene->replicated ? slot * owner : 0) + cursor[owner];
      cursor[owner] += padded_cells(scene->global[g].dims);
    }
  }
  {
    const size_t need = scene->replicated ? slot * P : std::max<size_t>(brick_bytes, 256);
    for (size_t i = 0; i < ctx->brick_pools.size(); ++i)
      if (ctx->brick_pools[i].bytes >= need && ctx->brick_pools[i].bytes <= 2 * need + (1u << 20)) {
        scene->bricks = ctx->brick_pools[i];
        ctx->brick_pools.erase(ctx->brick_pools.begin() + static_cast<long>(i));
        break;
      }
    CUDA_TRY(ctx, scene->bricks.reserve(need));
  }
  tr.mark("upload_setup");
  CUDA_TRY(ctx, take_event(ctx, &scene->up0));
  CUDA_TRY(ctx, take_event(ctx, &scene->up1));
  for (int b = 0; b < nboxes; ++b) {
    const int64_t cells = static_cast<int64_t>(std::max(boxes[b].dims[0], 0)) *
                          std::max(boxes[b].dims[1], 0) * std::max(boxes[b].dims[2], 0);
    scene->fab_cells += cells;
    scene->fab_bytes += cells * static_cast<int64_t>(sizeof(double));
  }
  // asynchronous only where a frame can use it: one rank (a sharded scene is allgathered whole)
  const bool async_groups = want_async && P == 1;
  if (drain_async_upload(ctx) != AMRVR_OK) return AMRVR_ERR_CUDA;  // an earlier one still reads the staging buffer
  cudaStream_t qstream = async_groups ? ctx->quant : ctx->stream;
  if (async_groups) {
    CUDA_TRY(ctx, cudaEventRecord(ctx->copy_start, ctx->stream));
    CUDA_TRY(ctx, cudaStreamWaitEvent(qstream, ctx->copy_start, 0));
  }
  CUDA_TRY(ctx, cudaEventRecord(scene->up0, qstream));
  {
    const int st = upload_bricks(ctx, nboxes, boxes, *transform, range_min, range_max, *scene, async_groups,
                                 reuse_staged);
    if (st != AMRVR_OK) return st;
  }
  CUDA_TRY(ctx, cudaEventRecord(scene->up1, qstream));
  tr.mark("upload_enqueue");
  if (async_groups) {
    // returns with the copies in flight: the caller keeps the FABs alive until
    // amrvr_scene_wait or the first frame of this scene has returned
    scene->pending = !scene->group_ready.empty();
    ctx->async_upload_inflight = scene->pending;
    ctx->async_upload_owner = scene->pending ? scene.get() : nullptr;
    *out = scene.release();
    return AMRVR_OK;
  }
  if (scene->replicated)  // in place: every rank contributes its own slot
    NCCL_TRY(ctx, nccl_api().AllGather(scene->bricks.as<char>() + slot * ctx->rank, scene->bricks.ptr,
                                       slot, ncclChar, ctx->comm, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));  // host FABs may be reused on return
  tr.mark("upload_sync");
  *out = scene.release();
  return AMRVR_OK;
}
}  // namespace

extern "C" int amrvr_stripe_plan_host(int nranks, int nboxes, const int32_t* rects, int32_t* tile_base,
                                      int32_t* ry_first, uint64_t* layer_off, int32_t* queue_tile_begin,
                                      uint64_t* pool_floats) {
  if (nranks < 1 || nranks > kMaxPeers || nboxes < 0 || (nboxes > 0 && rects == nullptr) ||
      tile_base == nullptr || ry_first == nullptr || layer_off == nullptr || queue_tile_begin == nullptr ||
      pool_floats == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  std::vector<int32_t> ry0(static_cast<size_t>(nboxes)), rh(static_cast<size_t>(nboxes)), rw(static_cast<size_t>(nboxes));
  for (int b = 0; b < nboxes; ++b) {
    ry0[static_cast<size_t>(b)] = rects[4 * b + 1];
    rw[static_cast<size_t>(b)] = rects[4 * b + 2];
    rh[static_cast<size_t>(b)] = rects[4 * b + 3];
    if (rw[static_cast<size_t>(b)] <= 0 || rh[static_cast<size_t>(b)] <= 0 || ry0[static_cast<size_t>(b)] < 0)
      return fail(nullptr, AMRVR_ERR_INVALID, "empty footprint or negative row");
  }
  std::vector<QueueItem> items(static_cast<size_t>(nranks) * std::max(nboxes, 1));
  stripe_plan(nranks, nboxes, ry0.data(), rh.data(), rw.data(), items.data(), queue_tile_begin, pool_floats);
  for (size_t i = 0; i < static_cast<size_t>(nranks) * nboxes; ++i) {
    tile_base[i] = items[i].tile_base;
    ry_first[i] = items[i].ry_first;
    layer_off[i] = items[i].layer_off;
  }
  return AMRVR_OK;
}

extern "C" int64_t amrvr_stripe_layer_row_host(int nranks, int rank, int row0, int y) {
  if (nranks < 1 || rank < 0 || rank >= nranks) return -1;
  if (((y >> 4) % nranks) != rank || (y >> 4) < stripe_first_strip(nranks, rank, row0)) return -1;
  return stripe_layer_row(nranks, rank, row0, y);
}

extern "C" int amrvr_scene_is_replicated(const amrvr_scene* scene) {
  return (scene != nullptr && scene->replicated) ? 1 : 0;
}

extern "C" int amrvr_scene_upload_stats(const amrvr_scene* scene, float* device_ms, int64_t* fab_bytes,
                                        int64_t* cells) {
  if (scene == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "scene is NULL");
  if (device_ms) {
    *device_ms = 0.0f;
    if (scene->up0 != nullptr && scene->up1 != nullptr &&
        cudaEventElapsedTime(device_ms, scene->up0, scene->up1) != cudaSuccess) {
      (void)cudaGetLastError();
      return fail(nullptr, AMRVR_ERR_CUDA, "event timing failed");
    }
  }
  if (fab_bytes) *fab_bytes = scene->fab_bytes;
  if (cells) *cells = scene->fab_cells;
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// scene construction (f1) and histogram (f4)
// ---------------------------------------------------------------------------
namespace {

// min over ranks of `mins` and sum over ranks of `sums` (two NCCL allreduces on
// the ctx stream, one readback).  Maxima travel negated.
int allreduce_min_sum(amrvr_ctx* ctx, double* mins, int nmin, int64_t* sums, int nsum) {
  if (ctx->nranks == 1) return AMRVR_OK;
  NcclApi& nc = nccl_api();
  const size_t bytes = static_cast<size_t>(nmin) * sizeof(double) + static_cast<size_t>(nsum) * sizeof(int64_t);
  CUDA_TRY(ctx, ctx->gather_counts.reserve(std::max<size_t>(bytes, 256)));
  double* dmin = ctx->gather_counts.as<double>();
  int64_t* dsum = reinterpret_cast<int64_t*>(dmin + nmin);
  CUDA_TRY(ctx, cudaMemcpyAsync(dmin, mins, nmin * sizeof(double), cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(dsum, sums, nsum * sizeof(int64_t), cudaMemcpyHostToDevice, ctx->stream));
  NCCL_TRY(ctx, nc.AllReduce(dmin, dmin, static_cast<size_t>(nmin), ncclDouble, ncclMin, ctx->comm, ctx->stream));
  NCCL_TRY(ctx, nc.AllReduce(dsum, dsum, static_cast<size_t>(nsum), ncclInt64, ncclSum, ctx->comm, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(mins, dmin, nmin * sizeof(double), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(sums, dsum, nsum * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return AMRVR_OK;
}

}  // namespace

extern "C" int amrvr_scene_build(amrvr_ctx* ctx, int nlevels, const amrvr_level* levels,
                                 int min_level, int max_level, int component, int log_scale_input,
                                 int normalize_to_data_range, amrvr_box* boxes_out, int capacity,
                                 amrvr_scene_summary* summary) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, nlevels >= 0 && (nlevels == 0 || levels != nullptr) && summary != nullptr &&
                   capacity >= 0 && (capacity == 0 || boxes_out != nullptr) && component >= 0,
          "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  std::memset(summary, 0, sizeof(*summary));
  const double big = std::numeric_limits<double>::max();
  const double inf = std::numeric_limits<double>::infinity();

  // SceneBuilder.cpp:134-196: boxes of the requested levels, corners from the level geometry
  std::vector<amrvr_box> local;
  double lmin[3] = {big, big, big}, lmax[3] = {-big, -big, -big};
  for (int level = min_level; level <= max_level; ++level) {
    if (level < 0 || level >= nlevels) continue;
    const amrvr_level& L = levels[level];
    REQUIRE(ctx, L.nfabs >= 0 && (L.nfabs == 0 || L.fabs != nullptr), "bad level");
    for (int f = 0; f < L.nfabs; ++f) {
      const amrvr_fab& F = L.fabs[f];
      const int nx = F.hi[0] - F.lo[0] + 1, ny = F.hi[1] - F.lo[1] + 1, nz = F.hi[2] - F.lo[2] + 1;
      if (nx <= 0 || ny <= 0 || nz <= 0) continue;
      amrvr_box B;
      std::memset(&B, 0, sizeof(B));
      for (int d = 0; d < 3; ++d) {
        B.min_corner[d] = L.prob_lo[d] + static_cast<double>(F.lo[d]) * L.cell_size[d];
        B.max_corner[d] = L.prob_lo[d] + static_cast<double>(F.hi[d] + 1) * L.cell_size[d];
        B.lo[d] = F.lo[d];
        B.begin[d] = F.begin[d];
        lmin[d] = std::min(lmin[d], B.min_corner[d]);
        lmax[d] = std::max(lmax[d], B.max_corner[d]);
      }
      B.dims[0] = nx;
      B.dims[1] = ny;
      B.dims[2] = nz;
      B.data = F.data;
      B.jstride = F.jstride;
      B.kstride = F.kstride;
      B.nstride = F.nstride;
      B.component = component;
      local.push_back(B);
    }
  }
  REQUIRE(ctx, static_cast<int>(local.size()) <= capacity, "boxes_out is too small");

  // scalar statistics of the local cells on the GPU (K2; reduceLocalScalarStats :53-94)
  double smin = inf, smax = -inf, spos = inf;
  int64_t finite = 0;
  if (!local.empty()) {
    const int st = amrvr_brick_stats(ctx, static_cast<int>(local.size()), local.data(), &smin, &smax,
                                     &spos, &finite);
    if (st != AMRVR_OK) return st;
  }
  const double send_min = (finite > 0) ? smin : inf;   // :312-319
  const double send_max = (finite > 0) ? smax : -inf;
  const double send_pos = (spos > 0.0 && std::isfinite(spos)) ? spos : inf;  // :361-366

  // one min-allreduce (maxima negated) + one sum-allreduce replace :210-221,
  // :281-292, :324-341, :367-385
  double mins[11] = {lmin[0], lmin[1], lmin[2], -lmax[0], -lmax[1], -lmax[2], send_min, -send_max,
                     send_pos, 0.0, 0.0};
  int64_t sums[2] = {finite, (send_pos < inf) ? 1 : 0};
  {
    const int st = allreduce_min_sum(ctx, mins, 9, sums, 2);
    if (st != AMRVR_OK) return st;
  }
  double gmin[3] = {mins[0], mins[1], mins[2]}, gmax[3] = {-mins[3], -mins[4], -mins[5]};
  const double global_scalar_min = mins[6], global_scalar_max = -mins[7], global_pos_min = mins[8];
  const int64_t global_finite = sums[0], global_pos_count = sums[1];

  // :230-257 global rescale to a unit shortest side
  double min_extent = big;
  for (int d = 0; d < 3; ++d) {
    const double len = std::fabs(gmax[d] - gmin[d]);
    if (len > 0.0 && std::isfinite(len)) min_extent = std::min(min_extent, len);
  }
  double scale = 1.0;
  if (min_extent > 0.0 && std::isfinite(min_extent)) scale = 1.0 / min_extent;
  if (!std::isfinite(scale) || !(scale > 0.0)) scale = 1.0;
  if (scale != 1.0) {
    for (amrvr_box& B : local)
      for (int d = 0; d < 3; ++d) {
        B.min_corner[d] *= scale;
        B.max_corner[d] *= scale;
      }
    // min/max commute with the (monotone, identical) multiplication, so the second
    // pair of allreduces of the reference (:281-292) gives exactly these values
    for (int d = 0; d < 3; ++d) {
      gmin[d] *= scale;
      gmax[d] *= scale;
    }
  }
  if (gmin[0] > gmax[0] || gmin[1] > gmax[1] || gmin[2] > gmax[2])
    return fail(ctx, AMRVR_ERR_RUNTIME, "no data");  // options.noDataError
  const double ext[3] = {gmax[0] - gmin[0], gmax[1] - gmin[1], gmax[2] - gmin[2]};
  const double max_extent = std::max(ext[0], std::max(ext[1], ext[2]));
  const double pad = (max_extent > 0.0) ? max_extent * 0.05 : 1.0;
  for (int d = 0; d < 3; ++d) {
    summary->bounds_min[d] = gmin[d] - pad;
    summary->bounds_max[d] = gmax[d] + pad;
  }
  summary->global_scale = scale;

  auto make_range = [](double lo, double hi, float* out) {  // makeScalarRange :103-109
    if (lo == hi) hi = lo + 1.0;
    out[0] = static_cast<float>(lo);
    out[1] = static_cast<float>(hi);
  };
  if (global_finite > 0 && std::isfinite(global_scalar_min) && std::isfinite(global_scalar_max)) {
    make_range(global_scalar_min, global_scalar_max, summary->original_range);
    summary->has_original_range = 1;
  }
  double pmin = global_scalar_min, pmax = global_scalar_max;
  amrvr_scalar_transform& T = summary->transform;
  T.log_scale_input = log_scale_input ? 1 : 0;
  T.normalize_to_unit_range = 0;
  T.positive_floor = 0.0;
  T.inverse_normalization_span = 1.0;
  T.inverse_processed_span = 1.0;
  T.processed_max = 1.0;
  T.normalization_max = 1.0;
  if (log_scale_input) {
    if (global_pos_count <= 0 || !std::isfinite(global_pos_min) || !(global_pos_min > 0.0))
      return fail(ctx, AMRVR_ERR_RUNTIME,
                  "Log scaling requested but no positive scalar values were found.");
    T.positive_floor = global_pos_min;
    pmin = std::log(global_pos_min);
    pmax = std::log(std::max(global_scalar_max, global_pos_min));
  }
  if (!std::isfinite(pmin) || !std::isfinite(pmax))
    return fail(ctx, AMRVR_ERR_RUNTIME, "invalid scalar range");  // options.invalidScalarError
  if (pmin == pmax) pmax = pmin + 1.0;
  make_range(pmin, pmax, summary->processed_range);
  T.processed_min = pmin;
  T.processed_max = pmax;
  T.inverse_processed_span = 1.0 / (pmax - pmin);
  T.normalization_min = pmin;
  T.normalization_max = pmax;
  T.inverse_normalization_span = T.inverse_processed_span;
  summary->scalar_range[0] = summary->processed_range[0];
  summary->scalar_range[1] = summary->processed_range[1];
  if (normalize_to_data_range) {  // SetSceneNormalizationRange :427-443
    const double span = pmax - pmin;
    if (!(span > 0.0) || !std::isfinite(span))
      return fail(ctx, AMRVR_ERR_RUNTIME,
                  "Failed to establish a finite scalar range for color mapping.");
    T.normalize_to_unit_range = 1;
    T.normalization_min = pmin;
    T.normalization_max = pmax;
    T.inverse_normalization_span = 1.0 / span;
    summary->scalar_range[0] = 0.0f;
    summary->scalar_range[1] = 1.0f;
  }
  for (size_t b = 0; b < local.size(); ++b) boxes_out[b] = local[b];
  summary->nboxes = static_cast<int32_t>(local.size());
  return AMRVR_OK;
}

extern "C" int amrvr_scene_histogram(amrvr_ctx* ctx, int nboxes, const amrvr_box* boxes,
                                     const amrvr_scalar_transform* transform, float range_min,
                                     float range_max, int bin_count, uint64_t* counts_host,
                                     uint64_t* sample_count) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, bin_count > 0, "binCount must be positive");  // :448-450
  REQUIRE(ctx, nboxes >= 0 && (nboxes == 0 || boxes != nullptr) && transform != nullptr &&
                   counts_host != nullptr,
          "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  for (int b = 0; b < bin_count; ++b) counts_host[b] = 0;
  if (sample_count) *sample_count = 0;
  const float width = range_max - range_min;
  if (!(width > 0.0f) || !std::isfinite(width)) return AMRVR_OK;  // :469-471
  std::vector<StagedBox> staged;
  const int st = stage_boxes(ctx, nboxes, boxes, staged);
  if (st != AMRVR_OK) return st;
  const size_t bytes = static_cast<size_t>(bin_count) * sizeof(unsigned long long);
  CUDA_TRY(ctx, ctx->counter.reserve(std::max<size_t>(bytes, 256)));
  unsigned long long* counts = ctx->counter.as<unsigned long long>();
  CUDA_TRY(ctx, cudaMemsetAsync(counts, 0, bytes, ctx->stream));
  for (int b = 0; b < nboxes; ++b) {
    if (staged[b].dev == nullptr) continue;
    LAUNCH_TRY(ctx, launch_brick_histogram(staged[b].dev, staged[b].jstride, staged[b].kstride,
                                           boxes[b].dims[0], boxes[b].dims[1], boxes[b].dims[2],
                                           *transform, range_min, range_max, bin_count, counts,
                                           ctx->stream, &ctx->launches));
  }
  if (ctx->nranks > 1)  // the two MPI_Allreduce of :557-570 (the total is the sum of the bins)
    NCCL_TRY(ctx, nccl_api().AllReduce(counts, counts, static_cast<size_t>(bin_count), ncclUint64,
                                       ncclSum, ctx->comm, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(counts_host, counts, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  uint64_t total = 0;
  for (int b = 0; b < bin_count; ++b) total += counts_host[b];
  if (sample_count) *sample_count = total;
  return AMRVR_OK;
}

extern "C" int amrvr_scene_free(amrvr_ctx* ctx, amrvr_scene* scene) {
  if (scene == nullptr) return AMRVR_OK;
  if (ctx != nullptr) {
    cudaSetDevice(ctx->device);
    if (scene->pending) {
      cudaStreamSynchronize(ctx->copy);
      cudaStreamSynchronize(ctx->quant);
    }
    cudaStreamSynchronize(ctx->stream);
    async_upload_done(ctx, scene);  // waited for above if it was still pending
  }
  for (cudaEvent_t e : scene->group_ready) {
    if (ctx != nullptr && ctx->event_pool.size() < 64) ctx->event_pool.push_back(e);
    else cudaEventDestroy(e);
  }
  if (ctx != nullptr && scene->bricks.ptr != nullptr && ctx->brick_pools.size() < 2) {
    ctx->brick_pools.push_back(scene->bricks);  // cudaMalloc/cudaFree cost more than the quantise
    scene->bricks = DeviceBuffer{};
  }
  scene->bricks.release();
  for (cudaEvent_t e : {scene->up0, scene->up1}) {
    if (e == nullptr) continue;
    if (ctx != nullptr && ctx->event_pool.size() < 16) ctx->event_pool.push_back(e);
    else cudaEventDestroy(e);
  }
  delete scene;
  return AMRVR_OK;
}

extern "C" int amrvr_scene_global_boxes(const amrvr_scene* scene, int* nglobal,
                                        const amrvr_box_record** records) {
  if (scene == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "scene is NULL");
  if (nglobal) *nglobal = static_cast<int>(scene->global.size());
  if (records) *records = scene->global.data();
  return AMRVR_OK;
}

extern "C" int amrvr_scene_read_brick(amrvr_ctx* ctx, const amrvr_scene* scene, int local_box,
                                      uint8_t* dst) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene != nullptr && dst != nullptr, "NULL argument");
  REQUIRE(ctx, local_box >= 0 && local_box < static_cast<int>(scene->local.size()),
          "local_box out of range");
  const LocalBox& L = scene->local[local_box];
  const size_t cells = static_cast<size_t>(std::max(L.dims[0], 0)) * std::max(L.dims[1], 0) *
                       std::max(L.dims[2], 0);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (wait_upload(ctx, scene) != AMRVR_OK) return AMRVR_ERR_CUDA;
  CUDA_TRY(ctx, cudaMemcpyAsync(dst, scene->bricks.as<uint8_t>() + L.brick_off, cells,
                                cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// ordering
// ---------------------------------------------------------------------------

extern "C" int amrvr_order_boxes(amrvr_ctx* ctx, const amrvr_scene* scene,
                                 const amrvr_camera* camera, float aspect, int use_graph,
                                 int* rank_order, int* graph_failed) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene != nullptr && camera != nullptr && rank_order != nullptr, "NULL argument");
  return amrvr_order_records(ctx, ctx->nranks, static_cast<int>(scene->global.size()),
                             scene->global.data(), camera, aspect, use_graph, rank_order,
                             graph_failed);
}

extern "C" int amrvr_order_records(amrvr_ctx* ctx, int nranks, int nglobal,
                                   const amrvr_box_record* records, const amrvr_camera* camera,
                                   float aspect, int use_graph, int* rank_order,
                                   int* graph_failed) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, nranks >= 1 && nglobal >= 0 && (nglobal == 0 || records != nullptr) &&
                   camera != nullptr && rank_order != nullptr,
          "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (nranks == 1) {  // one piece, one owner: the graph cannot change the answer
    rank_order[0] = 0;
    if (graph_failed) *graph_failed = 0;
    return AMRVR_OK;
  }
  if (nglobal < ctx->order_device_min) {
    // a few hundred boxes: the pair test and the key sort take tens of
    // microseconds on the host, less than one kernel launch + readback
    const bool ok = order_ranks(nranks, nglobal, records, *camera, aspect, use_graph != 0,
                                *host_order_backend(), rank_order);
    if (graph_failed) *graph_failed = ok ? 0 : 1;
    return AMRVR_OK;
  }
  DeviceOrderBackend backend(ctx);
  const bool ok = order_ranks(nranks, nglobal, records, *camera, aspect, use_graph != 0, backend,
                              rank_order);
  if (backend.status != 0) return fail(ctx, backend.status, "device ordering failed");
  if (graph_failed) *graph_failed = ok ? 0 : 1;
  return AMRVR_OK;
}

extern "C" int amrvr_order_boxes_host(int nranks, int nglobal, const amrvr_box_record* records,
                                      const amrvr_camera* camera, float aspect, int use_graph,
                                      int* rank_order, int* graph_failed) {
  if (nranks < 1 || nglobal < 0 || (nglobal > 0 && records == nullptr) || camera == nullptr ||
      rank_order == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  const bool ok = order_ranks(nranks, nglobal, records, *camera, aspect, use_graph != 0,
                              *host_order_backend(), rank_order);
  if (graph_failed) *graph_failed = ok ? 0 : 1;
  return AMRVR_OK;
}

extern "C" int amrvr_layer_order_host(int nglobal, const amrvr_box_record* records,
                                      const amrvr_camera* camera, float* depth_hints, int* order) {
  if (nglobal < 0 || (nglobal > 0 && (records == nullptr || order == nullptr)) || camera == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  layer_order(nglobal, records, *camera, depth_hints, order);
  return AMRVR_OK;
}

extern "C" int amrvr_piece_range(int num_pixels, int piece, int num_pieces, int* begin, int* end) {
  if (num_pieces < 1 || piece < 0 || piece >= num_pieces || begin == nullptr || end == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  piece_range(num_pixels, piece, num_pieces, begin, end);
  return AMRVR_OK;
}

extern "C" int amrvr_reference_sample_distance(int nglobal, const amrvr_box_record* records,
                                               const double bounds_min[3],
                                               const double bounds_max[3], float* out) {
  if (nglobal < 0 || (nglobal > 0 && records == nullptr) || bounds_min == nullptr ||
      bounds_max == nullptr || out == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  *out = reference_sample_distance(nglobal, records, bounds_min, bounds_max);
  return AMRVR_OK;
}

extern "C" int amrvr_color_table(float alpha_scale, float normalization_factor, float range_min,
                                 float range_max, const amrvr_colormap_pt* pts, int npts,
                                 float* table_rgba) {
  if (table_rgba == nullptr || npts < 0 || (npts > 0 && pts == nullptr))
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  build_color_table(alpha_scale, normalization_factor, range_min, range_max, pts, npts, table_rgba);
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// ray-march
// ---------------------------------------------------------------------------

namespace {

// buildColorTable is a pure function of (alpha scale, normalisation factor,
// scalar range, control points); an animation re-uses the same few tables
// every frame, so they are memoised per context (most recent first, <= 16).
const float* cached_color_table(amrvr_ctx* ctx, float alpha_scale, float norm, float range_min,
                                float range_max, const amrvr_colormap_pt* pts, int npts) {
  std::vector<unsigned char> key(4 * sizeof(float) + static_cast<size_t>(npts) * sizeof(amrvr_colormap_pt));
  const float head[4] = {alpha_scale, norm, range_min, range_max};
  std::memcpy(key.data(), head, sizeof(head));
  if (npts > 0) std::memcpy(key.data() + sizeof(head), pts, static_cast<size_t>(npts) * sizeof(amrvr_colormap_pt));
  auto& cache = ctx->table_cache;
  for (size_t i = 0; i < cache.size(); ++i)
    if (cache[i].key == key) {
      if (i != 0) std::rotate(cache.begin(), cache.begin() + static_cast<long>(i), cache.begin() + static_cast<long>(i) + 1);
      return cache[0].rgba.data();
    }
  amrvr_ctx::TableEntry entry;
  entry.key = std::move(key);
  entry.rgba.resize(4 * AMRVR_COLOR_TABLE_SIZE);
  build_color_table(alpha_scale, norm, range_min, range_max, pts, npts, entry.rgba.data());
  cache.insert(cache.begin(), std::move(entry));
  if (cache.size() > 16) cache.pop_back();
  return cache[0].rgba.data();
}


}  // namespace

namespace {

// The per-box part of a MarchBox that does not depend on the frame: corners, cell
// sizes, exactness of the reciprocal, epsilon (VolumePainter.cpp:658-692).
void fill_march_const(const double* min_corner, const double* max_corner, const int32_t* dims,
                      uint64_t brick_off, MarchBox& M) {
  bool pow2 = true;
  for (int d = 0; d < 3; ++d) {
    M.lo[d] = static_cast<float>(min_corner[d]);
    M.hi[d] = static_cast<float>(max_corner[d]);
    M.n[d] = dims[d];
    M.cell[d] = (M.hi[d] - M.lo[d]) / static_cast<float>(dims[d]);  // VolumePainter.cpp:676-678
    M.inv_cell_tiny[d] = std::ldexp(1.0f / M.cell[d], -149);
    // 2^-149/cell must itself be a (subnormal) float: cell a power of two <= 1
    pow2 = pow2 && is_pow2_float(M.cell[d]) && M.inv_cell_tiny[d] > 0.0f &&
           std::ldexp(M.inv_cell_tiny[d], 149) == 1.0f / M.cell[d];
  }
  // :680-683 extent, float, x then y then z
  const float ex = M.hi[0] - M.lo[0], ey = M.hi[1] - M.lo[1], ez = M.hi[2] - M.lo[2];
  M.eps = std::sqrt(ex * ex + ey * ey + ez * ez) * 0.0001f;
  M.exact_recip = pow2 ? 1 : 0;
  M.brick_off = brick_off;
}

// Pinned staging block of the per-frame parameters: one host->device copy per
// frame.  The previous copy out of it must have completed before it is rewritten.
int param_staging(amrvr_ctx* ctx, size_t bytes, unsigned char** out) {
  if (ctx->param_copied == nullptr)
    CUDA_TRY(ctx, cudaEventCreateWithFlags(&ctx->param_copied, cudaEventDisableTiming));
  else
    CUDA_TRY(ctx, cudaEventSynchronize(ctx->param_copied));
  if (ctx->param_host_bytes < bytes) {
    if (ctx->param_host != nullptr) CUDA_TRY(ctx, cudaFreeHost(ctx->param_host));
    ctx->param_host = nullptr;
    ctx->param_host_bytes = 0;
    const size_t want = (bytes + 65535) & ~static_cast<size_t>(65535);
    CUDA_TRY(ctx, cudaMallocHost(reinterpret_cast<void**>(&ctx->param_host), want));
    ctx->param_host_bytes = want;
  }
  *out = ctx->param_host;
  return AMRVR_OK;
}

constexpr size_t kParamHeader = 64;  // u64 words: [samples | pushed bytes | stolen tiles | tile counter (u32) | CTA busy ns | ~first start | last end | table staging ns]

}  // namespace

extern "C" int amrvr_raymarch(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_camera* camera,
                              int width, int height, float box_transparency,
                              float reference_sample_distance_, const amrvr_colormap_pt* pts,
                              int npts, int count_samples, amrvr_layers** out) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, out != nullptr, "out is NULL");
  *out = nullptr;
  REQUIRE(ctx, scene != nullptr && camera != nullptr, "NULL argument");
  REQUIRE(ctx, width > 0 && height > 0, "width and height must be positive");
  REQUIRE(ctx, npts >= 0 && (npts == 0 || pts != nullptr), "bad colour map");
  REQUIRE(ctx, static_cast<int64_t>(width) * height < (1LL << 31), "frame too large");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  if (wait_upload(ctx, scene) != AMRVR_OK) return AMRVR_ERR_CUDA;

  const int n = static_cast<int>(scene->local.size());
  std::unique_ptr<amrvr_layers> L(new amrvr_layers);
  L->width = width;
  L->height = height;
  L->camera = *camera;
  L->rects.resize(n);
  L->offs.resize(n);
  L->counted = count_samples != 0;

  CameraBasis basis;
  make_camera_basis(*camera, basis);
  MarchFrame frame{};
  for (int d = 0; d < 3; ++d) {
    frame.eye[d] = basis.eye[d];
    frame.fwd[d] = basis.fwd[d];
    frame.right[d] = basis.right[d];
    frame.up[d] = basis.up[d];
  }
  frame.aspect = static_cast<float>(width) / static_cast<float>(std::max(height, 1));
  frame.tan_half = std::tan(camera->fov_y_degrees * 0.5f * 3.14159265358979323846f / 180.0f);
  frame.inv_w = 1.0f / static_cast<float>(width);
  frame.inv_h = 1.0f / static_cast<float>(height);
  frame.width = width;
  frame.height = height;
  frame.nboxes = n;

  L->global_rects.resize(scene->global.size());
  box_footprints(static_cast<int>(scene->global.size()), scene->global.data(), basis, *camera, width,
                 height, L->global_rects.data());
  const float alpha_scale = std::min(std::max(1.0f - box_transparency, 0.0f), 1.0f);
  if (scene->march_const.size() != static_cast<size_t>(n)) {
    std::vector<MarchBox>& mc = const_cast<amrvr_scene*>(scene)->march_const;
    mc.assign(static_cast<size_t>(n), MarchBox{});
    for (int b = 0; b < n; ++b) {
      const LocalBox& B = scene->local[b];
      fill_march_const(B.min_corner, B.max_corner, B.dims, B.brick_off, mc[b]);
    }
  }
  std::vector<MarchBox> mb(scene->march_const);
  std::vector<float> table_norm;  // one table per distinct normalisation factor
  std::vector<float> tables;
  uint64_t floats = 0;
  int tiles = 0;
  int64_t rays = 0;
  for (int b = 0; b < n; ++b) {
    const LocalBox& B = scene->local[b];
    MarchBox& M = mb[b];
    float step, norm;
    box_sampling(B.min_corner, B.max_corner, B.dims, reference_sample_distance_, &step, &norm);
    int which = -1;
    for (size_t t = 0; t < table_norm.size(); ++t)
      if (std::memcmp(&table_norm[t], &norm, sizeof(float)) == 0) which = static_cast<int>(t);
    if (which < 0) {
      which = static_cast<int>(table_norm.size());
      table_norm.push_back(norm);
      tables.resize(tables.size() + 4 * AMRVR_COLOR_TABLE_SIZE);
      std::memcpy(tables.data() + static_cast<size_t>(which) * 4 * AMRVR_COLOR_TABLE_SIZE,
                  cached_color_table(ctx, alpha_scale, norm, scene->range_min, scene->range_max, pts,
                                     npts),
                  4 * AMRVR_COLOR_TABLE_SIZE * sizeof(float));
    }
    M.table = which;
    M.step = step;
    M.margin = inside_margin(M, basis.eye);
    const Rect r = L->global_rects[static_cast<size_t>(scene->first_global + b)];
    L->rects[b] = r;
    M.rx0 = r.x0;
    M.ry0 = r.y0;
    M.rw = r.w;
    M.rh = r.h;
    M.ry_first = r.y0;
    M.tiles_x = (r.w + 15) / 16;
    M.tile_base = tiles;
    tiles += M.tiles_x * ((r.h + 15) / 16);
    M.layer_off = floats;
    L->offs[b] = floats;
    floats += static_cast<uint64_t>(r.w) * r.h * 5;
    rays += static_cast<int64_t>(r.w) * r.h;
  }
  frame.ntiles = tiles;
  frame.strip_stride = 16;
  L->pool_floats = floats;
  L->stats.rays = rays;
  L->stats.fragment_pixels = rays;

  // layer pool: a recycled one when it is large enough
  const size_t need = std::max<size_t>(floats * sizeof(float), 256);
  for (size_t i = 0; L->pool.ptr == nullptr && i < ctx->layer_pools.size(); ++i)
    if (ctx->layer_pools[i].bytes >= need) {
      L->pool = ctx->layer_pools[i];
      ctx->layer_pools.erase(ctx->layer_pools.begin() + static_cast<long>(i));
      break;
    }
  if (L->pool.ptr == nullptr) CUDA_TRY(ctx, L->pool.reserve(need));

  // one parameter block, one host->device copy: [header | tables | boxes]
  const size_t tables_bytes = tables.size() * sizeof(float);
  const size_t boxes_bytes = mb.size() * sizeof(MarchBox);
  const size_t block_bytes = kParamHeader + tables_bytes + boxes_bytes;
  unsigned char* block = nullptr;
  {
    const int st = param_staging(ctx, block_bytes, &block);
    if (st != AMRVR_OK) return st;
  }
  std::memset(block, 0, kParamHeader);  // counters start at zero
  if (tables_bytes) std::memcpy(block + kParamHeader, tables.data(), tables_bytes);
  if (boxes_bytes) std::memcpy(block + kParamHeader + tables_bytes, mb.data(), boxes_bytes);
  CUDA_TRY(ctx, ctx->march_boxes.reserve(block_bytes));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->march_boxes.ptr, block, block_bytes, cudaMemcpyHostToDevice,
                                ctx->stream));
  CUDA_TRY(ctx, cudaEventRecord(ctx->param_copied, ctx->stream));
  unsigned char* dev_block = ctx->march_boxes.as<unsigned char>();
  unsigned long long* counter = reinterpret_cast<unsigned long long*>(dev_block);
  MarchQueues queues{};  // one queue, one pool: this rank's own boxes
  queues.nqueues = 1;
  queues.tile_begin[1] = tiles;
  queues.counter[0] = reinterpret_cast<unsigned*>(dev_block + 24);
  queues.pool[0] = L->pool.as<float>();
  queues.stats = counter;
  CUDA_TRY(ctx, take_event(ctx, &L->t0));
  CUDA_TRY(ctx, take_event(ctx, &L->t1));
  CUDA_TRY(ctx, cudaEventRecord(L->t0, ctx->stream));
  if (n > 0)
    LAUNCH_TRY(ctx, launch_raymarch(frame, queues,
                                    reinterpret_cast<const MarchBox*>(dev_block + kParamHeader + tables_bytes),
                                    scene->bricks.as<uint8_t>(),
                                    reinterpret_cast<const float*>(dev_block + kParamHeader),
                                    count_samples != 0, ctx->stream, &ctx->launches));
  CUDA_TRY(ctx, cudaEventRecord(L->t1, ctx->stream));
  if (count_samples) {
    unsigned long long host = 0;
    CUDA_TRY(ctx, cudaMemcpyAsync(&host, counter, 8, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    L->stats.samples = static_cast<int64_t>(host);
  }
  *out = L.release();
  return AMRVR_OK;
}

extern "C" int amrvr_layers_free(amrvr_ctx* ctx, amrvr_layers* layers) {
  if (layers == nullptr) return AMRVR_OK;
  if (layers->pool_shared) layers->pool = DeviceBuffer{};  // not owned
  if (ctx != nullptr) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    if (layers->pool.ptr != nullptr && ctx->layer_pools.size() < 4) {
      ctx->layer_pools.push_back(layers->pool);
      layers->pool = DeviceBuffer{};
    }
  }
  layers->pool.release();
  for (cudaEvent_t e : {layers->t0, layers->t1}) {
    if (e == nullptr) continue;
    if (ctx != nullptr && ctx->event_pool.size() < 16) ctx->event_pool.push_back(e);
    else cudaEventDestroy(e);
  }
  delete layers;
  return AMRVR_OK;
}

extern "C" int amrvr_layers_count(const amrvr_layers* layers) {
  return layers ? static_cast<int>(layers->rects.size()) : 0;
}

extern "C" int amrvr_layers_stats(const amrvr_layers* layers, amrvr_march_stats* stats) {
  if (layers == nullptr || stats == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "NULL argument");
  amrvr_layers* L = const_cast<amrvr_layers*>(layers);
  if (!L->stats_ready) {
    if (cudaEventSynchronize(L->t1) != cudaSuccess ||
        cudaEventElapsedTime(&L->stats.kernel_ms, L->t0, L->t1) != cudaSuccess)
      return fail(nullptr, AMRVR_ERR_CUDA, "event timing failed");
    L->stats_ready = true;
  }
  *stats = L->stats;
  return AMRVR_OK;
}

extern "C" int amrvr_layers_rect(const amrvr_layers* layers, int layer, int* x0, int* y0, int* w,
                                 int* h) {
  if (layers == nullptr || layer < 0 || layer >= static_cast<int>(layers->rects.size()))
    return fail(nullptr, AMRVR_ERR_INVALID, "layer out of range");
  const Rect& r = layers->rects[layer];
  if (x0) *x0 = r.x0;
  if (y0) *y0 = r.y0;
  if (w) *w = r.w;
  if (h) *h = r.h;
  return AMRVR_OK;
}

extern "C" int amrvr_layers_read(amrvr_ctx* ctx, const amrvr_layers* layers, int layer,
                                 float* dst5) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, layers != nullptr && dst5 != nullptr, "NULL argument");
  REQUIRE(ctx, layer >= 0 && layer < static_cast<int>(layers->rects.size()), "layer out of range");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const Rect& r = layers->rects[layer];
  const int W = layers->width, H = layers->height;
  const size_t npix = static_cast<size_t>(W) * H;
  for (size_t p = 0; p < npix; ++p) {  // ImageRGBAFloatColorDepthSort::clear
    float* o = dst5 + 5 * p;
    o[0] = o[1] = o[2] = o[3] = 0.0f;
    o[4] = INFINITY;
  }
  const size_t count = static_cast<size_t>(r.w) * r.h * 5;
  if (count == 0) return AMRVR_OK;
  std::vector<float> host(count);
  CUDA_TRY(ctx, cudaMemcpyAsync(host.data(), layers->pool.as<float>() + layers->offs[layer],
                                count * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  for (int y = 0; y < r.h; ++y) {
    const float* row = host.data() + static_cast<size_t>(y) * r.w * 5;
    for (int x = 0; x < r.w; ++x) {
      const int px = r.x0 + x, py = r.y0 + y;
      if (px >= W || py >= H) continue;
      float* o = dst5 + 5 * (static_cast<size_t>(py) * W + px);
      o[0] = row[4 * x + 0];
      o[1] = row[4 * x + 1];
      o[2] = row[4 * x + 2];
      o[3] = row[4 * x + 3];
      o[4] = row[4 * r.w + x];
    }
  }
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// composite
// ---------------------------------------------------------------------------
namespace {

// piece_dst != nullptr: write the piece there instead of ctx->piece (single-rank
// frames fold straight into the frame buffer)
int composite_impl(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_layers* layers,
                   const amrvr_camera* camera, const int* rank_order, int* region_begin,
                   int* region_end, float** piece_out, int64_t* exchanged_bytes,
                   float* piece_dst = nullptr, const FoldBytes* bytes = nullptr) {
  const int P = ctx->nranks, me = ctx->rank;
  const int W = layers->width, H = layers->height;
  {
    std::vector<char> seen(static_cast<size_t>(P), 0);
    for (int i = 0; i < P; ++i) {
      REQUIRE(ctx, rank_order[i] >= 0 && rank_order[i] < P && !seen[rank_order[i]],
              "rank_order is not a permutation");
      seen[rank_order[i]] = 1;
    }
  }
  FramePlan plan;
  const bool same_camera = std::memcmp(&layers->camera, camera, sizeof(amrvr_camera)) == 0;
  make_frame_plan(*scene, *camera, W, H, P, rank_order, plan,
                  same_camera ? &layers->global_rects : nullptr);
  const int n = static_cast<int>(scene->global.size());
  REQUIRE(ctx, static_cast<int>(layers->rects.size()) == static_cast<int>(scene->local.size()),
          "layers do not belong to this scene");
  for (size_t l = 0; l < layers->rects.size(); ++l) {
    const Rect& a = layers->rects[l];
    const Rect& b = plan.rects[scene->first_global + static_cast<int>(l)];
    if (a.x0 != b.x0 || a.y0 != b.y0 || a.w != b.w || a.h != b.h)
      return fail(ctx, AMRVR_ERR_INVALID, "camera differs from the one the layers were marched with");
  }
  const int my_begin = plan.begin[me], my_end = plan.end[me];

  // exchange plan (P > 1): slices of my layers other pieces need, slices of
  // other ranks' layers my piece needs
  std::vector<uint64_t> send_off(P + 1, 0), recv_off(P + 1, 0);
  std::vector<CopyJob> jobs;
  std::vector<uint64_t> recv_slot(static_cast<size_t>(n), 0);  // float offset in recv_buf per global box
  // With peer access the slices are not packed and sent: one kernel reads them out of the layer pool and
  // stores them straight into the destination rank's receive pool over NVLink (fused pack + push), between
  // two flag barriers.  AMRVR_EXCHANGE=nccl (no peer mapping) keeps pack + grouped ncclSend/ncclRecv.
  const bool push = P > 1 && ctx->peer_ok;
  float* recv_base = nullptr;
  if (P > 1) {
    std::vector<uint64_t> send_len(P, 0), recv_len(P, 0);
    // len[q * P + o]: floats rank q receives from rank o (every rank derives the whole matrix)
    std::vector<uint64_t> len(static_cast<size_t>(P) * P, 0);
    for (int g = 0; g < n; ++g) {
      const int owner = scene->global[g].owner;
      const Rect& r = plan.rects[g];
      int s0, s1;
      for (int q = 0; q < P; ++q) {
        if (q == owner || !slice_rows(r, plan.begin[q], plan.end[q], W, &s0, &s1)) continue;
        len[static_cast<size_t>(q) * P + owner] += static_cast<uint64_t>(s1 - s0) * r.w * 5;
      }
    }
    uint64_t max_recv = 0;
    std::vector<uint64_t> my_off_at(P, 0);  // where my slices start in rank q's receive pool
    for (int q = 0; q < P; ++q) {
      uint64_t total = 0;
      for (int o = 0; o < P; ++o) {
        if (o == me) my_off_at[q] = total;
        total += len[static_cast<size_t>(q) * P + o];
      }
      max_recv = std::max(max_recv, total);
      send_len[q] = len[static_cast<size_t>(q) * P + me];
      recv_len[q] = len[static_cast<size_t>(me) * P + q];
    }
    for (int q = 0; q < P; ++q) {
      send_off[q + 1] = send_off[q] + send_len[q];
      recv_off[q + 1] = recv_off[q] + recv_len[q];
    }
    std::vector<uint64_t> send_cur(send_off.begin(), send_off.end() - 1);
    std::vector<uint64_t> recv_cur(recv_off.begin(), recv_off.end() - 1);
    const uint32_t kChunk = 16384;  // floats per copy job
    for (int g = 0; g < n; ++g) {
      const int owner = scene->global[g].owner;
      const Rect& r = plan.rects[g];
      int s0, s1;
      if (owner == me) {
        const uint64_t base = layers->offs[g - scene->first_global];
        for (int q = 0; q < P; ++q) {
          if (q == me || !slice_rows(r, plan.begin[q], plan.end[q], W, &s0, &s1)) continue;
          uint64_t src = base + static_cast<uint64_t>(s0 - r.y0) * r.w * 5;
          uint64_t left = static_cast<uint64_t>(s1 - s0) * r.w * 5;
          while (left > 0) {
            const uint32_t c = static_cast<uint32_t>(std::min<uint64_t>(left, kChunk));
            jobs.push_back(push ? CopyJob{src, my_off_at[q] + (send_cur[q] - send_off[q]), c,
                                          static_cast<uint32_t>(q)}
                                : CopyJob{src, send_cur[q], c, 0});
            src += c;
            send_cur[q] += c;
            left -= c;
          }
        }
      } else if (slice_rows(r, my_begin, my_end, W, &s0, &s1)) {
        recv_slot[g] = recv_cur[owner];
        recv_cur[owner] += static_cast<uint64_t>(s1 - s0) * r.w * 5;
      }
    }
    if (!jobs.empty()) {
      CUDA_TRY(ctx, ctx->copy_jobs.reserve(jobs.size() * sizeof(CopyJob)));
      CUDA_TRY(ctx, cudaMemcpyAsync(ctx->copy_jobs.ptr, jobs.data(), jobs.size() * sizeof(CopyJob),
                                    cudaMemcpyHostToDevice, ctx->stream));
    }
    if (push) {
      // the pool is as large as the largest receiver needs, so every rank grows it in the same call
      bool mapped = true;
      const int st = sym_reserve(ctx, ctx->sym_recv, std::max<uint64_t>(max_recv, 4) * sizeof(float), &mapped);
      if (st != AMRVR_OK) return st;
      if (!mapped) return fail(ctx, AMRVR_ERR_CUDA, "peer mapping of the receive pool failed");
      recv_base = static_cast<float*>(ctx->sym_recv.local);
      for (auto& e : ctx->seam_ev)
        if (e == nullptr) CUDA_TRY(ctx, cudaEventCreate(&e));
      CUDA_TRY(ctx, cudaEventRecord(ctx->seam_ev[0], ctx->stream));
      // nobody stores into a pool its owner may still be folding from (the previous composite)
      int bst = peer_barrier(ctx, 3);
      if (bst != AMRVR_OK) return bst;
      CUDA_TRY(ctx, cudaEventRecord(ctx->seam_ev[1], ctx->stream));
      LAUNCH_TRY(ctx, launch_push(ctx->copy_jobs.as<CopyJob>(), static_cast<int>(jobs.size()),
                                  layers->pool.as<float>(), peer_ptrs(ctx->sym_recv), ctx->stream,
                                  &ctx->launches));
      CUDA_TRY(ctx, cudaEventRecord(ctx->seam_ev[2], ctx->stream));
      // every peer's slices have arrived (the barrier kernel fences before it raises its flags)
      bst = peer_barrier(ctx, 3);
      if (bst != AMRVR_OK) return bst;
      CUDA_TRY(ctx, cudaEventRecord(ctx->seam_ev[3], ctx->stream));
      ctx->seam_timed = true;
    } else {
      CUDA_TRY(ctx, ctx->send_buf.reserve(std::max<uint64_t>(send_off[P], 4) * sizeof(float)));
      CUDA_TRY(ctx, ctx->recv_buf.reserve(std::max<uint64_t>(recv_off[P], 4) * sizeof(float)));
      recv_base = ctx->recv_buf.as<float>();
      LAUNCH_TRY(ctx, launch_pack(ctx->copy_jobs.as<CopyJob>(), static_cast<int>(jobs.size()),
                                  layers->pool.as<float>(), ctx->send_buf.as<float>(), ctx->stream,
                                  &ctx->launches));
      NcclApi& nc = nccl_api();
      NCCL_TRY(ctx, nc.GroupStart());
      for (int q = 0; q < P; ++q) {
        if (q == me) continue;
        if (send_len[q] > 0)
          NCCL_TRY(ctx, nc.Send(ctx->send_buf.as<float>() + send_off[q], send_len[q], ncclFloat, q,
                                ctx->comm, ctx->stream));
        if (recv_len[q] > 0)
          NCCL_TRY(ctx, nc.Recv(ctx->recv_buf.as<float>() + recv_off[q], recv_len[q], ncclFloat, q,
                                ctx->comm, ctx->stream));
      }
      NCCL_TRY(ctx, nc.GroupEnd());
    }
    if (exchanged_bytes) *exchanged_bytes = static_cast<int64_t>(send_off[P] * sizeof(float));
  } else if (exchanged_bytes) {
    *exchanged_bytes = 0;
  }

  // fold table in blend order, only layers that reach my piece
  std::vector<FoldLayer> fold;
  fold.reserve(static_cast<size_t>(n));
  for (int pos = 0; pos < n; ++pos) {
    const int g = plan.order[pos];
    const Rect& r = plan.rects[g];
    int s0, s1;
    if (!slice_rows(r, my_begin, my_end, W, &s0, &s1)) continue;
    FoldLayer F{};
    if (scene->global[g].owner == me) {
      F.base = layers->pool.as<float>() + layers->offs[g - scene->first_global] +
               static_cast<uint64_t>(s0 - r.y0) * r.w * 5;
    } else {
      F.base = recv_base + recv_slot[g];
    }
    F.x0 = r.x0;
    F.w = r.w;
    F.row0 = s0;
    F.row1 = s1;
    F.run = plan.run_of[g];
    fold.push_back(F);
  }
  const int npix = my_end - my_begin;
  if (piece_dst == nullptr) {
    CUDA_TRY(ctx, ctx->piece.reserve(std::max<size_t>(static_cast<size_t>(std::max(npix, 0)) * 20, 256)));
    piece_dst = ctx->piece.as<float>();
  }
  if (!fold.empty()) {
    CUDA_TRY(ctx, ctx->fold_layers.reserve(fold.size() * sizeof(FoldLayer)));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->fold_layers.ptr, fold.data(), fold.size() * sizeof(FoldLayer),
                                  cudaMemcpyHostToDevice, ctx->stream));
  }
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev[6], ctx->stream));
  CUDA_TRY(ctx, ctx->fold_bins.reserve(std::max<size_t>(
                    fold_scratch_bytes(static_cast<int>(fold.size()), W, my_begin, my_end), 256)));
  LAUNCH_TRY(ctx, launch_fold_bin(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), W,
                                  my_begin, my_end, ctx->fold_bins.ptr, ctx->stream, &ctx->launches));
  LAUNCH_TRY(ctx, launch_fold(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), W,
                              my_begin, my_end, piece_dst, ctx->fold_bins.ptr, ctx->stream, &ctx->launches,
                              bytes));
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev[7], ctx->stream));
  if (region_begin) *region_begin = my_begin;
  if (region_end) *region_end = my_end;
  if (piece_out) *piece_out = piece_dst;
  return AMRVR_OK;
}

}  // namespace

extern "C" int amrvr_ctx_last_composite_ms(amrvr_ctx* ctx, float* ms5) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, ms5 != nullptr, "ms5 is NULL");
  REQUIRE(ctx, ctx->seam_timed, "no composite has pushed slices to peers on this ctx yet");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  const cudaEvent_t mark[6] = {ctx->seam_ev[0], ctx->seam_ev[1], ctx->seam_ev[2], ctx->seam_ev[3],
                               ctx->ev[6], ctx->ev[7]};
  for (int i = 0; i < 5; ++i) CUDA_TRY(ctx, cudaEventElapsedTime(&ms5[i], mark[i], mark[i + 1]));
  return AMRVR_OK;
}

extern "C" int amrvr_composite(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_layers* layers,
                               const amrvr_camera* camera, const int* rank_order,
                               int* region_begin, int* region_end, float** piece5_device) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene && layers && camera && rank_order, "NULL argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  return composite_impl(ctx, scene, layers, camera, rank_order, region_begin, region_end,
                        piece5_device, nullptr);
}

extern "C" int amrvr_exchange_plan_host(int nranks, int rank, int nglobal,
                                        const amrvr_box_record* records,
                                        const amrvr_camera* camera, int width, int height,
                                        const int* rank_order, int64_t* send_floats,
                                        int64_t* recv_floats, int* region_begin, int* region_end) {
  if (nranks < 1 || rank < 0 || rank >= nranks || nglobal < 0 ||
      (nglobal > 0 && records == nullptr) || camera == nullptr || width <= 0 || height <= 0 ||
      rank_order == nullptr || send_floats == nullptr || recv_floats == nullptr)
    return fail(nullptr, AMRVR_ERR_INVALID, "bad argument");
  std::vector<char> seen(static_cast<size_t>(nranks), 0);
  for (int i = 0; i < nranks; ++i) {
    if (rank_order[i] < 0 || rank_order[i] >= nranks || seen[rank_order[i]])
      return fail(nullptr, AMRVR_ERR_INVALID, "rank_order is not a permutation");
    seen[rank_order[i]] = 1;
  }
  amrvr_scene shadow;
  shadow.global.assign(records, records + nglobal);
  FramePlan plan;
  make_frame_plan(shadow, *camera, width, height, nranks, rank_order, plan);
  for (int q = 0; q < nranks; ++q) send_floats[q] = recv_floats[q] = 0;
  for (int g = 0; g < nglobal; ++g) {
    const int owner = records[g].owner;
    const Rect& r = plan.rects[g];
    int s0, s1;
    if (owner == rank) {
      for (int q = 0; q < nranks; ++q)
        if (q != rank && slice_rows(r, plan.begin[q], plan.end[q], width, &s0, &s1))
          send_floats[q] += static_cast<int64_t>(s1 - s0) * r.w * 5;
    } else if (owner >= 0 && owner < nranks &&
               slice_rows(r, plan.begin[rank], plan.end[rank], width, &s0, &s1)) {
      recv_floats[owner] += static_cast<int64_t>(s1 - s0) * r.w * 5;
    }
  }
  if (region_begin) *region_begin = plan.begin[rank];
  if (region_end) *region_end = plan.end[rank];
  return AMRVR_OK;
}

// The reference's own calling convention for the composite: every layer is a
// dense W*H*5-float host image (ImageRGBAFloatColorDepthSort) with a depth hint
// (LayeredVolumeImage).  Used by the C++ adapter of DirectSendBase::compose;
// the fast path (amrvr_composite) keeps layers on the device as footprints.
extern "C" int amrvr_composite_dense(amrvr_ctx* ctx, int width, int height, int nlayers,
                                     const float* const* layers5_host, const float* depth_hints,
                                     const int* rank_order, int* region_begin, int* region_end,
                                     float* piece5_host) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, width > 0 && height > 0 && nlayers >= 0, "bad image size or layer count");
  REQUIRE(ctx, static_cast<int64_t>(width) * height < (1LL << 31), "frame too large");
  REQUIRE(ctx, nlayers == 0 || (layers5_host != nullptr && depth_hints != nullptr), "NULL layers");
  REQUIRE(ctx, rank_order != nullptr, "rank_order is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int P = ctx->nranks, me = ctx->rank;
  const int npix = width * height;
  std::vector<int> begin(P), end(P);
  {
    std::vector<char> seen(static_cast<size_t>(P), 0);
    for (int piece = 0; piece < P; ++piece) {
      const int r = rank_order[piece];
      REQUIRE(ctx, r >= 0 && r < P && !seen[r], "rank_order is not a permutation");
      seen[r] = 1;
      piece_range(npix, piece, P, &begin[r], &end[r]);
    }
  }
  // DirectSendBase.cpp:329-361: every rank learns every layer count and hint
  std::vector<int> counts(P, 0);
  std::vector<float> hints;  // rank-major
  counts[me] = nlayers;
  if (P == 1) {
    hints.assign(depth_hints, depth_hints + nlayers);
  } else {
    NcclApi& nc = nccl_api();
    DeviceBuffer tmp;
    CUDA_TRY(ctx, tmp.reserve(static_cast<size_t>(P + 1) * sizeof(int)));
    int* cd = tmp.as<int>();
    CUDA_TRY(ctx, cudaMemcpyAsync(cd + P, &nlayers, sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    NCCL_TRY(ctx, nc.AllGather(cd + P, cd, sizeof(int), ncclChar, ctx->comm, ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(counts.data(), cd, sizeof(int) * P, cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    const int most = std::max(1, *std::max_element(counts.begin(), counts.end()));
    DeviceBuffer hb;
    CUDA_TRY(ctx, hb.reserve(static_cast<size_t>(most) * (P + 1) * sizeof(float)));
    float* hd = hb.as<float>();
    if (nlayers > 0)
      CUDA_TRY(ctx, cudaMemcpyAsync(hd + static_cast<size_t>(most) * P, depth_hints,
                                    sizeof(float) * nlayers, cudaMemcpyHostToDevice, ctx->stream));
    NCCL_TRY(ctx, nc.AllGather(hd + static_cast<size_t>(most) * P, hd, sizeof(float) * most, ncclChar,
                               ctx->comm, ctx->stream));
    std::vector<float> all(static_cast<size_t>(most) * P);
    CUDA_TRY(ctx, cudaMemcpyAsync(all.data(), hd, all.size() * sizeof(float), cudaMemcpyDeviceToHost,
                                  ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    for (int r = 0; r < P; ++r)
      hints.insert(hints.end(), all.begin() + static_cast<size_t>(most) * r,
                   all.begin() + static_cast<size_t>(most) * r + counts[r]);
    tmp.release();
    hb.release();
  }
  struct Entry {
    float depth;
    int owner, local;
  };
  std::vector<Entry> order;
  for (int r = 0, k = 0; r < P; ++r)
    for (int l = 0; l < counts[r]; ++l, ++k) order.push_back(Entry{hints[k], r, l});
  std::sort(order.begin(), order.end(), [](const Entry& a, const Entry& b) {  // :378-388
    if (a.depth == b.depth) {
      if (a.owner == b.owner) return a.local < b.local;
      return a.owner < b.owner;
    }
    return a.depth < b.depth;
  });

  // my layers go to the device whole (interleaved, as they are); each piece's
  // pixel range is then a contiguous float range of every layer
  const size_t layer_floats = static_cast<size_t>(npix) * 5;
  CUDA_TRY(ctx, ctx->send_buf.reserve(std::max<size_t>(layer_floats * nlayers, 4) * sizeof(float)));
  for (int l = 0; l < nlayers; ++l) {
    REQUIRE(ctx, layers5_host[l] != nullptr, "NULL layer");
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->send_buf.as<float>() + layer_floats * l, layers5_host[l],
                                  layer_floats * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
  }
  const size_t mine = static_cast<size_t>(end[me] - begin[me]) * 5;
  std::vector<size_t> recv_base(P + 1, 0);  // float offset of rank r's first layer slice
  for (int r = 0; r < P; ++r) recv_base[r + 1] = recv_base[r] + (r == me ? 0 : mine * counts[r]);
  CUDA_TRY(ctx, ctx->recv_buf.reserve(std::max<size_t>(recv_base[P], 4) * sizeof(float)));
  if (P > 1) {
    NcclApi& nc = nccl_api();
    NCCL_TRY(ctx, nc.GroupStart());
    for (int q = 0; q < P; ++q) {
      if (q == me) continue;
      const size_t theirs = static_cast<size_t>(end[q] - begin[q]) * 5;
      for (int l = 0; l < nlayers && theirs > 0; ++l)
        NCCL_TRY(ctx, nc.Send(ctx->send_buf.as<float>() + layer_floats * l + static_cast<size_t>(begin[q]) * 5,
                              theirs, ncclFloat, q, ctx->comm, ctx->stream));
      for (int l = 0; l < counts[q] && mine > 0; ++l)
        NCCL_TRY(ctx, nc.Recv(ctx->recv_buf.as<float>() + recv_base[q] + mine * l, mine, ncclFloat, q,
                              ctx->comm, ctx->stream));
    }
    NCCL_TRY(ctx, nc.GroupEnd());
  }
  std::vector<FoldLayer> fold;
  int run = -1, prev_owner = -1;
  const int row0 = begin[me] / width;
  const int64_t lead = static_cast<int64_t>(begin[me]) - static_cast<int64_t>(row0) * width;  // pixels
  for (size_t pos = 0; pos < order.size(); ++pos) {
    const Entry& e = order[pos];
    if (pos == 0 || e.owner != prev_owner) ++run;  // :400-407
    prev_owner = e.owner;
    if (mine == 0) continue;
    FoldLayer F{};
    const float* slice = (e.owner == me)
                             ? ctx->send_buf.as<float>() + layer_floats * e.local +
                                   static_cast<size_t>(begin[me]) * 5
                             : ctx->recv_buf.as<float>() + recv_base[e.owner] + mine * e.local;
    F.base = slice - lead * 5;  // pixel (0,row0) of a virtual full-width slice
    F.x0 = 0;
    F.w = width;
    F.row0 = row0;
    F.row1 = (end[me] - 1) / width + 1;
    F.run = run;
    F.interleaved = 1;
    fold.push_back(F);
  }
  CUDA_TRY(ctx, ctx->piece.reserve(std::max<size_t>(mine * sizeof(float), 256)));
  if (!fold.empty()) {
    CUDA_TRY(ctx, ctx->fold_layers.reserve(fold.size() * sizeof(FoldLayer)));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->fold_layers.ptr, fold.data(), fold.size() * sizeof(FoldLayer),
                                  cudaMemcpyHostToDevice, ctx->stream));
  }
  CUDA_TRY(ctx, ctx->fold_bins.reserve(std::max<size_t>(
                    fold_scratch_bytes(static_cast<int>(fold.size()), width, begin[me], end[me]), 256)));
  LAUNCH_TRY(ctx, launch_fold_bin(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), width,
                                  begin[me], end[me], ctx->fold_bins.ptr, ctx->stream, &ctx->launches));
  LAUNCH_TRY(ctx, launch_fold(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), width,
                              begin[me], end[me], ctx->piece.as<float>(), ctx->fold_bins.ptr, ctx->stream, &ctx->launches));
  if (piece5_host != nullptr && mine > 0)
    CUDA_TRY(ctx, cudaMemcpyAsync(piece5_host, ctx->piece.ptr, mine * sizeof(float),
                                  cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (region_begin) *region_begin = begin[me];
  if (region_end) *region_end = end[me];
  return AMRVR_OK;
}

extern "C" int amrvr_composite_virtual(amrvr_ctx* ctx, const amrvr_scene* scene,
                                       const amrvr_layers* layers, const amrvr_camera* camera,
                                       int nvirtual, const int* box_owner, const int* rank_order,
                                       float* full5_device, int* piece_owner_host) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene && layers && camera && box_owner && rank_order && full5_device,
          "NULL argument");
  REQUIRE(ctx, ctx->nranks == 1, "virtual composite needs a single-rank context");
  REQUIRE(ctx, nvirtual >= 1, "nvirtual must be positive");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const int n = static_cast<int>(scene->local.size());
  REQUIRE(ctx, static_cast<int>(layers->rects.size()) == n, "layers do not belong to this scene");
  const int W = layers->width, H = layers->height;
  // the global table P ranks would have allgathered: owner-major, local order kept
  amrvr_scene shadow;
  std::vector<int> source;  // shadow index -> local layer
  for (int r = 0; r < nvirtual; ++r) {
    int local = 0;
    for (int l = 0; l < n; ++l) {
      REQUIRE(ctx, box_owner[l] >= 0 && box_owner[l] < nvirtual, "box_owner out of range");
      if (box_owner[l] != r) continue;
      amrvr_box_record rec = scene->global[scene->first_global + l];
      rec.owner = r;
      rec.local_index = local++;
      shadow.global.push_back(rec);
      source.push_back(l);
    }
  }
  FramePlan plan;
  make_frame_plan(shadow, *camera, W, H, nvirtual, rank_order, plan);
  std::vector<FoldLayer> fold;
  for (int pos = 0; pos < n; ++pos) {
    const int g = plan.order[pos];
    const Rect& r = plan.rects[g];
    if (r.w <= 0 || r.h <= 0) continue;
    FoldLayer F{};
    F.base = layers->pool.as<float>() + layers->offs[source[g]];
    F.x0 = r.x0;
    F.w = r.w;
    F.row0 = r.y0;
    F.row1 = r.y0 + r.h;
    F.run = plan.run_of[g];
    fold.push_back(F);
  }
  if (!fold.empty()) {
    CUDA_TRY(ctx, ctx->fold_layers.reserve(fold.size() * sizeof(FoldLayer)));
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->fold_layers.ptr, fold.data(), fold.size() * sizeof(FoldLayer),
                                  cudaMemcpyHostToDevice, ctx->stream));
  }
  CUDA_TRY(ctx, ctx->fold_bins.reserve(std::max<size_t>(
                    fold_scratch_bytes(static_cast<int>(fold.size()), W, 0, W * H), 256)));
  LAUNCH_TRY(ctx, launch_fold_bin(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), W, 0,
                                  W * H, ctx->fold_bins.ptr, ctx->stream, &ctx->launches));
  LAUNCH_TRY(ctx, launch_fold(ctx->fold_layers.as<FoldLayer>(), static_cast<int>(fold.size()), W,
                              0, W * H, full5_device, ctx->fold_bins.ptr, ctx->stream, &ctx->launches));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (piece_owner_host != nullptr)
    for (int r = 0; r < nvirtual; ++r)
      for (int p = plan.begin[r]; p < plan.end[r]; ++p) piece_owner_host[p] = r;
  return AMRVR_OK;
}

namespace {

// spans[2r], spans[2r+1] = pixel range of rank r's piece
int gather_impl(amrvr_ctx* ctx, const float* piece5_device, const std::vector<int>& spans, int root,
                float* full5_device) {
  const int P = ctx->nranks, me = ctx->rank;
  const int region_begin = spans[2 * me], region_end = spans[2 * me + 1];
  const size_t mine = static_cast<size_t>(region_end - region_begin) * 5;
  if (P > 1) {
    NcclApi& nc = nccl_api();
    NCCL_TRY(ctx, nc.GroupStart());
    if (me == root) {
      for (int r = 0; r < P; ++r) {
        const size_t len = static_cast<size_t>(spans[2 * r + 1] - spans[2 * r]) * 5;
        if (r == me || len == 0) continue;
        NCCL_TRY(ctx, nc.Recv(full5_device + static_cast<size_t>(spans[2 * r]) * 5, len, ncclFloat, r,
                              ctx->comm, ctx->stream));
      }
    } else if (mine > 0) {
      NCCL_TRY(ctx, nc.Send(piece5_device, mine, ncclFloat, root, ctx->comm, ctx->stream));
    }
    NCCL_TRY(ctx, nc.GroupEnd());
  }
  if (me == root && mine > 0)
    CUDA_TRY(ctx, cudaMemcpyAsync(full5_device + static_cast<size_t>(region_begin) * 5,
                                  piece5_device, mine * sizeof(float), cudaMemcpyDeviceToDevice,
                                  ctx->stream));
  return AMRVR_OK;
}

}  // namespace

extern "C" int amrvr_gather(amrvr_ctx* ctx, const float* piece5_device, int region_begin,
                            int region_end, int num_pixels, int root, float* full5_device) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  const int P = ctx->nranks, me = ctx->rank;
  REQUIRE(ctx, root >= 0 && root < P, "root out of range");
  REQUIRE(ctx, region_begin >= 0 && region_end >= region_begin && region_end <= num_pixels,
          "bad region");
  REQUIRE(ctx, me != root || full5_device != nullptr, "root needs a destination");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  std::vector<int> spans(static_cast<size_t>(2 * P));
  spans[2 * me] = region_begin;
  spans[2 * me + 1] = region_end;
  if (P > 1) {
    // every rank learns every region (ImageColorOnly.hpp:220-270 gathers sizes first)
    NcclApi& nc = nccl_api();
    DeviceBuffer tmp;
    CUDA_TRY(ctx, tmp.reserve(static_cast<size_t>(2 * (P + 1)) * sizeof(int)));
    int* all_dev = tmp.as<int>();
    const int span[2] = {region_begin, region_end};
    CUDA_TRY(ctx, cudaMemcpyAsync(all_dev + 2 * P, span, sizeof(span), cudaMemcpyHostToDevice,
                                  ctx->stream));
    NCCL_TRY(ctx, nc.AllGather(all_dev + 2 * P, all_dev, sizeof(span), ncclChar, ctx->comm,
                               ctx->stream));
    CUDA_TRY(ctx, cudaMemcpyAsync(spans.data(), all_dev, sizeof(int) * 2 * P, cudaMemcpyDeviceToHost,
                                  ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    tmp.release();
  }
  return gather_impl(ctx, piece5_device, spans, root, full5_device);
}

extern "C" int amrvr_downsample(amrvr_ctx* ctx, const float* full5_device, int width, int height,
                                int sqrt_aa, float* out5_device) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, full5_device && out5_device && width > 0 && height > 0, "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  LAUNCH_TRY(ctx, launch_downsample(full5_device, width, height, sqrt_aa, out5_device, ctx->stream,
                                    &ctx->launches));
  return AMRVR_OK;
}

extern "C" int amrvr_quantize_rgb8(amrvr_ctx* ctx, const float* full5_device, int width, int height,
                                   uint8_t* rgb8_device) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, full5_device && rgb8_device && width > 0 && height > 0, "bad argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  LAUNCH_TRY(ctx, launch_quantize_rgb8(full5_device, width, height, rgb8_device, ctx->stream,
                                       &ctx->launches));
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// memory helpers
// ---------------------------------------------------------------------------

extern "C" int amrvr_device_alloc(amrvr_ctx* ctx, size_t bytes, void** out) {
  if (ctx == nullptr || out == nullptr) return fail(ctx, AMRVR_ERR_INVALID, "NULL argument");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMalloc(out, std::max<size_t>(bytes, 1)));
  return AMRVR_OK;
}
extern "C" int amrvr_device_free(amrvr_ctx* ctx, void* ptr) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  CUDA_TRY(ctx, cudaFree(ptr));
  return AMRVR_OK;
}
extern "C" int amrvr_copy_to_host(amrvr_ctx* ctx, void* dst_host, const void* src_device,
                                  size_t bytes) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMemcpyAsync(dst_host, src_device, bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return AMRVR_OK;
}
extern "C" int amrvr_copy_to_device(amrvr_ctx* ctx, void* dst_device, const void* src_host,
                                    size_t bytes) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaMemcpyAsync(dst_device, src_host, bytes, cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return AMRVR_OK;
}
extern "C" int amrvr_host_alloc_pinned(size_t bytes, void** out) {
  if (out == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "out is NULL");
  CUDA_TRY(nullptr, cudaMallocHost(out, std::max<size_t>(bytes, 1)));
  return AMRVR_OK;
}
extern "C" int amrvr_host_free_pinned(void* ptr) {
  CUDA_TRY(nullptr, cudaFreeHost(ptr));
  return AMRVR_OK;
}
extern "C" int amrvr_synchronize(amrvr_ctx* ctx) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  return AMRVR_OK;
}

// ---------------------------------------------------------------------------
// whole frame
// ---------------------------------------------------------------------------
namespace {

struct FrameGeometry {
  int W, H, RW, RH, sqrt_aa, n;
  double tmin[3], tmax[3];  // tight bounds of all boxes (VolumeRenderer.cpp:791-849)
  float ref_step, aspect;
};

// Overlay edges in device memory (ctx->copy_jobs), or nullptr when the frame has none.
int upload_overlay_edges(amrvr_ctx* ctx, const amrvr_frame_params* params, const amrvr_camera* camera,
                         const FrameGeometry& G, const OverlayEdges** out) {
  *out = nullptr;
  if (!params->draw_bounding_box || G.n <= 0) return AMRVR_OK;
  OverlayEdges edges;
  make_overlay_edges(G.tmin, G.tmax, *camera, 1, G.W, G.H, edges);  // sqrtAA = 1 at :1313
  CUDA_TRY(ctx, ctx->copy_jobs.reserve(sizeof(OverlayEdges)));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->copy_jobs.ptr, &edges, sizeof(edges), cudaMemcpyHostToDevice,
                                ctx->stream));
  *out = ctx->copy_jobs.as<OverlayEdges>();
  return AMRVR_OK;
}

// The root's steps after the gather (VolumeRenderer.cpp:1302-1327): downsample,
// wireframe, float frame and bytes to the host.  ctx->full holds the RW x RH frame.
int post_root(amrvr_ctx* ctx, const amrvr_frame_params* params, const amrvr_camera* camera,
              const FrameGeometry& G, float* rgba5_host, uint8_t* rgb8_host) {
  const int W = G.W, H = G.H;
  const size_t out_floats = static_cast<size_t>(W) * H * 5;
  float* image = ctx->full.as<float>();
  if (G.sqrt_aa > 1) {
    CUDA_TRY(ctx, ctx->small.reserve(out_floats * sizeof(float)));
    LAUNCH_TRY(ctx, launch_downsample(ctx->full.as<float>(), W, H, G.sqrt_aa, ctx->small.as<float>(),
                                      ctx->stream, &ctx->launches));
    image = ctx->small.as<float>();
  }
  const bool want_bytes = rgb8_host != nullptr || params->rgb8_on_device;
  const bool fused_bytes = want_bytes && rgba5_host == nullptr;  // overlay + bytes in one pass
  if (params->draw_bounding_box && G.n > 0 && !fused_bytes) {
    OverlayEdges edges;
    make_overlay_edges(G.tmin, G.tmax, *camera, 1, W, H, edges);  // sqrtAA = 1 at :1313
    LAUNCH_TRY(ctx, launch_overlay(edges, 1, W, H, image, ctx->stream, &ctx->launches));
  }
  if (rgba5_host != nullptr)
    CUDA_TRY(ctx, cudaMemcpyAsync(rgba5_host, image, out_floats * sizeof(float), cudaMemcpyDeviceToHost,
                                  ctx->stream));
  if (want_bytes) {
    CUDA_TRY(ctx, ctx->rgb8.reserve(static_cast<size_t>(W) * H * 3));
    if (fused_bytes) {
      // nobody reads the float frame: the overlay is applied on the way to the bytes
      const OverlayEdges* edges_dev = nullptr;
      const int st = upload_overlay_edges(ctx, params, camera, G, &edges_dev);
      if (st != AMRVR_OK) return st;
      LAUNCH_TRY(ctx, launch_post_piece(image, 0, W * H, W, H, edges_dev, ctx->rgb8.as<uint8_t>(),
                                        ctx->stream, &ctx->launches));
    } else {
      LAUNCH_TRY(ctx, launch_quantize_rgb8(image, W, H, ctx->rgb8.as<uint8_t>(), ctx->stream,
                                           &ctx->launches));
    }
    ctx->last_rgb8 = ctx->rgb8.as<uint8_t>();
    if (rgb8_host != nullptr)
      CUDA_TRY(ctx, cudaMemcpyAsync(rgb8_host, ctx->rgb8.ptr, static_cast<size_t>(W) * H * 3,
                                    cudaMemcpyDeviceToHost, ctx->stream));
  }
  return AMRVR_OK;
}

// The visibility ordering of the frame (BuildVisibilityOrderedGroup).  A few
// hundred boxes are pure host work and run on the helper thread beside the
// march enqueue; larger tables use the device kernels on the aux stream, also
// from the helper thread.  join_order() must be called before the result is read.
struct OrderJob {
  std::vector<int> rank_order;
  int graph_failed = 0;
  int status = AMRVR_OK;
  float host_ms = 0.0f;
  bool on_thread = false;
};

void start_order(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_camera* camera, float aspect,
                 bool use_graph, OrderJob* job) {
  const int P = ctx->nranks;
  const int n = static_cast<int>(scene->global.size());
  job->rank_order.assign(static_cast<size_t>(P), 0);
  if (P == 1) return;  // one piece, one owner
  job->on_thread = true;
  const auto t0 = std::chrono::steady_clock::now();
  const amrvr_box_record* recs = scene->global.data();
  const amrvr_camera cam_copy = *camera;
  const bool device = n >= ctx->order_device_min;
  if (device) {  // grow the device buffers here: cudaMalloc on the helper thread would stall the frame
    const size_t cap = static_cast<size_t>(2 * n + 64);
    if (ctx->order_boxes.reserve(cap * sizeof(OrderBox)) != cudaSuccess ||
        ctx->order_scratch_a.reserve(cap * 16) != cudaSuccess ||
        ctx->order_scratch_b.reserve(cap * 16) != cudaSuccess ||
        ctx->order_priority.reserve(cap * 4) != cudaSuccess ||
        ctx->order_edges.reserve((16 * cap + 64) * sizeof(OrderEdge) + 16) != cudaSuccess) {
      job->status = AMRVR_ERR_CUDA;
      job->on_thread = false;
      return;
    }
  }
  ctx->worker.submit([=]() {
    bool ok;
    if (device) {
      cudaSetDevice(ctx->device);
      DeviceOrderBackend backend(ctx);
      ok = order_ranks(P, n, recs, cam_copy, aspect, use_graph, backend, job->rank_order.data());
      if (backend.status != 0) job->status = backend.status;
    } else {
      ok = order_ranks(P, n, recs, cam_copy, aspect, use_graph, *host_order_backend(),
                       job->rank_order.data());
    }
    job->graph_failed = ok ? 0 : 1;
    job->host_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
  });
}

int join_order(amrvr_ctx* ctx, OrderJob* job) {
  if (job->on_thread) {
    ctx->worker.wait();
    job->on_thread = false;
  }
  if (job->status != AMRVR_OK) return fail(ctx, job->status, "visibility ordering failed");
  ctx->last_rank_order = job->rank_order;
  ctx->last_graph_failed = job->graph_failed;
  return AMRVR_OK;
}

// A rank that gives up in the middle of a sharded frame raises its flags on every
// peer (so that nobody waits for it until the time-out) and their error words.
void abort_sharded_frame(amrvr_ctx* ctx) {
  if (!(ctx->peer_ok && ctx->nranks > 1)) return;
  const PeerPtrs ctl = peer_ptrs(ctx->sym_flags);
  std::vector<uint32_t> row(static_cast<size_t>(kMaxPeers), 0);
  const uint32_t epoch = ctx->barrier_epoch, bad = 2u;
  for (int r = 0; r < ctx->nranks; ++r) {
    uint32_t* base = static_cast<uint32_t*>(ctl.p[r]);
    cudaMemcpyAsync(base + kCtlError, &bad, 4, cudaMemcpyHostToDevice, ctx->stream);
    for (int which = 0; which < 2; ++which)
      cudaMemcpyAsync(base + which * kMaxPeers + ctx->rank, &epoch, 4, cudaMemcpyHostToDevice, ctx->stream);
  }
  cudaStreamSynchronize(ctx->stream);
  (void)cudaGetLastError();
  ctx->poisoned = true;
}

// One rank, or every GPU of an NVLink box with the bricks replicated: the whole
// frame is a handful of launches per GPU and no collective call.
//   copy    parameter block, first part (tables, one MarchBox per box, queue items)
//   march   one queue per GPU: the tiles of the 16-row frame strips that GPU folds
//           (stripe_plan); own queue first, then tiles of the others with the
//           fragments stored into their GPU's pool over NVLink; last CTA raises
//           "march finished" on every GPU
//   (beside the march, second stream: the host builds the fold table, it is copied
//    and binned into per-band layer lists)
//   fold    waits for every GPU's flag, folds this GPU's strips from LOCAL memory in
//           the reference's order and run structure, writes overlay + bytes straight
//           into the root's frame (and the float rows beside them); last CTA raises
//           "rows finished"
//   wait    every GPU waits for all of them (the root's frame is complete; pools
//           and queue counters are free for the next frame)
// Which pixels a rank folds does not change a pixel of the gathered frame, and
// neither does which rank the REFERENCE would let own piece p -- rank_order[p]
// from the visibility graph -- so the ordering runs beside the
// march on the helper thread and is reported (amrvr_ctx_last_rank_order);
// amrvr_composite, the Compositor::compose seam, deals the reference's contiguous
// pieces (getPieceRange, DirectSendBase.cpp:59-74) by it.
int frame_fused(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_camera* camera,
                const amrvr_frame_params* params, const amrvr_colormap_pt* pts, int npts, int root,
                float* rgba5_host, uint8_t* rgb8_host, amrvr_frame_stats* stats,
                const FrameGeometry& G, Trace& trace) {
  const bool sharded = ctx->nranks > 1;
  const int P = ctx->nranks, me = ctx->rank;
  const int W = G.W, H = G.H, RW = G.RW, RH = G.RH, n = G.n;
  const int npix = RW * RH;
  cudaEvent_t* ev = ctx->ev;
  if (ctx->poisoned)
    return fail(ctx, AMRVR_ERR_RUNTIME,
                "context unusable: an earlier frame lost a peer at a barrier; destroy and recreate it");
  if (ctx->frame_stats_host == nullptr)
    CUDA_TRY(ctx, cudaMallocHost(reinterpret_cast<void**>(&ctx->frame_stats_host), 128));

  // The visibility ordering runs beside the march on the helper thread.  Small tables: it is
  // started right after the march is enqueued (waking the thread costs microseconds of the
  // frame's head).  Tables that go to the device back end take about as long as the march
  // itself (C5: 4096 boxes), so they start first.
  OrderJob order;
  const bool order_first = n >= ctx->order_device_min;
  if (order_first) start_order(ctx, scene, camera, G.aspect, params->use_visibility_graph != 0, &order);
  // every return path joins the helper thread
  auto bail = [&](int status) {
    if (order.on_thread) ctx->worker.wait();
    if (sharded) abort_sharded_frame(ctx);
    return status;
  };
  if (sharded) ++ctx->barrier_epoch;
  const uint32_t epoch = ctx->barrier_epoch;

  // ---- geometry of the frame: footprints, tables, per-box march constants
  CameraBasis basis;
  make_camera_basis(*camera, basis);
  MarchFrame frame{};
  for (int d = 0; d < 3; ++d) {
    frame.eye[d] = basis.eye[d];
    frame.fwd[d] = basis.fwd[d];
    frame.right[d] = basis.right[d];
    frame.up[d] = basis.up[d];
  }
  frame.aspect = static_cast<float>(RW) / static_cast<float>(std::max(RH, 1));
  frame.tan_half = std::tan(camera->fov_y_degrees * 0.5f * 3.14159265358979323846f / 180.0f);
  frame.inv_w = 1.0f / static_cast<float>(RW);
  frame.inv_h = 1.0f / static_cast<float>(RH);
  frame.width = RW;
  frame.height = RH;
  std::vector<Rect> rects(static_cast<size_t>(n));
  box_footprints(n, scene->global.data(), basis, *camera, RW, RH, rects.data());
  if (scene->march_const_global.size() != static_cast<size_t>(n)) {
    std::vector<MarchBox>& mc = const_cast<amrvr_scene*>(scene)->march_const_global;
    mc.assign(static_cast<size_t>(n), MarchBox{});
    for (int g = 0; g < n; ++g)
      fill_march_const(scene->global[g].min_corner, scene->global[g].max_corner, scene->global[g].dims,
                       scene->global_brick_off[g], mc[g]);
  }
  trace.mark("footprints");
  const float alpha_scale = std::min(std::max(1.0f - params->box_transparency, 0.0f), 1.0f);
  std::vector<float> table_norm, tables;  // one table per distinct normalisation factor
  std::vector<MarchBox> per_box(scene->march_const_global);
  for (int g = 0; g < n; ++g) {
    const amrvr_box_record& B = scene->global[g];
    float step, norm;
    box_sampling(B.min_corner, B.max_corner, B.dims, G.ref_step, &step, &norm);
    int which = -1;
    for (size_t t = 0; t < table_norm.size(); ++t)
      if (std::memcmp(&table_norm[t], &norm, sizeof(float)) == 0) which = static_cast<int>(t);
    if (which < 0) {
      which = static_cast<int>(table_norm.size());
      table_norm.push_back(norm);
      tables.resize(tables.size() + 4 * AMRVR_COLOR_TABLE_SIZE);
      std::memcpy(tables.data() + static_cast<size_t>(which) * 4 * AMRVR_COLOR_TABLE_SIZE,
                  cached_color_table(ctx, alpha_scale, norm, scene->range_min, scene->range_max, pts, npts),
                  4 * AMRVR_COLOR_TABLE_SIZE * sizeof(float));
    }
    per_box[g].table = which;
    per_box[g].step = step;
    per_box[g].margin = inside_margin(per_box[g], basis.eye);
  }
  // boxes by table, so that a launch moves through the AMR levels once
  std::vector<int> by_table(static_cast<size_t>(n));
  for (int g = 0; g < n; ++g) by_table[g] = g;
  std::stable_sort(by_table.begin(), by_table.end(),
                   [&](int a, int b) { return per_box[a].table < per_box[b].table; });
  // a scene whose upload is still under way (one rank): boxes by copy group first, one
  // march launch per group, each waiting only for its own group's bricks
  const bool grouped = !sharded && scene->pending && !scene->group_ready.empty() &&
                       scene->box_group.size() == static_cast<size_t>(n);
  const int ngroups = grouped ? static_cast<int>(scene->group_ready.size()) : 0;
  if (grouped)
    std::stable_sort(by_table.begin(), by_table.end(),
                     [&](int a, int b) { return scene->box_group[a] < scene->box_group[b]; });
  else if (wait_upload(ctx, scene) != AMRVR_OK)
    return bail(AMRVR_ERR_CUDA);
  std::vector<int> group_tile_begin(static_cast<size_t>(ngroups) + 1, 0);

  trace.mark("boxconst");
  // ---- march items.  One rank: one queue, a box's tiles cover its footprint from its top
  // row.  Several ranks: the frame is STRIPED -- GPU q folds the 16-row strips k = q, q + P,
  // ... of the frame and queue q holds, of every box, the tiles of those strips (QueueItem),
  // so every GPU gets an even share of every footprint: the march and the fold balance
  // without any tile changing hands, and stealing only evens out what is left.  Which
  // pixels a rank folds is bookkeeping: the gathered frame does not depend on it (each
  // pixel is folded once, in the reference's global layer order with its run structure;
  // pixel ownership in the reference's sense is what amrvr_composite reports).
  std::vector<MarchBox> items;
  items.reserve(static_cast<size_t>(n));
  std::vector<int> my_item(static_cast<size_t>(n), -1);  // global box -> its item
  std::vector<int> item_box;                              // item -> global box
  for (int k = 0; k < n; ++k) {
    const int g = by_table[k];
    const Rect& r = rects[g];
    if (r.w <= 0 || r.h <= 0) continue;
    MarchBox M = per_box[g];
    M.rx0 = r.x0;
    M.ry0 = r.y0;
    M.rw = r.w;
    M.rh = r.h;
    M.tiles_x = (r.w + 15) / 16;
    M.ry_first = r.y0;
    M.dest = 0;
    my_item[g] = static_cast<int>(items.size());
    item_box.push_back(g);
    items.push_back(M);
  }
  const int nb = static_cast<int>(items.size());
  auto by_item_box = [&](int i) { return item_box[static_cast<size_t>(i)]; };
  std::vector<QueueItem> qitems(sharded ? static_cast<size_t>(P) * nb : 0);
  std::vector<uint64_t> my_layer_off(static_cast<size_t>(nb), 0);
  std::vector<char> mine_has_rows(static_cast<size_t>(nb), 0);
  MarchQueues queues{};
  queues.nqueues = P;
  queues.mine = sharded ? me : 0;
  std::vector<uint64_t> pool_floats(static_cast<size_t>(P), 0);
  int tiles = 0;
  int64_t my_rays = 0;
  if (!sharded) {
    queues.tile_begin[0] = 0;
    for (int i = 0; i < nb; ++i) {
      MarchBox& M = items[static_cast<size_t>(i)];
      M.tile_base = tiles;
      tiles += M.tiles_x * ((M.rh + 15) / 16);
      if (grouped) group_tile_begin[static_cast<size_t>(scene->box_group[by_item_box(i)]) + 1] = tiles;
      M.layer_off = pool_floats[0];
      pool_floats[0] += static_cast<uint64_t>(M.rw) * M.rh * 5;
      my_layer_off[static_cast<size_t>(i)] = M.layer_off;
      mine_has_rows[static_cast<size_t>(i)] = 1;
      my_rays += static_cast<int64_t>(M.rw) * M.rh;
    }
  } else {
    std::vector<int32_t> ry0(static_cast<size_t>(nb)), rh(static_cast<size_t>(nb)), rw(static_cast<size_t>(nb));
    for (int i = 0; i < nb; ++i) {
      ry0[static_cast<size_t>(i)] = items[static_cast<size_t>(i)].ry0;
      rh[static_cast<size_t>(i)] = items[static_cast<size_t>(i)].rh;
      rw[static_cast<size_t>(i)] = items[static_cast<size_t>(i)].rw;
    }
    stripe_plan(P, nb, ry0.data(), rh.data(), rw.data(), qitems.data(), queues.tile_begin, pool_floats.data());
    tiles = queues.tile_begin[P];
    for (int i = 0; i < nb; ++i) {  // what this rank folds of every box
      const MarchBox& M = items[static_cast<size_t>(i)];
      const int y0 = M.ry0, y1 = M.ry0 + M.rh;
      const int ns = stripe_count(P, me, y0, y1);
      my_layer_off[static_cast<size_t>(i)] = qitems[static_cast<size_t>(me) * nb + i].layer_off;
      mine_has_rows[static_cast<size_t>(i)] = ns > 0;
      for (int k = stripe_first_strip(P, me, y0), j = 0; j < ns; ++j, k += P)
        my_rays += static_cast<int64_t>(M.rw) * (std::min(16 * k + 16, y1) - std::max(16 * k, y0));
    }
  }
  queues.tile_begin[P] = tiles;
  for (int g = 1; g <= ngroups; ++g)  // groups without a visible box: empty ranges
    group_tile_begin[g] = std::max(group_tile_begin[g], group_tile_begin[g - 1]);
  frame.nboxes = nb;
  frame.ntiles = tiles;
  frame.strip_stride = 16 * P;
  // the rows this rank folds: strips me, me + P, ... (all rows on one rank)
  const int strips_total = (RH + 15) / 16;
  const int my_strips = sharded ? ((strips_total > me) ? (strips_total - 1 - me) / P + 1 : 0) : strips_total;
  {  // the other queues, heaviest first
    std::vector<int> others;
    for (int p = 0; p < P; ++p)
      if (p != queues.mine) others.push_back(p);
    std::stable_sort(others.begin(), others.end(), [&](int a, int b) {
      return queues.tile_begin[a + 1] - queues.tile_begin[a] > queues.tile_begin[b + 1] - queues.tile_begin[b];
    });
    for (size_t i = 0; i < others.size(); ++i) queues.steal_order[i] = others[i];
  }
  trace.mark("plan");

  // ---- buffers (the peer-mapped ones grow collectively: every rank computes the same sizes)
  const bool bytes_mode = G.sqrt_aa == 1;  // the fold writes overlay + RGB8 itself
  const bool single_bytes = !sharded && bytes_mode && rgba5_host == nullptr &&
                            (rgb8_host != nullptr || params->rgb8_on_device);
  const size_t most_floats = *std::max_element(pool_floats.begin(), pool_floats.end());
  float* pool_local = nullptr;
  if (sharded) {
    bool ok_a = true, ok_b = true, ok_c = true;
    // the float piece of a rank: its strips one after the other, 16 rows each
    const size_t max_piece = static_cast<size_t>((strips_total + P - 1) / P) * 16 * RW;
    int st = sym_reserve(ctx, ctx->sym_pool, std::max<size_t>(most_floats * sizeof(float), 256), &ok_a);
    if (st == AMRVR_OK) st = sym_reserve(ctx, ctx->sym_piece, std::max<size_t>(max_piece * 20, 256), &ok_b);
    if (st == AMRVR_OK && bytes_mode)
      st = sym_reserve(ctx, ctx->sym_rgb8, static_cast<size_t>(W) * H * 3, &ok_c);
    if (st != AMRVR_OK) return bail(st);
    if (!ok_a || !ok_b || !ok_c)
      return bail(fail(ctx, AMRVR_ERR_CUDA, "peer-mapped frame buffers could not be grown on every rank"));
    pool_local = static_cast<float*>(ctx->sym_pool.local);
    for (int p = 0; p < P; ++p) {
      queues.pool[p] = static_cast<float*>(ctx->sym_pool.peer[p]);
      queues.ctl[p] = static_cast<unsigned*>(ctx->sym_flags.peer[p]);
      queues.counter[p] = queues.ctl[p] + kCtlTileCounter;
    }
    queues.me = me;
    queues.nranks = P;
    queues.epoch = epoch;
    queues.signal_row = 0;
  } else {
    if (cudaSuccess != ctx->frame_pool.reserve(std::max<size_t>(most_floats * sizeof(float), 256)))
      return bail(fail(ctx, AMRVR_ERR_CUDA, "fragment pool allocation failed"));
    pool_local = ctx->frame_pool.as<float>();
    queues.pool[0] = pool_local;
  }
  if (me == root || !sharded) {
    if (ctx->full.reserve(static_cast<size_t>(npix) * 20) != cudaSuccess)
      return bail(fail(ctx, AMRVR_ERR_CUDA, "frame buffer allocation failed"));
    if (single_bytes && ctx->rgb8.reserve(static_cast<size_t>(W) * H * 3) != cudaSuccess)
      return bail(fail(ctx, AMRVR_ERR_CUDA, "frame buffer allocation failed"));
  }

  // ---- one parameter block: [header | tables | items | fold table | overlay edges].  The
  // march needs the first three only: they are copied and the march is launched BEFORE
  // the host builds the fold table, which then travels while the march runs.
  const int my_begin = 0, my_end = npix;  // one rank: every pixel; striped: FoldStrips says which rows
  const FoldStrips strips{P, me, RH};
  const FoldStrips* my_strips_arg = sharded ? &strips : nullptr;
  const size_t tables_bytes = tables.size() * sizeof(float);
  const size_t items_bytes = items.size() * sizeof(MarchBox);
  const size_t qitems_bytes = qitems.size() * sizeof(QueueItem);
  const size_t fold_bytes_max = static_cast<size_t>(n) * sizeof(FoldLayer);
  const size_t counters_bytes = (static_cast<size_t>(ngroups) * sizeof(uint32_t) + 15) & ~size_t(15);
  const size_t off_items = kParamHeader + tables_bytes;
  const size_t off_qitems = off_items + items_bytes;      // striped frame: [queue][box] table
  const size_t off_counters = off_qitems + qitems_bytes;  // tile counters of the per-group launches
  const size_t off_fold = off_counters + counters_bytes;
  const size_t off_edges = off_fold + fold_bytes_max;
  const size_t block_bytes = off_edges + sizeof(OverlayEdges);
  unsigned char* block = nullptr;
  {
    const int st = param_staging(ctx, block_bytes, &block);
    if (st != AMRVR_OK) return bail(st);
  }
  std::memset(block, 0, kParamHeader);
  if (tables_bytes) std::memcpy(block + kParamHeader, tables.data(), tables_bytes);
  if (items_bytes) std::memcpy(block + off_items, items.data(), items_bytes);
  if (qitems_bytes) std::memcpy(block + off_qitems, qitems.data(), qitems_bytes);
  if (counters_bytes) std::memset(block + off_counters, 0, counters_bytes);
  if (ctx->march_boxes.reserve(block_bytes) != cudaSuccess ||
      ctx->fold_bins.reserve(std::max<size_t>(fold_scratch_bytes(n, RW, my_begin, my_end, my_strips_arg), 256)) !=
          cudaSuccess)
    return bail(fail(ctx, AMRVR_ERR_CUDA, "parameter block allocation failed"));
  unsigned char* dev_block = ctx->march_boxes.as<unsigned char>();
  queues.stats = reinterpret_cast<unsigned long long*>(dev_block);
  queues.queue_items = sharded ? reinterpret_cast<const QueueItem*>(dev_block + off_qitems) : nullptr;
  if (!sharded) queues.counter[0] = reinterpret_cast<unsigned*>(dev_block + 24);
  const FoldLayer* fold_dev = reinterpret_cast<const FoldLayer*>(dev_block + off_fold);
  trace.mark("tables");

  auto try_cuda = [&](cudaError_t e, const char* what) {
    return (e == cudaSuccess) ? AMRVR_OK
                              : fail(ctx, AMRVR_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
  };
  auto try_launch = [&](int e, const char* what) {
    return (e == 0) ? AMRVR_OK
                    : fail(ctx, AMRVR_ERR_CUDA,
                           std::string(what) + ": " + cudaGetErrorString(static_cast<cudaError_t>(e)));
  };
  // While FAB copies fill the host-to-device copy engine a cudaMemcpyAsync would queue
  // behind all of them: the march's parameters are then fetched by a small kernel that
  // reads the pinned staging block across PCIe itself.
  int st = grouped ? try_launch(launch_peer_copy(block, dev_block, off_fold,
                                                 static_cast<int>((off_fold / 16 + 255) / 256), ctx->stream,
                                                 &ctx->launches),
                                "parameter fetch")
                   : try_cuda(cudaMemcpyAsync(dev_block, block, off_fold, cudaMemcpyHostToDevice, ctx->stream),
                              "parameter copy");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ev[1], ctx->stream), "event");
  if (!grouped) {
    if (st == AMRVR_OK)
      st = try_launch(launch_raymarch(frame, queues, reinterpret_cast<const MarchBox*>(dev_block + off_items),
                                      scene->bricks.as<uint8_t>(),
                                      reinterpret_cast<const float*>(dev_block + kParamHeader),
                                      params->count_samples != 0, ctx->stream, &ctx->launches),
                      "march");
  } else {
    for (int g = 0; g < ngroups && st == AMRVR_OK; ++g) {
      // the bricks of group g are quantised; later groups are still on PCIe
      st = try_cuda(cudaStreamWaitEvent(ctx->stream, scene->group_ready[static_cast<size_t>(g)], 0), "wait");
      MarchQueues q = queues;
      q.tile_begin[0] = group_tile_begin[static_cast<size_t>(g)];
      q.tile_begin[1] = group_tile_begin[static_cast<size_t>(g) + 1];
      q.counter[0] = reinterpret_cast<unsigned*>(dev_block + off_counters) + g;
      MarchFrame fr = frame;
      fr.ntiles = q.tile_begin[1] - q.tile_begin[0];
      if (st == AMRVR_OK && fr.ntiles > 0)
        st = try_launch(launch_raymarch(fr, q, reinterpret_cast<const MarchBox*>(dev_block + off_items),
                                        scene->bricks.as<uint8_t>(),
                                        reinterpret_cast<const float*>(dev_block + kParamHeader),
                                        params->count_samples != 0, ctx->stream, &ctx->launches),
                        "march");
    }
  }
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ev[2], ctx->stream), "event");
  trace.mark("march_enqueue");
  if (!order_first) start_order(ctx, scene, camera, G.aspect, params->use_visibility_graph != 0, &order);

  // ---- fold table of my piece, in the reference's blend order with its run ids
  // (DirectSendBase.cpp:363-407); built while the march runs
  std::vector<int> blend_order(static_cast<size_t>(n));
  layer_order(n, scene->global.data(), *camera, nullptr, blend_order.data());
  std::vector<FoldLayer> fold;
  fold.reserve(static_cast<size_t>(n));
  {
    int run = -1, prev_owner = -1;
    for (int pos = 0; pos < n; ++pos) {
      const int g = blend_order[pos];
      if (pos == 0 || scene->global[g].owner != prev_owner) ++run;
      prev_owner = scene->global[g].owner;
      if (my_item[g] < 0 || !mine_has_rows[static_cast<size_t>(my_item[g])]) continue;
      const MarchBox& M = items[static_cast<size_t>(my_item[g])];
      FoldLayer F{};
      F.base = pool_local + my_layer_off[static_cast<size_t>(my_item[g])];
      F.x0 = M.rx0;
      F.w = M.rw;
      F.row0 = M.ry0;  // the footprint's rows; a striped layer stores this rank's strips of them
      F.row1 = M.ry0 + M.rh;
      F.run = run;
      fold.push_back(F);
    }
  }
  OverlayEdges edges{};
  const bool overlay = bytes_mode && params->draw_bounding_box && n > 0;
  if (overlay) make_overlay_edges(G.tmin, G.tmax, *camera, 1, W, H, edges);
  const size_t fold_bytes = fold.size() * sizeof(FoldLayer);
  if (fold_bytes) std::memcpy(block + off_fold, fold.data(), fold_bytes);
  std::memcpy(block + off_edges, &edges, sizeof(OverlayEdges));
  // on the plan stream, beside the march; the fold waits for it
  if (st == AMRVR_OK && ctx->bin_done == nullptr)
    st = try_cuda(cudaEventCreateWithFlags(&ctx->bin_done, cudaEventDisableTiming), "event");
  if (st == AMRVR_OK)
    st = try_cuda(cudaMemcpyAsync(dev_block + off_fold, block + off_fold, block_bytes - off_fold,
                                  cudaMemcpyHostToDevice, ctx->plan),
                  "fold table copy");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ctx->param_copied, ctx->plan), "event");
  if (st == AMRVR_OK)
    st = try_launch(launch_fold_bin(fold_dev, static_cast<int>(fold.size()), RW, my_begin, my_end,
                                    ctx->fold_bins.ptr, ctx->plan, &ctx->launches, my_strips_arg),
                    "fold_bin");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ctx->bin_done, ctx->plan), "event");
  if (st == AMRVR_OK) st = try_cuda(cudaStreamWaitEvent(ctx->stream, ctx->bin_done, 0), "wait");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ev[6], ctx->stream), "event");

  // ---- fold (+ bytes) of my piece
  FoldSync sync{};
  if (sharded) {
    sync.ctl = peer_ptrs(ctx->sym_flags);
    sync.me = me;
    sync.nranks = P;
    sync.wait_row = 0;
    sync.signal_row = 1;
    sync.epoch = epoch;
    sync.timeout_ns = 10ull * 1000 * 1000 * 1000;
  }
  FoldBytes fb{nullptr, H, nullptr};
  float* piece_dst;
  if (sharded) {
    piece_dst = static_cast<float*>(ctx->sym_piece.local);
    if (bytes_mode) fb.rgb8 = static_cast<uint8_t*>(ctx->sym_rgb8.peer[root]);
  } else {
    piece_dst = single_bytes ? nullptr : ctx->full.as<float>();
    if (single_bytes) fb.rgb8 = ctx->rgb8.as<uint8_t>();
  }
  if (fb.rgb8 != nullptr && overlay) fb.edges = reinterpret_cast<const OverlayEdges*>(dev_block + off_edges);
  if (st == AMRVR_OK)
    st = try_launch(launch_fold(fold_dev, static_cast<int>(fold.size()), RW, my_begin, my_end, piece_dst,
                                ctx->fold_bins.ptr, ctx->stream, &ctx->launches,
                                fb.rgb8 != nullptr ? &fb : nullptr, sharded ? &sync : nullptr, my_strips_arg),
                    "fold");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ev[3], ctx->stream), "event");
  if (st == AMRVR_OK && sharded)
    st = try_launch(launch_peer_wait(sync.ctl, me, P, 1, epoch, sync.timeout_ns, ctx->stream, &ctx->launches),
                    "wait");
  if (st == AMRVR_OK) st = try_cuda(cudaEventRecord(ev[4], ctx->stream), "event");
  if (st != AMRVR_OK) return bail(st);
  trace.mark("fold_enqueue");

  // ---- the root's frame
  ctx->last_rgb8 = nullptr;
  const bool is_root = !sharded || me == root;
  if (is_root) {
    const bool need_floats = rgba5_host != nullptr || !bytes_mode ||
                             (!sharded && !single_bytes);
    if (sharded && need_floats) {
      // the root pulls the float pieces over NVLink (ImageColorOnly::Gather)
      // rank p's piece holds its strips p, p + P, ... one after the other: one pitched copy per
      // rank puts them where they belong, then the last (partial) strip of the frame by itself
      const size_t strip_bytes = static_cast<size_t>(16) * RW * 20;
      for (int p = 0; p < P; ++p) {
        const int mine_p = (strips_total > p) ? (strips_total - 1 - p) / P + 1 : 0;
        if (mine_p <= 0) continue;
        const int last_k = p + (mine_p - 1) * P;  // the rank's last strip may be cut by the frame's end
        const int full_strips = (16 * last_k + 16 <= RH) ? mine_p : mine_p - 1;
        char* dst = reinterpret_cast<char*>(ctx->full.as<float>()) + static_cast<size_t>(p) * strip_bytes;
        const char* src = static_cast<const char*>(ctx->sym_piece.peer[p]);
        if (full_strips > 0 &&
            cudaMemcpy2DAsync(dst, strip_bytes * P, src, strip_bytes, strip_bytes, static_cast<size_t>(full_strips),
                              cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess)
          return bail(fail(ctx, AMRVR_ERR_CUDA, "piece gather failed"));
        if (full_strips < mine_p &&
            cudaMemcpyAsync(dst + static_cast<size_t>(full_strips) * strip_bytes * P,
                            src + static_cast<size_t>(full_strips) * strip_bytes,
                            static_cast<size_t>(RH - 16 * last_k) * RW * 20, cudaMemcpyDeviceToDevice,
                            ctx->stream) != cudaSuccess)
          return bail(fail(ctx, AMRVR_ERR_CUDA, "piece gather failed"));
      }
    }
    if (fb.rgb8 != nullptr && !(sharded && !bytes_mode)) {
      // the bytes are complete (the fold wrote overlay + RGB8)
      uint8_t* built = sharded ? static_cast<uint8_t*>(ctx->sym_rgb8.local) : ctx->rgb8.as<uint8_t>();
      ctx->last_rgb8 = built;
      if (rgba5_host != nullptr) {
        float* image = ctx->full.as<float>();
        if (params->draw_bounding_box && n > 0) {
          OverlayEdges e2;
          make_overlay_edges(G.tmin, G.tmax, *camera, 1, W, H, e2);
          if (launch_overlay(e2, 1, W, H, image, ctx->stream, &ctx->launches) != 0)
            return bail(fail(ctx, AMRVR_ERR_CUDA, "overlay failed"));
        }
        if (cudaMemcpyAsync(rgba5_host, image, static_cast<size_t>(W) * H * 20, cudaMemcpyDeviceToHost,
                            ctx->stream) != cudaSuccess)
          return bail(fail(ctx, AMRVR_ERR_CUDA, "frame readback failed"));
      }
      if (rgb8_host != nullptr &&
          cudaMemcpyAsync(rgb8_host, built, static_cast<size_t>(W) * H * 3, cudaMemcpyDeviceToHost,
                          ctx->stream) != cudaSuccess)
        return bail(fail(ctx, AMRVR_ERR_CUDA, "frame readback failed"));
    } else {
      st = post_root(ctx, params, camera, G, rgba5_host, rgb8_host);
      if (st != AMRVR_OK) return bail(st);
    }
  }
  if (cudaEventRecord(ev[5], ctx->stream) != cudaSuccess)
    return bail(fail(ctx, AMRVR_ERR_CUDA, "event record failed"));
  trace.mark("post_enqueue");

  // ---- statistics and the barrier time-out word come back with one small copy
  if (cudaMemcpyAsync(ctx->frame_stats_host, dev_block, kParamHeader, cudaMemcpyDeviceToHost, ctx->stream) !=
      cudaSuccess)
    return bail(fail(ctx, AMRVR_ERR_CUDA, "statistics readback failed"));
  ctx->frame_stats_host[8] = 0;
  if (sharded &&
      cudaMemcpyAsync(ctx->frame_stats_host + 8, static_cast<uint32_t*>(ctx->sym_flags.local) + kCtlError,
                      sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream) != cudaSuccess)
    return bail(fail(ctx, AMRVR_ERR_CUDA, "statistics readback failed"));
  if (cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
    if (order.on_thread) ctx->worker.wait();
    ctx->poisoned = sharded;
    return fail(ctx, AMRVR_ERR_CUDA, std::string("frame failed: ") + cudaGetErrorString(cudaGetLastError()));
  }
  trace.mark("final_sync");
  if (grouped) {  // every group was waited for on the stream
    const_cast<amrvr_scene*>(scene)->pending = false;
    async_upload_done(ctx, scene);
  }
  st = join_order(ctx, &order);
  if (st != AMRVR_OK) return st;
  if (sharded && static_cast<uint32_t>(ctx->frame_stats_host[8]) != 0) {
    ctx->poisoned = true;
    return fail(ctx, AMRVR_ERR_RUNTIME,
                static_cast<uint32_t>(ctx->frame_stats_host[8]) == 2u
                    ? "a peer rank gave up in the middle of the frame"
                    : "a peer GPU did not reach the frame barrier within 10 s");
  }
  if (stats != nullptr) {
    std::memset(stats, 0, sizeof(*stats));
    stats->samples = static_cast<int64_t>(ctx->frame_stats_host[0]);
    stats->exchanged_bytes = static_cast<int64_t>(ctx->frame_stats_host[1]);
    stats->stolen_tiles = static_cast<int64_t>(ctx->frame_stats_host[2]);
    stats->tiles = tiles;
    stats->fragment_pixels = my_rays;
    if (ctx->frame_stats_host[6] != 0) {  // the march ran: its time line on the device clock
      stats->march_span_ns = static_cast<int64_t>(ctx->frame_stats_host[6] - ~ctx->frame_stats_host[5]);
      stats->march_cta_busy_ns = static_cast<int64_t>(ctx->frame_stats_host[4]);
      stats->march_table_ns = static_cast<int64_t>(ctx->frame_stats_host[7]);
      stats->march_ctas = static_cast<int64_t>(ctx->frame_stats_host[3] >> 32);
    }
    stats->order_ms = order.host_ms;  // beside the march, on the helper thread: host wall time
    cudaEventElapsedTime(&stats->march_ms, ev[1], ev[2]);
    cudaEventElapsedTime(&stats->pre_ms, ev[0], ev[1]);
    cudaEventElapsedTime(&stats->fold_ms, ev[6], ev[3]);       // the fold kernel (it starts with the hand-shake)
    cudaEventElapsedTime(&stats->composite_ms, ev[2], ev[4]);  // fold table copy + bin + fold + final wait
    stats->gather_ms = 0.0f;
    cudaEventElapsedTime(&stats->post_ms, ev[4], ev[5]);
  }
  return AMRVR_OK;
}

// Ranks that cannot map each other's memory (or scenes too large to replicate):
// every rank marches its own boxes, the fragments travel as one grouped NCCL
// send/recv (alltoallv, DirectSendBase.cpp:76-177 without the empty frames), the
// pieces are folded where the visibility order puts them and gathered on the root.
int frame_exchange(amrvr_ctx* ctx, const amrvr_scene* scene, const amrvr_camera* camera,
                   const amrvr_frame_params* params, const amrvr_colormap_pt* pts, int npts, int root,
                   float* rgba5_host, uint8_t* rgb8_host, amrvr_frame_stats* stats,
                   const FrameGeometry& G, Trace& trace) {
  cudaEvent_t* ev = ctx->ev;
  OrderJob order;
  start_order(ctx, scene, camera, G.aspect, params->use_visibility_graph != 0, &order);
  amrvr_layers* layers = nullptr;
  int st = amrvr_raymarch(ctx, scene, camera, G.RW, G.RH, params->box_transparency, G.ref_step, pts, npts,
                          params->count_samples, &layers);
  {
    const int joined = join_order(ctx, &order);  // the helper thread writes into `order`
    if (st == AMRVR_OK) st = joined;
  }
  if (st != AMRVR_OK) {
    amrvr_layers_free(ctx, layers);
    return st;
  }
  CUDA_TRY(ctx, cudaEventRecord(ev[1], ctx->stream));
  trace.mark("march_enqueue");
  CUDA_TRY(ctx, cudaEventRecord(ev[2], ctx->stream));
  trace.mark("order");
  int begin = 0, end = 0;
  float* piece = nullptr;
  int64_t exchanged = 0;
  if (ctx->rank == root) CUDA_TRY(ctx, ctx->full.reserve(static_cast<size_t>(G.RW) * G.RH * 20));
  st = composite_impl(ctx, scene, layers, camera, order.rank_order.data(), &begin, &end, &piece,
                      &exchanged);
  if (st == AMRVR_OK) {
    CUDA_TRY(ctx, cudaEventRecord(ev[3], ctx->stream));
    trace.mark("composite_enqueue");
    std::vector<int> spans(static_cast<size_t>(2 * ctx->nranks));
    for (int pc = 0; pc < ctx->nranks; ++pc)
      piece_range(G.RW * G.RH, pc, ctx->nranks, &spans[2 * order.rank_order[pc]],
                  &spans[2 * order.rank_order[pc] + 1]);
    st = gather_impl(ctx, piece, spans, root, ctx->full.as<float>());
  }
  if (st != AMRVR_OK) {
    amrvr_layers_free(ctx, layers);
    return st;
  }
  CUDA_TRY(ctx, cudaEventRecord(ev[4], ctx->stream));
  ctx->last_rgb8 = nullptr;
  if (ctx->rank == root) {
    st = post_root(ctx, params, camera, G, rgba5_host, rgb8_host);
    if (st != AMRVR_OK) {
      amrvr_layers_free(ctx, layers);
      return st;
    }
  }
  CUDA_TRY(ctx, cudaEventRecord(ev[5], ctx->stream));
  trace.mark("post_enqueue");
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  trace.mark("final_sync");
  if (stats != nullptr) {
    std::memset(stats, 0, sizeof(*stats));
    amrvr_march_stats ms;
    if (amrvr_layers_stats(layers, &ms) == AMRVR_OK) {
      stats->samples = ms.samples;
      stats->march_ms = ms.kernel_ms;
      stats->fragment_pixels = ms.fragment_pixels;
    }
    stats->order_ms = order.host_ms;
    cudaEventElapsedTime(&stats->composite_ms, ev[2], ev[3]);
    cudaEventElapsedTime(&stats->gather_ms, ev[3], ev[4]);
    cudaEventElapsedTime(&stats->post_ms, ev[4], ev[5]);
    cudaEventElapsedTime(&stats->fold_ms, ev[6], ev[7]);
    stats->exchanged_bytes = exchanged;
  }
  amrvr_layers_free(ctx, layers);
  return AMRVR_OK;
}

}  // namespace

extern "C" int amrvr_render_frame(amrvr_ctx* ctx, const amrvr_scene* scene,
                                  const amrvr_camera* camera, const amrvr_frame_params* params,
                                  const amrvr_colormap_pt* pts, int npts, int root,
                                  float* rgba5_host, uint8_t* rgb8_host, amrvr_frame_stats* stats) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, scene && camera && params, "NULL argument");
  REQUIRE(ctx, params->width > 0 && params->height > 0, "width and height must be positive");
  REQUIRE(ctx, root >= 0 && root < ctx->nranks, "root out of range");
  REQUIRE(ctx, npts >= 0 && (npts == 0 || pts != nullptr), "bad colour map");
  FrameGeometry G{};
  G.sqrt_aa = static_cast<int>(std::lround(std::sqrt(static_cast<double>(params->antialiasing))));
  REQUIRE(ctx, params->antialiasing >= 1 && G.sqrt_aa * G.sqrt_aa == params->antialiasing,
          "antialiasing must be a positive perfect square");  // VolumeRenderer.cpp:425-437
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  G.W = params->width;
  G.H = params->height;
  G.RW = G.W * G.sqrt_aa;
  G.RH = G.H * G.sqrt_aa;
  REQUIRE(ctx, static_cast<int64_t>(G.RW) * G.RH < (1LL << 31), "frame too large");
  G.aspect = static_cast<float>(G.W) / static_cast<float>(G.H);
  G.n = static_cast<int>(scene->global.size());
  // tight bounds of all boxes (VolumeRenderer.cpp:791-849), also the fallback
  // bounds of the reference step
  for (int d = 0; d < 3 && G.n > 0; ++d) {
    double lo = scene->global[0].min_corner[d], hi = scene->global[0].max_corner[d];
    for (int g = 1; g < G.n; ++g) {
      lo = std::min(lo, scene->global[g].min_corner[d]);
      hi = std::max(hi, scene->global[g].max_corner[d]);
    }
    G.tmin[d] = static_cast<double>(static_cast<float>(lo));
    G.tmax[d] = static_cast<double>(static_cast<float>(hi));
  }
  G.ref_step = reference_sample_distance(G.n, scene->global.data(), G.tmin, G.tmax);

  Trace trace;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
  trace.mark("setup");
  const bool fused = ctx->nranks == 1 || (ctx->peer_ok && scene->replicated);
  return fused ? frame_fused(ctx, scene, camera, params, pts, npts, root, rgba5_host, rgb8_host, stats, G,
                             trace)
               : frame_exchange(ctx, scene, camera, params, pts, npts, root, rgba5_host, rgb8_host,
                                stats, G, trace);
}

// ---------------------------------------------------------------------------
// NVLink probe
// ---------------------------------------------------------------------------

// Collective.  Every rank copies `bytes` between its own buffer and the buffer
// of rank (rank + peer_offset) mod nranks with a plain load/store kernel --
// write != 0: local -> peer (posted stores, what the march does with stolen
// tiles); write == 0: peer -> local (loads) -- `iters` times, all ranks at once,
// and reports its own rate.  This is the measured per-GPU, per-direction link
// figure the composite numbers are quoted against.
extern "C" int amrvr_peer_bandwidth(amrvr_ctx* ctx, size_t bytes, int peer_offset, int write, int iters,
                                    float* gbytes_per_s) {
  if (ctx == nullptr) return fail(nullptr, AMRVR_ERR_INVALID, "ctx is NULL");
  REQUIRE(ctx, gbytes_per_s != nullptr && bytes >= 4096 && iters >= 1, "bad argument");
  REQUIRE(ctx, ctx->nranks > 1 && ctx->peer_ok, "needs several ranks with peer access");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  bool ok = true;
  const size_t want = 2 * bytes;  // [0, bytes): source of this rank; [bytes, 2*bytes): landing zone
  const int st = sym_reserve(ctx, ctx->sym_pool, want, &ok);
  if (st != AMRVR_OK) return st;
  REQUIRE(ctx, ok, "peer-mapped buffer could not be grown");
  const int P = ctx->nranks;
  const int peer = ((ctx->rank + peer_offset) % P + P) % P;
  char* mine = static_cast<char*>(ctx->sym_pool.local);
  char* theirs = static_cast<char*>(ctx->sym_pool.peer[peer]);
  const void* src = write ? mine : theirs;
  void* dst = write ? theirs + bytes : mine + bytes;
  int st2 = peer_barrier(ctx, 2);
  if (st2 != AMRVR_OK) return st2;
  cudaEvent_t e0 = ctx->ev[6], e1 = ctx->ev[7];
  LAUNCH_TRY(ctx, launch_peer_copy(src, dst, bytes, 148 * 8, ctx->stream, &ctx->launches));  // warm-up
  st2 = peer_barrier(ctx, 2);
  if (st2 != AMRVR_OK) return st2;
  CUDA_TRY(ctx, cudaEventRecord(e0, ctx->stream));
  for (int i = 0; i < iters; ++i)
    LAUNCH_TRY(ctx, launch_peer_copy(src, dst, bytes, 148 * 8, ctx->stream, &ctx->launches));
  CUDA_TRY(ctx, cudaEventRecord(e1, ctx->stream));
  st2 = peer_barrier(ctx, 2);
  if (st2 != AMRVR_OK) return st2;
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  float ms = 0.0f;
  CUDA_TRY(ctx, cudaEventElapsedTime(&ms, e0, e1));
  *gbytes_per_s = (ms > 0.0f) ? static_cast<float>(static_cast<double>(bytes) * iters / (ms * 1e-3) / 1e9) : 0.0f;
  return AMRVR_OK;
}
