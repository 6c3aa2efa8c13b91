// B200Runtime.cpp -- see B200Runtime.hpp.
#include "B200Runtime.hpp"

#include <cstdlib>
#include <map>
#include <stdexcept>
#include <string>
#include <vector>

namespace amrvr_b200 {

namespace {

int pickDevice(int rank) {
  if (const char* env = std::getenv("AMRVR_DEVICE")) return std::atoi(env);
  int count = 0;
  amrvr_device_count(&count);
  if (count <= 0)
    throw std::runtime_error(
        "amrvr-b200: no CUDA device visible; the B200 render path has no CPU fallback");
  return rank % count;
}

struct ThreadContexts {
  amrvr_ctx* local = nullptr;
  std::map<MPI_Comm, amrvr_ctx*> group;
  ~ThreadContexts() {
    for (auto& entry : group) amrvr_ctx_destroy(entry.second);
    if (local != nullptr) amrvr_ctx_destroy(local);
  }
};
thread_local ThreadContexts t_contexts;

}  // namespace

void check(amrvr_ctx* ctx, int status, const char* where) {
  if (status == AMRVR_OK) return;
  const std::string message = std::string(where) + ": " + amrvr_last_error(ctx);
  if (status == AMRVR_ERR_INVALID) throw std::invalid_argument(message);
  throw std::runtime_error(message);
}

amrvr_ctx* localContext(int rank) {
  if (t_contexts.local == nullptr)
    check(nullptr, amrvr_ctx_create(pickDevice(rank), 0, 1, nullptr, &t_contexts.local),
          "amrvr_ctx_create");
  return t_contexts.local;
}

amrvr_ctx* groupContext(MPI_Comm communicator) {
  auto found = t_contexts.group.find(communicator);
  if (found != t_contexts.group.end()) return found->second;
  int rank = 0, size = 1;
  MPI_Comm_rank(communicator, &rank);
  MPI_Comm_size(communicator, &size);
  // rank 0's NCCL id reaches everyone through an allgather (a broadcast in all but name)
  std::vector<unsigned char> mine(AMRVR_NCCL_ID_BYTES, 0);
  std::vector<unsigned char> all(static_cast<std::size_t>(AMRVR_NCCL_ID_BYTES) * size, 0);
  int status = AMRVR_OK;
  if (rank == 0 && size > 1) status = amrvr_nccl_unique_id(mine.data());
  if (size > 1)
    MPI_Allgather(mine.data(), AMRVR_NCCL_ID_BYTES, MPI_BYTE, all.data(), AMRVR_NCCL_ID_BYTES,
                  MPI_BYTE, communicator);
  check(nullptr, status, "amrvr_nccl_unique_id");
  amrvr_ctx* ctx = nullptr;
  check(nullptr, amrvr_ctx_create(pickDevice(rank), rank, size, size > 1 ? all.data() : nullptr, &ctx),
        "amrvr_ctx_create");
  t_contexts.group[communicator] = ctx;
  return ctx;
}

amrvr_box toAbi(const amrVolumeRenderer::volume::AmrBox& box) {
  amrvr_box out{};
  for (int d = 0; d < 3; ++d) {
    out.min_corner[d] = box.minCorner[d];
    out.max_corner[d] = box.maxCorner[d];
    out.dims[d] = box.cellDimensions[d];
    out.lo[d] = box.validBox.smallEnd()[d];
  }
  out.data = box.values.p;
  out.jstride = box.values.jstride;
  out.kstride = box.values.kstride;
  out.nstride = box.values.nstride;
  out.begin[0] = box.values.begin.x;
  out.begin[1] = box.values.begin.y;
  out.begin[2] = box.values.begin.z;
  out.component = box.component;
  return out;
}

amrvr_scalar_transform toAbi(const amrVolumeRenderer::volume::ScalarTransform& t) {
  amrvr_scalar_transform out{};
  out.log_scale_input = t.logScaleInput ? 1 : 0;
  out.normalize_to_unit_range = t.normalizeToUnitRange ? 1 : 0;
  out.positive_floor = t.positiveFloor;
  out.processed_min = t.processedMin;
  out.processed_max = t.processedMax;
  out.inverse_processed_span = t.inverseProcessedSpan;
  out.normalization_min = t.normalizationMin;
  out.normalization_max = t.normalizationMax;
  out.inverse_normalization_span = t.inverseNormalizationSpan;
  return out;
}

amrvr_camera toAbi(const amrVolumeRenderer::volume::CameraParameters& c) {
  amrvr_camera out{};
  for (int d = 0; d < 3; ++d) {
    out.eye[d] = c.eye[d];
    out.look_at[d] = c.lookAt[d];
    out.up[d] = c.up[d];
  }
  out.fov_y_degrees = c.fovYDegrees;
  out.near_plane = c.nearPlane;
  out.far_plane = c.farPlane;
  return out;
}

}  // namespace amrvr_b200
