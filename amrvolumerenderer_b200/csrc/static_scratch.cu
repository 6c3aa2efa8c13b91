// amrvr_brick_stats_static: the K2 statistics pass (brick_stats_kernel + brick_stats_final_kernel,
// kernels.cu) for callers that must not allocate device memory.
//
// Who that is.  amrvr_brick_stats needs an amrvr_ctx, and a ctx grows device buffers on demand
// (cudaMalloc for the per-CTA partial records, a staging buffer for host FABs).  An in-situ host such
// as an AMReX simulation usually cannot afford that: its arena takes nearly all of the GPU's memory
// when the run starts, so a cudaMalloc issued by a library in the middle of a time step either fails
// for lack of memory or, with the cudaFree that has to follow it, makes the whole device wait.
// Codes that forbid allocation inside their time loop for that reason (or that run this call from
// a place where the CUDA allocator may not be entered at all) get the same numbers here from
// memory that exists before they call: the partial records live in the file-scope
// __device__ array below, which the CUDA runtime places on every GPU that loads this library and
// keeps there until the process exits (about 2 MiB per GPU, used or not).
//
// What follows from one shared array per GPU:
//   * sizes are bounded: at most AMRVR_BRICK_STATS_STATIC_MAX_BOXES boxes per call; the CTAs per box
//     shrink so that all partial records fit (never below 16 per box);
//   * the array keeps whatever the previous call wrote.  Nothing in it has to start at zero: every
//     record the final kernel reads was written by a CTA of the same call, and the result slot is
//     written before it is copied out -- so nothing is cleared, and stale records are never read;
//   * two calls that are on the same GPU at the same time (different streams or different threads)
//     write the same records and race.  One call at a time per GPU; the call itself waits for its
//     stream before it returns, so back-to-back calls from one thread are safe.
//   * the first call on a GPU asks the runtime where the array lives on that GPU
//     (cudaGetSymbolAddress -- the array's name in host code is only a handle, not a device
//     address) and remembers the answer.  Make that first call outside any stream capture.
#include <cuda_runtime.h>

#include <algorithm>
#include <cstring>
#include <mutex>
#include <string>

#include "../../include/amrvr.h"
#include "internal.hpp"

namespace {

constexpr int kMaxParts = 65536;     // partial records (4 doubles each): 2 MiB
constexpr int kCtasPerBox = 148 * 2;  // what amrvr_brick_stats launches per box
constexpr int kMaxDevices = 64;

// [0, 4*kMaxParts): one {min, max, min positive, finite count} record per CTA;
// [4*kMaxParts, 4*kMaxParts + 4): the reduced result of the call.
static __device__ __align__(256) double g_stats_scratch[4 * static_cast<size_t>(kMaxParts) + 4];

std::mutex g_setup_lock;
double* g_scratch_on_device[kMaxDevices] = {};  // device address of g_stats_scratch, per GPU

int cuda_fail(const char* what, cudaError_t e) {
  return amrvr::report_error(AMRVR_ERR_CUDA, std::string(what) + ": " + cudaGetErrorString(e));
}

// Device address of the scratch array on `device`, which must be the current device.
int scratch_address(int device, double** out) {
  std::lock_guard<std::mutex> hold(g_setup_lock);
  if (g_scratch_on_device[device] == nullptr) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_stats_scratch);
    if (e != cudaSuccess) return cuda_fail("cudaGetSymbolAddress(g_stats_scratch)", e);
    if (p == nullptr)
      return amrvr::report_error(AMRVR_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_scratch_on_device[device] = p;
  }
  *out = g_scratch_on_device[device];
  return AMRVR_OK;
}

}  // namespace

extern "C" int amrvr_brick_stats_static(int cuda_device, void* stream, int nboxes,
                                        const amrvr_box* boxes, double* min_value,
                                        double* max_value, double* min_positive,
                                        int64_t* finite_count) {
  using amrvr::report_error;
  if (nboxes < 0 || (nboxes > 0 && boxes == nullptr))
    return report_error(AMRVR_ERR_INVALID, "amrvr_brick_stats_static: bad box list");
  if (nboxes > AMRVR_BRICK_STATS_STATIC_MAX_BOXES)
    return report_error(AMRVR_ERR_INVALID,
                        "amrvr_brick_stats_static: " + std::to_string(nboxes) +
                            " boxes exceed the fixed scratch array (at most " +
                            std::to_string(AMRVR_BRICK_STATS_STATIC_MAX_BOXES) +
                            " per call); split the list or use amrvr_brick_stats");
  if (cuda_device < 0 || cuda_device >= kMaxDevices)
    return report_error(AMRVR_ERR_INVALID, "amrvr_brick_stats_static: device index out of range");
  cudaError_t e = cudaSetDevice(cuda_device);
  if (e != cudaSuccess) return cuda_fail("cudaSetDevice", e);

  // The FABs must already be where a kernel can read them: staging a host FAB takes a buffer.
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    if (B.dims[0] <= 0 || B.dims[1] <= 0 || B.dims[2] <= 0) continue;
    if (B.data == nullptr)
      return report_error(AMRVR_ERR_INVALID, "amrvr_brick_stats_static: box data pointer is NULL");
    cudaPointerAttributes attr;
    e = cudaPointerGetAttributes(&attr, B.data);
    if (e != cudaSuccess) (void)cudaGetLastError();
    if (e != cudaSuccess ||
        !(attr.type == cudaMemoryTypeDevice || attr.type == cudaMemoryTypeManaged))
      return report_error(AMRVR_ERR_INVALID,
                          "amrvr_brick_stats_static: box " + std::to_string(b) +
                              " is not in device memory; host FABs need a staging buffer, which this "
                              "entry point may not allocate (use amrvr_brick_stats)");
  }

  double* scratch = nullptr;
  const int st = scratch_address(cuda_device, &scratch);
  if (st != AMRVR_OK) return st;

  const int ctas_cap = std::min(kCtasPerBox, kMaxParts / std::max(nboxes, 1));
  int64_t launches = 0;
  int nparts = 0;
  for (int b = 0; b < nboxes; ++b) {
    const amrvr_box& B = boxes[b];
    if (B.dims[0] <= 0 || B.dims[1] <= 0 || B.dims[2] <= 0) continue;
    const long long cells = 1LL * B.dims[0] * B.dims[1] * B.dims[2];
    const int ctas = static_cast<int>(std::min<long long>(ctas_cap, (cells + 255) / 256));
    // first valid cell of the Array4 view, as amrvr_brick_stats addresses a device FAB
    const double* first = B.data + (B.lo[0] - B.begin[0]) + (B.lo[1] - B.begin[1]) * B.jstride +
                          (B.lo[2] - B.begin[2]) * B.kstride + B.component * B.nstride;
    const int le = amrvr::launch_brick_stats(first, B.jstride, B.kstride, B.dims[0], B.dims[1],
                                             B.dims[2], scratch + 4 * static_cast<size_t>(nparts),
                                             ctas, stream, &launches);
    if (le != 0) return cuda_fail("brick_stats_kernel", static_cast<cudaError_t>(le));
    nparts += ctas;
  }
  double* out4 = scratch + 4 * static_cast<size_t>(kMaxParts);
  const int le = amrvr::launch_brick_stats_final(scratch, nparts, out4, stream, &launches);
  if (le != 0) return cuda_fail("brick_stats_final_kernel", static_cast<cudaError_t>(le));

  double host4[4];
  e = cudaMemcpyAsync(host4, out4, sizeof(host4), cudaMemcpyDeviceToHost,
                      static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail("cudaMemcpyAsync(result)", e);
  e = cudaStreamSynchronize(static_cast<cudaStream_t>(stream));
  if (e != cudaSuccess) return cuda_fail("cudaStreamSynchronize", e);
  if (min_value) *min_value = host4[0];
  if (max_value) *max_value = host4[1];
  if (min_positive) *min_positive = host4[2];
  if (finite_count) std::memcpy(finite_count, &host4[3], 8);
  return AMRVR_OK;
}
