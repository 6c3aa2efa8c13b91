// B200Runtime.hpp -- glue shared by the C++ adapters that keep the reference's
// own entry points (VolumePainter::paint, BuildVisibilityOrderedGroup,
// DirectSendBase::compose) and forward to the C ABI of libamrvr_b200.so.
//
// These files are compiled INSIDE the reference's tree (they include the
// reference's headers) in place of Common/VolumePainter.cpp,
// Common/VisibilityOrdering.cpp and DirectSend/Base/DirectSendBase.cpp; see
// INTEGRATION.md.  They hold no arithmetic of their own.
#ifndef AMRVR_B200_RUNTIME_HPP
#define AMRVR_B200_RUNTIME_HPP

#include <Common/VolumeTypes.hpp>
#include <mpi.h>

#include <amrvr.h>

namespace amrvr_b200 {

// Single-rank context of the calling thread (one per rank: MPI ranks are
// processes in production, threads under the test shim).  The GPU is
// AMRVR_DEVICE when set, else rank % visible devices.  Throws
// std::runtime_error when no sm_100 GPU is usable -- there is no CPU fallback.
amrvr_ctx* localContext(int rank);

// Context whose NCCL communicator spans `communicator` (NCCL rank == MPI rank).
// The NCCL id travels over MPI.  Collective on first use per communicator.
amrvr_ctx* groupContext(MPI_Comm communicator);

// Status of an ABI call -> the exception type the reference would have thrown:
// AMRVR_ERR_INVALID -> std::invalid_argument, anything else -> std::runtime_error.
void check(amrvr_ctx* ctx, int status, const char* where);

amrvr_box toAbi(const amrVolumeRenderer::volume::AmrBox& box);
amrvr_scalar_transform toAbi(const amrVolumeRenderer::volume::ScalarTransform& transform);
amrvr_camera toAbi(const amrVolumeRenderer::volume::CameraParameters& camera);

}  // namespace amrvr_b200

#endif  // AMRVR_B200_RUNTIME_HPP
