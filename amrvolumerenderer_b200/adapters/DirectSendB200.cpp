// DirectSendB200.cpp -- defines DirectSendBase (the reference's Compositor,
// DirectSend/Base/DirectSendBase.hpp:17-45) on top of libamrvr_b200.so.
// Compile it INSTEAD of DirectSend/Base/DirectSendBase.cpp.
//
// Same contract as the reference (DirectSendBase.cpp:285-458): collective over
// `communicator`; `group` carries the visibility order (member i owns image
// piece i); the input image is borrowed and not modified; the result is a
// caller-owned image of the layers' own type holding this rank's fully blended
// piece with regionBegin/regionEnd set, ready for ImageFull::Gather.  The
// R dense DirectSend rounds become one NCCL alltoallv of piece slices plus one
// ordered fold kernel (amrvr_composite_dense).
//
// Only layered images (LayeredImageInterface) are composited here -- the volume
// pipeline passes nothing else (VolumeRenderer.cpp:1228-1249).  The generic
// two-group entry point is kept for link compatibility and reports that.
#include <Common/ImageRGBAFloatColorDepthSort.hpp>
#include <Common/LayeredImageInterface.hpp>
#include <DirectSend/Base/DirectSendBase.hpp>

#include <numeric>
#include <stdexcept>
#include <vector>

#include "B200Runtime.hpp"

DirectSendBase::DirectSendBase() = default;

std::unique_ptr<Image> DirectSendBase::compose(Image* localImage,
                                               MPI_Group group,
                                               MPI_Comm communicator) {
  auto* layered = dynamic_cast<LayeredImageInterface*>(localImage);
  if (layered == nullptr)
    throw std::runtime_error(
        "DirectSendBase (B200): only layered volume images are composited on the GPU path.");
  return this->composeLayered(localImage, *layered, group, communicator);
}

std::unique_ptr<Image> DirectSendBase::compose(Image*, MPI_Group, MPI_Group, MPI_Comm) {
  throw std::runtime_error(
      "DirectSendBase (B200): the generic send/receive-group compose is not part of the "
      "volume path; use compose(image, group, communicator) with a layered image.");
}

std::unique_ptr<Image> DirectSendBase::composeLayered(Image* layeredImage,
                                                      LayeredImageInterface& layers,
                                                      MPI_Group group,
                                                      MPI_Comm communicator) {
  int commSize = 1, groupSize = 0;
  MPI_Comm_size(communicator, &commSize);
  MPI_Group_size(group, &groupSize);
  if (groupSize != commSize)
    throw std::runtime_error("DirectSendBase (B200): the group must span the communicator.");

  // group position -> communicator rank == NCCL rank
  MPI_Group commGroup = MPI_GROUP_NULL;
  MPI_Comm_group(communicator, &commGroup);
  std::vector<int> positions(static_cast<std::size_t>(groupSize)), rankOrder(positions.size());
  std::iota(positions.begin(), positions.end(), 0);
  MPI_Group_translate_ranks(group, groupSize, positions.data(), commGroup, rankOrder.data());
  MPI_Group_free(&commGroup);

  const int width = layeredImage->getWidth(), height = layeredImage->getHeight();
  const int layerCount = layers.getLayerCount();
  std::vector<const float*> buffers;
  std::vector<float> hints;
  for (int l = 0; l < layerCount; ++l) {
    auto* layer = dynamic_cast<ImageRGBAFloatColorDepthSort*>(layers.getLayer(l));
    if (layer == nullptr || layer->getRegionBegin() != 0 ||
        layer->getRegionEnd() != width * height)
      throw std::runtime_error(
          "DirectSendBase (B200): layers must be full-frame ImageRGBAFloatColorDepthSort images.");
    buffers.push_back(layer->getColorBuffer(0));
    hints.push_back(layers.getLayerDepthHint(l));
  }

  amrvr_ctx* ctx = amrvr_b200::groupContext(communicator);
  int begin = 0, end = 0;
  // first call sizes the piece, the result image then receives it in place
  std::vector<float> piece(static_cast<std::size_t>(width) * height / static_cast<std::size_t>(commSize) * 5 +
                           static_cast<std::size_t>(width) * height % static_cast<std::size_t>(commSize) * 5 + 5);
  amrvr_b200::check(ctx,
                    amrvr_composite_dense(ctx, width, height, layerCount, buffers.data(), hints.data(),
                                          rankOrder.data(), &begin, &end, piece.data()),
                    "amrvr_composite_dense");
  std::unique_ptr<Image> result = layers.createEmptyLayer(begin, end);
  auto* typed = dynamic_cast<ImageRGBAFloatColorDepthSort*>(result.get());
  if (typed == nullptr)
    throw std::runtime_error("DirectSendBase (B200): unexpected layer prototype type.");
  std::copy(piece.begin(), piece.begin() + static_cast<std::size_t>(end - begin) * 5,
            typed->getColorBuffer(0));
  return result;
}
