// VolumePainterB200.cpp -- defines VolumePainter::paint with the reference's
// signature (Common/VolumePainter.hpp:20-31) on top of libamrvr_b200.so.
// Compile it INSTEAD of Common/VolumePainter.cpp.
//
// Same contract as the reference (Common/VolumePainter.cpp:548-961): the image
// must be an ImageRGBAFloatColorDepthSort (else std::runtime_error); `bounds`,
// `numProcs` and `antialiasing` are ignored exactly as there; any failure is
// logged as "VolumePainter error on rank N: ..." and rethrown; the call returns
// when the layer is in the caller's host image.  The box is quantised on the
// GPU (brick_quantise_kernel), marched (raymarch_bricks_kernel) and read back
// dense -- the reference's API shape forces the dense host image; the fast
// path that keeps footprint layers on the device is amrvr_render_frame.
#include <Common/ImageRGBAFloatColorDepthSort.hpp>
#include <Common/VolumePainter.hpp>

#include <iostream>
#include <stdexcept>
#include <vector>

#include "B200Runtime.hpp"

VolumePainter::VolumePainter() = default;
VolumePainter::~VolumePainter() = default;

void VolumePainter::paint(const amrVolumeRenderer::volume::AmrBox& box,
                          const amrVolumeRenderer::volume::VolumeBounds& bounds,
                          const amrVolumeRenderer::volume::ScalarTransform& scalarTransform,
                          const std::pair<float, float>& scalarRange,
                          int rank,
                          int numProcs,
                          float boxTransparency,
                          int antialiasing,
                          float referenceSampleDistance,
                          ImageFull& image,
                          const amrVolumeRenderer::volume::CameraParameters& camera,
                          const amrVolumeRenderer::volume::ColorMap* colorMap) {
  static_cast<void>(bounds);
  static_cast<void>(numProcs);
  static_cast<void>(antialiasing);

  auto* target = dynamic_cast<ImageRGBAFloatColorDepthSort*>(&image);
  if (target == nullptr) {
    throw std::runtime_error("VolumePainter expects ImageRGBAFloatColorDepthSort images.");
  }

  amrvr_scene* scene = nullptr;
  amrvr_layers* layers = nullptr;
  amrvr_ctx* ctx = nullptr;
  try {
    ctx = amrvr_b200::localContext(rank);
    const amrvr_box abiBox = amrvr_b200::toAbi(box);
    const amrvr_scalar_transform abiTransform = amrvr_b200::toAbi(scalarTransform);
    const amrvr_camera abiCamera = amrvr_b200::toAbi(camera);
    amrvr_b200::check(ctx,
                      amrvr_scene_upload(ctx, 1, &abiBox, &abiTransform, scalarRange.first,
                                         scalarRange.second, &scene),
                      "amrvr_scene_upload");
    std::vector<amrvr_colormap_pt> points;
    if (colorMap != nullptr)
      for (const auto& p : *colorMap)
        points.push_back(amrvr_colormap_pt{p.value, p.red, p.green, p.blue, p.alpha});
    amrvr_b200::check(ctx,
                      amrvr_raymarch(ctx, scene, &abiCamera, image.getWidth(), image.getHeight(),
                                     boxTransparency, referenceSampleDistance,
                                     points.empty() ? nullptr : points.data(),
                                     static_cast<int>(points.size()), 0, &layers),
                      "amrvr_raymarch");
    // the reference fills the whole frame: transparent, +inf depth outside the box
    if (target->getRegionBegin() != 0 ||
        target->getRegionEnd() != image.getWidth() * image.getHeight())
      throw std::runtime_error("VolumePainter expects a full-frame image.");
    amrvr_b200::check(ctx, amrvr_layers_read(ctx, layers, 0, target->getColorBuffer(0)),
                      "amrvr_layers_read");
    amrvr_layers_free(ctx, layers);
    amrvr_scene_free(ctx, scene);
  } catch (const std::exception& error) {
    if (layers != nullptr) amrvr_layers_free(ctx, layers);
    if (scene != nullptr) amrvr_scene_free(ctx, scene);
    std::cerr << "VolumePainter error on rank " << rank << ": " << error.what() << std::endl;
    throw;
  }
}
