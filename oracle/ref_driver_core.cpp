// oracle/ref_driver.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// C entry points (for ctypes) around the reference's OWN compiled sources
// (linked in from /root/reference by oracle/Makefile, never copied):
//   ref_paint          -> VolumePainter::paint            (Common/VolumePainter.cpp:548-961)
//   ref_order          -> BuildVisibilityOrderedGroup     (Common/VisibilityOrdering.cpp:63-632)
//   ref_compose        -> DirectSendBase::compose + Gather (DirectSend/Base/DirectSendBase.cpp:285-458,
//                                                           Common/ImageColorOnly.hpp:220-270)
//   ref_render_scene   -> VolumeRenderer::renderScene     (VolumeRenderer/VolumeRenderer.cpp:947-1101)
//                         i.e. the real renderSingleTrial incl. depth hints, downsample,
//                         bounding-box overlay and SavePPM byte quantisation
//   ref_build_scene    -> detail::BuildSceneGeometry      (VolumeRenderer/SceneBuilder.cpp:113-425)
// Ranks are threads of this process (shim/fake_mpi.cpp).  Only tests/,
// __graft_entry__.smoke() and bench.py's CPU-baseline legs may call this.
#include <AMReX_MultiFab.H>
#include <mpi.h>

#include <Common/ImageRGBAFloatColorDepthSort.hpp>
#include <Common/LayeredVolumeImage.hpp>
#include <Common/SavePPM.hpp>
#include <Common/VisibilityOrdering.hpp>
#include <Common/VolumePainter.hpp>
#include <Common/VolumeTypes.hpp>
#include <DirectSend/Base/DirectSendBase.hpp>
#include <VolumeRenderer/SceneBuilder.hpp>
#include <VolumeRenderer/VolumeRenderer.hpp>

#include "../include/amrvr.h"

#include <chrono>
#include <cstring>
#include <iostream>
#include <memory>
#include <mutex>
#include <optional>
#include <sstream>
#include <string>
#include <vector>

using amrVolumeRenderer::volume::AmrBox;
using amrVolumeRenderer::volume::CameraParameters;
using amrVolumeRenderer::volume::ColorMap;
using amrVolumeRenderer::volume::ColorMapControlPoint;
using amrVolumeRenderer::volume::ScalarTransform;
using amrVolumeRenderer::volume::VolumeBounds;

// The reference's SavePNG needs libpng (absent here).  VolumeRenderer.cpp
// references the symbol; the oracle only ever asks for ".ppm", and a ".png"
// request is served by the reference's own PPM writer (same bytes rule,
// Common/Color.hpp:86-90, same row order).
bool SavePNG(const Image& image, const std::string& filename) {
  return SavePPM(image, filename);
}

namespace {

thread_local std::string g_error;

AmrBox toAmrBox(const amrvr_box& b) {
  AmrBox box;
  box.minCorner =
      amrex::RealVect(b.min_corner[0], b.min_corner[1], b.min_corner[2]);
  box.maxCorner =
      amrex::RealVect(b.max_corner[0], b.max_corner[1], b.max_corner[2]);
  box.cellDimensions = amrex::IntVect(b.dims[0], b.dims[1], b.dims[2]);
  box.validBox = amrex::Box(
      amrex::IntVect(b.lo[0], b.lo[1], b.lo[2]),
      amrex::IntVect(b.lo[0] + b.dims[0] - 1, b.lo[1] + b.dims[1] - 1,
                     b.lo[2] + b.dims[2] - 1));
  amrex::Array4<amrex::Real const> view;
  view.p = b.data;
  view.jstride = b.jstride;
  view.kstride = b.kstride;
  view.nstride = b.nstride;
  view.begin = amrex::Dim3{b.begin[0], b.begin[1], b.begin[2]};
  view.end = amrex::Dim3{b.lo[0] + b.dims[0], b.lo[1] + b.dims[1],
                         b.lo[2] + b.dims[2]};
  view.ncomp = b.component + 1;
  box.values = view;
  box.component = b.component;
  return box;
}

ScalarTransform toTransform(const amrvr_scalar_transform& t) {
  ScalarTransform s;
  s.logScaleInput = t.log_scale_input != 0;
  s.normalizeToUnitRange = t.normalize_to_unit_range != 0;
  s.positiveFloor = t.positive_floor;
  s.processedMin = t.processed_min;
  s.processedMax = t.processed_max;
  s.inverseProcessedSpan = t.inverse_processed_span;
  s.normalizationMin = t.normalization_min;
  s.normalizationMax = t.normalization_max;
  s.inverseNormalizationSpan = t.inverse_normalization_span;
  return s;
}

void fromTransform(const ScalarTransform& s, amrvr_scalar_transform* t) {
  t->log_scale_input = s.logScaleInput ? 1 : 0;
  t->normalize_to_unit_range = s.normalizeToUnitRange ? 1 : 0;
  t->positive_floor = s.positiveFloor;
  t->processed_min = s.processedMin;
  t->processed_max = s.processedMax;
  t->inverse_processed_span = s.inverseProcessedSpan;
  t->normalization_min = s.normalizationMin;
  t->normalization_max = s.normalizationMax;
  t->inverse_normalization_span = s.inverseNormalizationSpan;
}

CameraParameters toCamera(const amrvr_camera& c) {
  return CameraParameters{
      amrex::RealVect(c.eye[0], c.eye[1], c.eye[2]),
      amrex::RealVect(c.look_at[0], c.look_at[1], c.look_at[2]),
      amrex::RealVect(c.up[0], c.up[1], c.up[2]),
      c.fov_y_degrees,
      c.near_plane,
      c.far_plane};
}

std::optional<ColorMap> toColorMap(const amrvr_colormap_pt* pts, int npts) {
  if (pts == nullptr || npts <= 0) return std::nullopt;
  ColorMap map;
  for (int i = 0; i < npts; ++i)
    map.push_back(ColorMapControlPoint{pts[i].value, pts[i].red, pts[i].green,
                                       pts[i].blue, pts[i].alpha});
  return map;
}

template <class F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::invalid_argument& e) {
    g_error = e.what();
    return AMRVR_ERR_INVALID;
  } catch (const std::exception& e) {
    g_error = e.what();
    return AMRVR_ERR_RUNTIME;
  }
}

// Redirect std::cout/cerr into a buffer for the duration of a call (the
// reference prints its stage timings on rank 0 only).
struct CaptureStdout {
  std::ostringstream buffer;
  std::streambuf* oldOut;
  std::streambuf* oldErr;
  CaptureStdout() {
    oldOut = std::cout.rdbuf(buffer.rdbuf());
    oldErr = std::cerr.rdbuf(buffer.rdbuf());
  }
  ~CaptureStdout() {
    std::cout.rdbuf(oldOut);
    std::cerr.rdbuf(oldErr);
  }
};

void copyLog(const std::string& text, char* log, int logCapacity) {
  if (log == nullptr || logCapacity <= 0) return;
  const std::size_t n =
      std::min(text.size(), static_cast<std::size_t>(logCapacity - 1));
  std::memcpy(log, text.data(), n);
  log[n] = '\0';
}

}  // namespace

extern "C" {

const char* ref_last_error() { return g_error.c_str(); }
void ref_set_error(const char* message) { g_error = message ? message : ""; }

// One box -> one dense layer (W*H*5 floats) through the real painter.
int ref_paint(const amrvr_box* box, const amrvr_scalar_transform* transform,
              float rangeMin, float rangeMax, float boxTransparency,
              float referenceSampleDistance, int width, int height,
              const amrvr_camera* camera, const amrvr_colormap_pt* pts,
              int npts, float* out5, long long* samples) {
  return guarded([&] {
    CaptureStdout quiet;
    const AmrBox amrBox = toAmrBox(*box);
    VolumeBounds bounds{amrBox.minCorner, amrBox.maxCorner};
    const ScalarTransform st = toTransform(*transform);
    const CameraParameters cam = toCamera(*camera);
    const std::optional<ColorMap> map = toColorMap(pts, npts);
    ImageRGBAFloatColorDepthSort image(width, height);
    VolumePainter painter;
    const long long before = amrex::shim::array4_reads;
    painter.paint(amrBox, bounds, st, {rangeMin, rangeMax}, 0, 1,
                  boxTransparency, 1, referenceSampleDistance, image, cam,
                  map ? &*map : nullptr);
    if (samples) *samples = amrex::shim::array4_reads - before;
    std::memcpy(out5, image.getColorBuffer(0),
                sizeof(float) * 5 * static_cast<std::size_t>(width) * height);
  });
}

// rank_order[i] = world rank at position i of the ordered group, as produced
// by the real BuildVisibilityOrderedGroup running collectively on `nranks`
// thread-ranks, each holding its own boxes of the global table.
static int refOrderImpl(int nranks, int nglobal,
                        const amrvr_box_record* records,
                        const amrvr_camera* camera, float aspect, int useGraph,
                        bool writeGraph, int* rankOrder, char* log,
                        int logCapacity);

int ref_order(int nranks, int nglobal, const amrvr_box_record* records,
              const amrvr_camera* camera, float aspect, int useGraph,
              int* rankOrder, char* log, int logCapacity) {
  return refOrderImpl(nranks, nglobal, records, camera, aspect, useGraph, false,
                      rankOrder, log, logCapacity);
}

// The same call with writeVisibilityGraph = true: rank 0 writes
// visibility_graph_<n>.dot into the CURRENT DIRECTORY, n from a counter that
// lives as long as this library is loaded (VisibilityOrdering.cpp:319-352).
int ref_order_write_graph(int nranks, int nglobal,
                          const amrvr_box_record* records,
                          const amrvr_camera* camera, float aspect,
                          int* rankOrder, char* log, int logCapacity) {
  return refOrderImpl(nranks, nglobal, records, camera, aspect, 1, true,
                      rankOrder, log, logCapacity);
}

static int refOrderImpl(int nranks, int nglobal,
                        const amrvr_box_record* records,
                        const amrvr_camera* camera, float aspect, int useGraph,
                        bool writeGraph, int* rankOrder, char* log,
                        int logCapacity) {
  return guarded([&] {
    CaptureStdout capture;
    const CameraParameters cam = toCamera(*camera);
    std::vector<std::vector<int>> orders(nranks);
    fake_mpi_run(nranks, [&](int rank) {
      std::vector<AmrBox> local;
      for (int i = 0; i < nglobal; ++i) {
        if (records[i].owner != rank) continue;
        AmrBox box;
        box.minCorner = amrex::RealVect(records[i].min_corner[0],
                                        records[i].min_corner[1],
                                        records[i].min_corner[2]);
        box.maxCorner = amrex::RealVect(records[i].max_corner[0],
                                        records[i].max_corner[1],
                                        records[i].max_corner[2]);
        box.cellDimensions = amrex::IntVect(
            records[i].dims[0], records[i].dims[1], records[i].dims[2]);
        local.push_back(box);
      }
      MPI_Group world = MPI_GROUP_NULL;
      MPI_Comm_group(MPI_COMM_WORLD, &world);
      MPI_Group ordered = BuildVisibilityOrderedGroup(
          cam, aspect, world, rank, nranks, useGraph != 0, writeGraph, local,
          MPI_COMM_WORLD);
      std::vector<int> positions(nranks), translated(nranks);
      for (int i = 0; i < nranks; ++i) positions[i] = i;
      MPI_Group_translate_ranks(ordered, nranks, positions.data(), world,
                                translated.data());
      orders[rank] = translated;
      MPI_Group_free(&ordered);
      MPI_Group_free(&world);
    });
    for (int r = 1; r < nranks; ++r)
      if (orders[r] != orders[0])
        throw std::runtime_error("ref_order: ranks disagree on the order");
    std::copy(orders[0].begin(), orders[0].end(), rankOrder);
    copyLog(capture.buffer.str(), log, logCapacity);
  });
}

// Sort-last composite of dense layers through the real DirectSendBase, then
// the real Gather to rank 0.  layerOwner[l] / layerData[l] / depthHints[l]
// describe layer l; layers of one owner keep their relative order (= local
// index).  pieceOwner (may be NULL) receives, per pixel, the world rank whose
// composited piece held it.
int ref_compose(int nranks, int width, int height, int nlayers,
                const int* layerOwner, const float* const* layerData,
                const float* depthHints, const int* rankOrder, float* full5,
                int* pieceOwner, double* seconds) {
  return guarded([&] {
    CaptureStdout quiet;
    const std::size_t pixelFloats =
        static_cast<std::size_t>(width) * height * 5;
    std::vector<double> elapsed(nranks, 0.0);
    std::mutex ownerMutex;
    fake_mpi_run(nranks, [&](int rank) {
      std::vector<std::unique_ptr<ImageRGBAFloatColorDepthSort>> layers;
      std::vector<float> hints;
      for (int l = 0; l < nlayers; ++l) {
        if (layerOwner[l] != rank) continue;
        auto image =
            std::make_unique<ImageRGBAFloatColorDepthSort>(width, height);
        std::memcpy(image->getColorBuffer(0), layerData[l],
                    sizeof(float) * pixelFloats);
        layers.push_back(std::move(image));
        hints.push_back(depthHints[l]);
      }
      auto prototype =
          std::make_unique<ImageRGBAFloatColorDepthSort>(width, height);
      LayeredVolumeImage layered(width, height, std::move(layers),
                                 std::move(hints), std::move(prototype));
      MPI_Group world = MPI_GROUP_NULL;
      MPI_Comm_group(MPI_COMM_WORLD, &world);
      MPI_Group ordered = MPI_GROUP_NULL;
      MPI_Group_incl(world, nranks, rankOrder, &ordered);

      MPI_Barrier(MPI_COMM_WORLD);
      const auto t0 = std::chrono::steady_clock::now();
      DirectSendBase compositor;
      Compositor* asInterface = &compositor;
      std::unique_ptr<Image> piece =
          asInterface->compose(&layered, ordered, MPI_COMM_WORLD);
      const auto t1 = std::chrono::steady_clock::now();
      elapsed[rank] = std::chrono::duration<double>(t1 - t0).count();

      auto* full = dynamic_cast<ImageFull*>(piece.get());
      if (full == nullptr)
        throw std::runtime_error("ref_compose: composite is not an ImageFull");
      if (pieceOwner) {
        std::lock_guard<std::mutex> lock(ownerMutex);
        for (int p = full->getRegionBegin(); p < full->getRegionEnd(); ++p)
          pieceOwner[p] = rank;
      }
      std::unique_ptr<ImageFull> gathered = full->Gather(0, MPI_COMM_WORLD);
      if (rank == 0) {
        auto* typed =
            dynamic_cast<ImageRGBAFloatColorDepthSort*>(gathered.get());
        std::memcpy(full5, typed->getColorBuffer(0),
                    sizeof(float) * pixelFloats);
      }
      MPI_Group_free(&ordered);
      MPI_Group_free(&world);
    });
    if (seconds) {
      *seconds = 0.0;
      for (double s : elapsed) *seconds = std::max(*seconds, s);
    }
  });
}

}  // extern "C"
