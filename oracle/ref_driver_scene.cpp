// oracle/ref_driver_scene.cpp -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
//
// Scene-level C entry points around the reference's own compiled sources:
//   ref_render_scene -> VolumeRenderer::renderScene     (VolumeRenderer.cpp:947-1101 -> :1103-1339)
//   ref_build_scene  -> detail::BuildSceneGeometry      (SceneBuilder.cpp:113-425)
//   ref_api_render   -> api::Render                     (VolumeRendererApi.cpp:257-395), single level
//   ref_api_histogram-> api::ComputeHistogram           (VolumeRendererApi.cpp:397-415), single level
// See ref_driver_core.cpp for the conventions.
#include <AMReX_Geometry.H>
#include <AMReX_MultiFab.H>
#include <mpi.h>

#include <Common/VolumeTypes.hpp>
#include <VolumeRenderer/SceneBuilder.hpp>
#include <VolumeRenderer/VolumeRenderer.hpp>
#include <VolumeRenderer/VolumeRendererApi.hpp>

#include "../include/amrvr.h"

#include <chrono>
#include <cstring>
#include <iostream>
#include <memory>
#include <optional>
#include <sstream>
#include <string>
#include <vector>

using amrVolumeRenderer::volume::AmrBox;
using amrVolumeRenderer::volume::CameraParameters;
using amrVolumeRenderer::volume::ColorMap;
using amrVolumeRenderer::volume::ColorMapControlPoint;
using amrVolumeRenderer::volume::ScalarTransform;

extern "C" const char* ref_last_error();

namespace {

// Shared with ref_driver_core.cpp through ref_set_error().
}  // namespace

extern "C" void ref_set_error(const char* message);

namespace {

struct Capture {
  std::ostringstream buffer;
  std::streambuf* oldOut;
  std::streambuf* oldErr;
  Capture() {
    oldOut = std::cout.rdbuf(buffer.rdbuf());
    oldErr = std::cerr.rdbuf(buffer.rdbuf());
  }
  ~Capture() {
    std::cout.rdbuf(oldOut);
    std::cerr.rdbuf(oldErr);
  }
};

void copyLog(const std::string& text, char* log, int capacity) {
  if (log == nullptr || capacity <= 0) return;
  const std::size_t n =
      std::min(text.size(), static_cast<std::size_t>(capacity - 1));
  std::memcpy(log, text.data(), n);
  log[n] = '\0';
}

template <class F>
int guarded(F&& f) {
  try {
    f();
    return 0;
  } catch (const std::invalid_argument& e) {
    ref_set_error(e.what());
    return AMRVR_ERR_INVALID;
  } catch (const std::exception& e) {
    ref_set_error(e.what());
    return AMRVR_ERR_RUNTIME;
  }
}

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

void fromAmrBox(const AmrBox& box, amrvr_box* b) {
  for (int d = 0; d < 3; ++d) {
    b->min_corner[d] = box.minCorner[d];
    b->max_corner[d] = box.maxCorner[d];
    b->dims[d] = box.cellDimensions[d];
    b->lo[d] = box.validBox.smallEnd()[d];
  }
  b->data = box.values.p;
  b->jstride = box.values.jstride;
  b->kstride = box.values.kstride;
  b->nstride = box.values.nstride;
  b->begin[0] = box.values.begin.x;
  b->begin[1] = box.values.begin.y;
  b->begin[2] = box.values.begin.z;
  b->component = box.component;
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

VolumeRenderer::RenderParameters toParameters(const amrvr_frame_params& p) {
  VolumeRenderer::RenderParameters r;
  r.width = p.width;
  r.height = p.height;
  r.boxTransparency = p.box_transparency;
  r.antialiasing = p.antialiasing;
  r.useVisibilityGraph = p.use_visibility_graph != 0;
  r.writeVisibilityGraph = false;
  return r;
}

}  // namespace

// A level of a block-structured AMR hierarchy as flat arrays (one process
// describes ALL ranks' boxes; `owner` says which thread-rank holds each FAB).
struct ref_level {
  double prob_lo[3];
  double cell_size[3];
  int32_t nboxes;
  int32_t ncomp;
  const int32_t* lo;        /* [nboxes*3] */
  const int32_t* hi;        /* [nboxes*3] inclusive */
  const int32_t* owner;     /* [nboxes] */
  const double* const* data; /* [nboxes] Fortran order, ncomp components */
};

// Result of the real BuildSceneGeometry for every rank, kept alive so the
// amrvr_box::data pointers stay valid.
struct ref_scene {
  int nranks = 1;
  std::vector<VolumeRenderer::SceneGeometry> perRank;
  std::vector<amrvr_box> boxes;  // owner-major, owner-local order
  std::vector<int32_t> owners;
};

extern "C" {

// The real renderScene on `nranks` thread-ranks.  camera == NULL selects the
// reference's automatic camera (VolumeRenderer.cpp:974-1023, mt19937(91021)).
// Writes outPath on rank 0 (".ppm").  samples: Array4 reads summed over ranks.
int ref_render_scene(int nranks, int nboxes, const amrvr_box* boxes,
                     const int32_t* owners,
                     const amrvr_scalar_transform* transform, float rangeMin,
                     float rangeMax, int hasBounds, const double* boundsMin,
                     const double* boundsMax, const amrvr_camera* camera,
                     const amrvr_frame_params* params,
                     const amrvr_colormap_pt* pts, int npts,
                     const char* outPath, int pinCores, long long* samples,
                     double* wallSeconds, char* log, int logCapacity) {
  return guarded([&] {
    Capture capture;
    const std::optional<ColorMap> map = toColorMap(pts, npts);
    const VolumeRenderer::RenderParameters parameters = toParameters(*params);
    std::vector<long long> reads(nranks, 0);
    const auto t0 = std::chrono::steady_clock::now();
    fake_mpi_run(
        nranks,
        [&](int rank) {
          VolumeRenderer::SceneGeometry geometry;
          for (int i = 0; i < nboxes; ++i)
            if (owners[i] == rank) geometry.localBoxes.push_back(toAmrBox(boxes[i]));
          if (hasBounds) {
            geometry.explicitBounds.minCorner =
                amrex::RealVect(boundsMin[0], boundsMin[1], boundsMin[2]);
            geometry.explicitBounds.maxCorner =
                amrex::RealVect(boundsMax[0], boundsMax[1], boundsMax[2]);
            geometry.hasExplicitBounds = true;
          }
          geometry.scalarRange = {rangeMin, rangeMax};
          geometry.hasScalarRange = true;
          geometry.scalarTransform = toTransform(*transform);
          const long long before = amrex::shim::array4_reads;
          VolumeRenderer renderer;
          int status = 0;
          if (camera != nullptr) {
            status = renderer.renderScene(outPath, parameters, geometry,
                                          toCamera(*camera), map);
          } else {
            status = renderer.renderScene(outPath, parameters, geometry, map);
          }
          reads[rank] = amrex::shim::array4_reads - before;
          if (status != 0)
            throw std::runtime_error("renderScene returned non-zero");
        },
        pinCores != 0);
    const auto t1 = std::chrono::steady_clock::now();
    if (wallSeconds)
      *wallSeconds = std::chrono::duration<double>(t1 - t0).count();
    if (samples) {
      *samples = 0;
      for (long long r : reads) *samples += r;
    }
    copyLog(capture.buffer.str(), log, logCapacity);
  });
}

// Registers the content of a plotfile under `path` for the shim's PlotFileData
// (oracle/shim/amrex_shim.hpp); owners in `levels` are ignored -- get() deals
// boxes to thread-ranks itself.
int ref_register_plotfile(const char* path, int nlevels, const ref_level* levels,
                          int nnames, const char* const* names) {
  return guarded([&] {
    amrex::shim::PlotContent content;
    for (int c = 0; c < nnames; ++c) content.names.push_back(names[c]);
    for (int l = 0; l < nlevels; ++l) {
      const ref_level& level = levels[l];
      if (level.ncomp != nnames) throw std::invalid_argument("component count differs from names");
      amrex::shim::PlotLevel out;
      for (int d = 0; d < 3; ++d) {
        out.cellSize[d] = level.cell_size[d];
        content.probLo[d] = level.prob_lo[d];
      }
      for (int b = 0; b < level.nboxes; ++b) {
        out.boxes.emplace_back(
            amrex::IntVect(level.lo[3 * b], level.lo[3 * b + 1], level.lo[3 * b + 2]),
            amrex::IntVect(level.hi[3 * b], level.hi[3 * b + 1], level.hi[3 * b + 2]));
        const std::size_t count =
            static_cast<std::size_t>(out.boxes.back().numPts()) * level.ncomp;
        out.data.emplace_back(level.data[b], level.data[b] + count);
      }
      content.levels.push_back(std::move(out));
    }
    amrex::shim::plotfiles()[path] = std::move(content);
  });
}

// Installs (nlevels > 0) or clears (nlevels == 0) the box cuts the shim's
// convexify hands back: boxes[l] = counts[l] boxes as lo[3], hi[3].
int ref_set_convexify_override(int nlevels, const int32_t* counts,
                               const int32_t* const* boxes) {
  return guarded([&] {
    amrex::shim::ConvexOverride& cuts = amrex::shim::convex_override();
    cuts.boxes.clear();
    cuts.active = nlevels > 0;
    for (int l = 0; l < nlevels; ++l) {
      std::vector<amrex::Box> list;
      for (int b = 0; b < counts[l]; ++b)
        list.emplace_back(
            amrex::IntVect(boxes[l][6 * b], boxes[l][6 * b + 1], boxes[l][6 * b + 2]),
            amrex::IntVect(boxes[l][6 * b + 3], boxes[l][6 * b + 4], boxes[l][6 * b + 5]));
      cuts.boxes.push_back(std::move(list));
    }
  });
}

// The reference's own VolumeRenderer::run(RunOptions) (VolumeRenderer.cpp:
// 1341-1576) on `nranks` thread-ranks over a registered plotfile: argument
// checks, loadPlotFileGeometry, scalar-range override, colour-map
// normalisation, automatic or explicit camera, renderScene, image file.
// customUp / scalarRange / camera / pts may be NULL (option not given).
int ref_run(int nranks, const char* plotfile, const amrvr_frame_params* params,
            const char* output, const char* variable, int minLevel, int maxLevel,
            int logScale, const double* customUp, const float* scalarRange,
            const amrvr_camera* camera, const amrvr_colormap_pt* pts, int npts,
            char* log, int logCapacity) {
  std::string text;
  const int status = guarded([&] {
    Capture capture;
    try {
      fake_mpi_run(
          nranks,
          [&](int) {
            VolumeRenderer::RunOptions options;
            options.parameters = toParameters(*params);
            if (customUp != nullptr) {
              options.parameters.cameraUp = amrex::RealVect(customUp[0], customUp[1], customUp[2]);
              options.parameters.useCustomUp = true;
            }
            options.outputFilename = output;
            options.plotfilePath = plotfile;
            options.variableName = (variable != nullptr) ? variable : "";
            options.minLevel = minLevel;
            options.maxLevel = maxLevel;
            options.logScaleInput = logScale != 0;
            if (scalarRange != nullptr)
              options.scalarRange = std::make_pair(scalarRange[0], scalarRange[1]);
            if (camera != nullptr) options.camera = toCamera(*camera);
            options.colorMap = toColorMap(pts, npts);
            VolumeRenderer renderer;
            if (renderer.run(options) != 0) throw std::runtime_error("run returned non-zero");
          },
          false);
    } catch (...) {
      text = capture.buffer.str();
      throw;
    }
    text = capture.buffer.str();
  });
  copyLog(text, log, logCapacity);
  return status;
}

// The real BuildSceneGeometry, collectively on `nranks` thread-ranks, from
// ALREADY-CONVEX levels (convexify is un-vendored AMReX).
int ref_build_scene(int nranks, int nlevels, const ref_level* levels,
                    int minLevel, int maxLevel, int component, int logScale,
                    int normalizeToDataRange, ref_scene** out) {
  return guarded([&] {
    Capture quiet;
    auto scene = std::make_unique<ref_scene>();
    scene->nranks = nranks;
    scene->perRank.resize(nranks);
    fake_mpi_run(nranks, [&](int rank) {
      auto owned = std::make_shared<amrex::Vector<amrex::MultiFab>>();
      amrex::Vector<amrVolumeRenderer::detail::LevelGridGeometry> grids;
      for (int l = 0; l < nlevels; ++l) {
        const ref_level& level = levels[l];
        std::vector<amrex::Box> boxList;
        std::vector<int> ownerList;
        for (int b = 0; b < level.nboxes; ++b) {
          boxList.emplace_back(
              amrex::IntVect(level.lo[3 * b], level.lo[3 * b + 1],
                             level.lo[3 * b + 2]),
              amrex::IntVect(level.hi[3 * b], level.hi[3 * b + 1],
                             level.hi[3 * b + 2]));
          ownerList.push_back(level.owner[b]);
        }
        amrex::MultiFab mf(boxList, ownerList, level.ncomp, rank);
        for (int b = 0; b < level.nboxes; ++b) {
          if (level.owner[b] != rank) continue;
          std::memcpy(mf.dataPtr(b), level.data[b],
                      sizeof(double) * static_cast<std::size_t>(
                                           mf.box(b).numPts()) *
                          level.ncomp);
        }
        owned->push_back(std::move(mf));
        amrVolumeRenderer::detail::LevelGridGeometry grid;
        for (int d = 0; d < 3; ++d) {
          grid.probLo[d] = level.prob_lo[d];
          grid.cellSize[d] = level.cell_size[d];
        }
        grids.push_back(grid);
      }
      scene->perRank[rank] = amrVolumeRenderer::detail::BuildSceneGeometry(
          owned, grids,
          amrVolumeRenderer::detail::SceneBuildOptions{
              minLevel, maxLevel, component, logScale != 0,
              normalizeToDataRange != 0, "no data", "invalid scalar range"});
    });
    for (int rank = 0; rank < nranks; ++rank) {
      for (const AmrBox& box : scene->perRank[rank].localBoxes) {
        amrvr_box flat;
        fromAmrBox(box, &flat);
        scene->boxes.push_back(flat);
        scene->owners.push_back(rank);
      }
    }
    *out = scene.release();
  });
}

void ref_scene_free(ref_scene* scene) { delete scene; }

int ref_scene_info(const ref_scene* scene, int* nboxes,
                   const amrvr_box** boxes, const int32_t** owners,
                   amrvr_scalar_transform* transform, float* rangeMin,
                   float* rangeMax, double* boundsMin, double* boundsMax,
                   float* processedRange, float* originalRange) {
  const VolumeRenderer::SceneGeometry& g = scene->perRank[0];
  *nboxes = static_cast<int>(scene->boxes.size());
  *boxes = scene->boxes.data();
  *owners = scene->owners.data();
  fromTransform(g.scalarTransform, transform);
  *rangeMin = g.scalarRange.first;
  *rangeMax = g.scalarRange.second;
  for (int d = 0; d < 3; ++d) {
    boundsMin[d] = g.explicitBounds.minCorner[d];
    boundsMax[d] = g.explicitBounds.maxCorner[d];
  }
  if (processedRange) {
    processedRange[0] = g.processedScalarRange.first;
    processedRange[1] = g.processedScalarRange.second;
  }
  if (originalRange) {
    originalRange[0] = g.originalScalarRange.first;
    originalRange[1] = g.originalScalarRange.second;
  }
  return 0;
}

// The real ComputeSceneHistogram (SceneBuilder.cpp:445-577), collectively on the
// scene's thread-ranks; counts as rank 0 sees them (all ranks hold the same sums).
int ref_scene_histogram(const ref_scene* scene, int binCount, uint64_t* counts,
                        uint64_t* sampleCount) {
  return guarded([&] {
    Capture quiet;
    std::vector<VolumeRenderer::ScalarHistogram> results(scene->nranks);
    fake_mpi_run(scene->nranks, [&](int rank) {
      results[rank] = amrVolumeRenderer::detail::ComputeSceneHistogram(
          scene->perRank[rank], binCount);
    });
    const VolumeRenderer::ScalarHistogram& h = results[0];
    for (int b = 0; b < binCount; ++b)
      counts[b] = h.binCounts[static_cast<std::size_t>(b)];
    if (sampleCount) *sampleCount = h.sampleCount;
  });
}

// The real api::Render for ONE level (config 1, Examples/RenderFromMultiFab.cpp)
// on `nranks` thread-ranks.  Optional arguments follow api::RenderOptions.
int ref_api_render(int nranks, const ref_level* level,
                   const amrvr_frame_params* params, int component,
                   int logScale, const float* scalarRange /*[2] or NULL*/,
                   const double* upVector /*[3] or NULL*/,
                   const amrvr_camera* camera /*or NULL*/,
                   const amrvr_colormap_pt* pts, int npts, const char* outPath,
                   long long* samples, char* log, int logCapacity) {
  return guarded([&] {
    Capture capture;
    std::vector<long long> reads(nranks, 0);
    fake_mpi_run(nranks, [&](int rank) {
      std::vector<amrex::Box> boxList;
      std::vector<int> ownerList;
      for (int b = 0; b < level->nboxes; ++b) {
        boxList.emplace_back(
            amrex::IntVect(level->lo[3 * b], level->lo[3 * b + 1],
                           level->lo[3 * b + 2]),
            amrex::IntVect(level->hi[3 * b], level->hi[3 * b + 1],
                           level->hi[3 * b + 2]));
        ownerList.push_back(level->owner[b]);
      }
      amrex::MultiFab mf(boxList, ownerList, level->ncomp, rank);
      for (int b = 0; b < level->nboxes; ++b) {
        if (level->owner[b] != rank) continue;
        std::memcpy(mf.dataPtr(b), level->data[b],
                    sizeof(double) *
                        static_cast<std::size_t>(mf.box(b).numPts()) *
                        level->ncomp);
      }
      amrVolumeRenderer::api::AmrData data;
      data.levelData.push_back(&mf);
      data.levelGeometry.push_back(amrex::Geometry(
          {level->prob_lo[0], level->prob_lo[1], level->prob_lo[2]},
          {level->cell_size[0], level->cell_size[1], level->cell_size[2]}));
      amrVolumeRenderer::api::RenderOptions options;
      options.width = params->width;
      options.height = params->height;
      options.boxTransparency = params->box_transparency;
      options.antialiasing = params->antialiasing;
      options.visibilityGraph = params->use_visibility_graph != 0;
      options.component = component;
      options.logScaleInput = logScale != 0;
      options.outputFilename = outPath;
      if (scalarRange)
        options.scalarRange = std::make_pair(scalarRange[0], scalarRange[1]);
      if (upVector)
        options.upVector =
            amrex::RealVect(upVector[0], upVector[1], upVector[2]);
      if (camera) options.camera = toCamera(*camera);
      options.colorMap = toColorMap(pts, npts);
      const long long before = amrex::shim::array4_reads;
      const int status = amrVolumeRenderer::api::Render(data, options);
      reads[rank] = amrex::shim::array4_reads - before;
      if (status != 0) throw std::runtime_error("api::Render returned non-zero");
    });
    if (samples) {
      *samples = 0;
      for (long long r : reads) *samples += r;
    }
    copyLog(capture.buffer.str(), log, logCapacity);
  });
}

// The real api::ComputeHistogram (VolumeRendererApi.cpp:397-415) for ONE level on
// `nranks` thread-ranks; counts and ranges as rank 0 sees them.
int ref_api_histogram(int nranks, const ref_level* level, int component, int logScale,
                      int binCount, uint64_t* counts, uint64_t* sampleCount,
                      float* ranges /*[6]: normalized, processed, original*/,
                      int* flags /*[2]: hasProcessedRange, hasOriginalRange*/) {
  return guarded([&] {
    Capture quiet;
    std::vector<VolumeRenderer::ScalarHistogram> results(nranks);
    fake_mpi_run(nranks, [&](int rank) {
      std::vector<amrex::Box> boxList;
      std::vector<int> ownerList;
      for (int b = 0; b < level->nboxes; ++b) {
        boxList.emplace_back(
            amrex::IntVect(level->lo[3 * b], level->lo[3 * b + 1], level->lo[3 * b + 2]),
            amrex::IntVect(level->hi[3 * b], level->hi[3 * b + 1], level->hi[3 * b + 2]));
        ownerList.push_back(level->owner[b]);
      }
      amrex::MultiFab mf(boxList, ownerList, level->ncomp, rank);
      for (int b = 0; b < level->nboxes; ++b) {
        if (level->owner[b] != rank) continue;
        std::memcpy(mf.dataPtr(b), level->data[b],
                    sizeof(double) * static_cast<std::size_t>(mf.box(b).numPts()) * level->ncomp);
      }
      amrVolumeRenderer::api::AmrData data;
      data.levelData.push_back(&mf);
      data.levelGeometry.push_back(amrex::Geometry(
          {level->prob_lo[0], level->prob_lo[1], level->prob_lo[2]},
          {level->cell_size[0], level->cell_size[1], level->cell_size[2]}));
      amrVolumeRenderer::api::HistogramOptions options;
      options.component = component;
      options.logScaleInput = logScale != 0;
      options.binCount = binCount;
      results[rank] = amrVolumeRenderer::api::ComputeHistogram(data, options);
    });
    const VolumeRenderer::ScalarHistogram& h = results[0];
    for (int b = 0; b < binCount && b < static_cast<int>(h.binCounts.size()); ++b)
      counts[b] = h.binCounts[static_cast<std::size_t>(b)];
    if (sampleCount) *sampleCount = h.sampleCount;
    if (ranges) {
      ranges[0] = h.normalizedRange.first;
      ranges[1] = h.normalizedRange.second;
      ranges[2] = h.processedRange.first;
      ranges[3] = h.processedRange.second;
      ranges[4] = h.originalRange.first;
      ranges[5] = h.originalRange.second;
    }
    if (flags) {
      flags[0] = h.hasProcessedRange ? 1 : 0;
      flags[1] = h.hasOriginalRange ? 1 : 0;
    }
  });
}

}  // extern "C"
