// VolumeRendererApiB200.cpp -- defines api::Render and api::ComputeHistogram with
// the reference's signatures (VolumeRenderer/VolumeRendererApi.hpp:22-56) on top
// of libamrvr_b200.so.  Compile it INSTEAD of VolumeRenderer/VolumeRendererApi.cpp.
//
// This is the seam where simulation data enters (the in-situ path,
// Examples/RenderFromMultiFab.cpp:20-58).  The reference's body is
//   validate -> amrex::convexify (copies into new MultiFabs) -> BuildSceneGeometry
//   -> renderScene: one dense W*H*5-float host layer per box, host blends, gather.
// Here the MultiFabs are only LOOKED at: every FAB is handed to the C ABI as a
// view (valid box + Array4 strides), amrvr_convexify_levels narrows the views to
// the cells no finer level covers, amrvr_scene_build reduces the statistics on
// the GPU, amrvr_scene_upload turns the views into byte bricks (the strided
// quantise is the masked copy) and amrvr_render_frame produces the bytes of the
// image -- no dense layer ever exists.  Argument checks, their exception types
// and texts, the status lines on stdout and the image file follow
// VolumeRendererApi.cpp:257-415 and VolumeRenderer.cpp:947-1101, 1296-1336.
#include <VolumeRenderer/VolumeRendererApi.hpp>

#include <Common/CameraUtils.hpp>

#include <algorithm>
#include <cctype>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iomanip>
#include <iostream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>

#include "B200Runtime.hpp"

namespace amrVolumeRenderer {
namespace api {
namespace {

using Vec3 = amrex::RealVect;

// ---- argument checks: VolumeRendererApi.cpp:149-255 (same conditions, same texts)

void validateScalarRange(const std::optional<std::pair<float, float>>& range) {
  if (!range) return;
  if (!std::isfinite(range->first) || !std::isfinite(range->second) || !(range->first < range->second))
    throw std::invalid_argument("scalar_range must contain two finite values with min < max");
}

void validateColorMap(const std::optional<VolumeRenderer::ColorMap>& colorMap) {
  if (!colorMap) return;
  if (colorMap->size() < 2)
    throw std::invalid_argument("color map must provide at least two control points");
  float previous = -std::numeric_limits<float>::infinity();
  for (const auto& point : *colorMap) {
    if (!std::isfinite(point.value))
      throw std::invalid_argument("color map control point values must be finite");
    if (point.value <= previous)
      throw std::invalid_argument("color map control point values must be strictly increasing");
    previous = point.value;
    const std::pair<float, const char*> parts[4] = {
        {point.red, "red"}, {point.green, "green"}, {point.blue, "blue"}, {point.alpha, "alpha"}};
    for (const auto& part : parts)
      if (!std::isfinite(part.first) || part.first < 0.0f || part.first > 1.0f)
        throw std::invalid_argument(std::string("color map ") + part.second +
                                    " components must be finite and within [0, 1]");
  }
}

void validateCamera(const std::optional<VolumeRenderer::CameraParameters>& camera) {
  if (!camera) return;
  const auto finite3 = [](const Vec3& v) {
    return std::isfinite(v[0]) && std::isfinite(v[1]) && std::isfinite(v[2]);
  };
  const Vec3 eye(camera->eye), lookAt(camera->lookAt), up(camera->up);
  if (!finite3(eye) || !finite3(lookAt) || !finite3(up))
    throw std::invalid_argument("camera vectors must have finite components");
  const Vec3 forward = lookAt - eye;
  const float forwardLength = static_cast<float>(forward.vectorLength());
  if (!(forwardLength > 0.0f) || !std::isfinite(forwardLength))
    throw std::invalid_argument("camera eye and look-at must be distinct");
  const float upLength = static_cast<float>(up.vectorLength());
  if (!(upLength > 0.0f) || !std::isfinite(upLength))
    throw std::invalid_argument("camera up vector must be non-zero");
  const float cross = static_cast<float>(forward.crossProduct(up).vectorLength());
  if (!(cross > 1e-6f))
    throw std::invalid_argument("camera up vector must not be parallel to the view direction");
  if (!std::isfinite(camera->fovYDegrees) || !(camera->fovYDegrees > 0.0f) || !(camera->fovYDegrees < 180.0f))
    throw std::invalid_argument("camera fov must be in (0, 180) degrees");
  if (!std::isfinite(camera->nearPlane) || !(camera->nearPlane > 0.0f))
    throw std::invalid_argument("camera near plane must be > 0");
  if (!std::isfinite(camera->farPlane) || !(camera->farPlane > camera->nearPlane))
    throw std::invalid_argument("camera far plane must exceed the near plane");
}

void validateUpVector(const std::optional<amrex::RealVect>& upVector) {
  if (!upVector) return;
  const float length = static_cast<float>(upVector->vectorLength());
  if (!(length > 0.0f) || !std::isfinite(length))
    throw std::invalid_argument("up_vector must be non-zero and finite");
}

// ---- the scene: views of the caller's MultiFabs -> convex pieces -> boxes + summary

struct LoadedScene {
  amrvr_ctx* ctx = nullptr;
  std::vector<std::vector<amrvr_fab>> fabs;  // per level: this rank's FABs as views
  std::vector<amrvr_fab> pieces;             // convexified views, level after level
  std::vector<amrvr_box> boxes;              // BuildSceneGeometry's localBoxes
  amrvr_scene_summary summary{};
  int rank = 0;
};

// loadMultiFabGeometry, VolumeRendererApi.cpp:44-148
LoadedScene loadMultiFabScene(const AmrData& data, int requestedMinLevel, int requestedMaxLevel,
                              int component, bool logScaleInput, bool normalizeToDataRange) {
  if (AMREX_SPACEDIM != 3)
    throw std::runtime_error("The volume renderer currently expects 3D data (AMREX_SPACEDIM=3).");
  if (data.levelData.empty())
    throw std::invalid_argument("levelData must include at least one level");
  if (data.levelGeometry.size() != data.levelData.size())
    throw std::invalid_argument("levelGeometry must match the size of levelData");

  const int finestLevel = static_cast<int>(data.levelData.size()) - 1;
  int minLevel = std::min(std::max(requestedMinLevel, 0), finestLevel);
  int maxLevel = requestedMaxLevel;
  if (maxLevel < 0 || maxLevel > finestLevel) maxLevel = finestLevel;
  if (minLevel > maxLevel) throw std::runtime_error("minLevel must not exceed maxLevel");
  if (maxLevel > 0 && data.refinementRatios.size() < static_cast<std::size_t>(maxLevel))
    throw std::invalid_argument("refinementRatios must provide ratios for each level transition");
  for (int level = 0; level <= maxLevel; ++level) {
    if (data.levelData[level] == nullptr)
      throw std::invalid_argument("levelData contains a null MultiFab pointer");
    if (component < 0 || component >= data.levelData[level]->nComp())
      throw std::invalid_argument("component index is out of bounds");
  }

  LoadedScene scene;
  MPI_Comm_rank(MPI_COMM_WORLD, &scene.rank);
  scene.ctx = amrvr_b200::groupContext(MPI_COMM_WORLD);
  const int nlevels = maxLevel + 1;
  scene.fabs.resize(static_cast<std::size_t>(nlevels));
  std::vector<amrvr_level> levels(static_cast<std::size_t>(nlevels));
  std::vector<std::vector<int32_t>> globalBoxes(static_cast<std::size_t>(nlevels));
  std::vector<const int32_t*> globalPtrs(static_cast<std::size_t>(nlevels));
  std::vector<int32_t> globalCounts(static_cast<std::size_t>(nlevels));
  std::vector<int32_t> ratios(static_cast<std::size_t>(3 * std::max(nlevels - 1, 1)), 1);
  for (int level = 0; level < nlevels; ++level) {
    const amrex::MultiFab& mf = *data.levelData[level];
    // this rank's FABs: valid box + the Array4 view MFIter hands out
    for (amrex::MFIter mfi(mf); mfi.isValid(); ++mfi) {
      const amrex::Box& box = mfi.validbox();
      const auto view = mf.const_array(mfi);
      amrvr_fab fab{};
      for (int d = 0; d < 3; ++d) {
        fab.lo[d] = box.smallEnd()[d];
        fab.hi[d] = box.bigEnd()[d];
      }
      fab.begin[0] = view.begin.x;
      fab.begin[1] = view.begin.y;
      fab.begin[2] = view.begin.z;
      fab.data = view.p;
      fab.jstride = view.jstride;
      fab.kstride = view.kstride;
      fab.nstride = view.nstride;
      scene.fabs[level].push_back(fab);
    }
    // the level's BoxArray, the same on every rank: what convexify subtracts
    for (int b = 0; b < static_cast<int>(mf.size()); ++b) {
      const amrex::Box& box = mf.box(b);
      for (int d = 0; d < 3; ++d) globalBoxes[level].push_back(box.smallEnd()[d]);
      for (int d = 0; d < 3; ++d) globalBoxes[level].push_back(box.bigEnd()[d]);
    }
    globalPtrs[level] = globalBoxes[level].data();
    globalCounts[level] = static_cast<int32_t>(mf.size());
    const auto probLo = data.levelGeometry[level].ProbLoArray();
    const auto cellSize = data.levelGeometry[level].CellSizeArray();
    for (int d = 0; d < 3; ++d) {
      levels[level].prob_lo[d] = probLo[d];
      levels[level].cell_size[d] = cellSize[d];
      if (level < maxLevel) ratios[3 * level + d] = data.refinementRatios[level][d];
    }
    levels[level].nfabs = static_cast<int32_t>(scene.fabs[level].size());
    levels[level].fabs = scene.fabs[level].data();
  }

  // amrex::convexify (:97-98) without moving a cell
  std::vector<int32_t> counts(static_cast<std::size_t>(nlevels), 0);
  std::size_t capacity = 64;
  for (;;) {
    scene.pieces.assign(capacity, amrvr_fab{});
    const int status = amrvr_convexify_levels(
        nlevels, levels.data(), globalPtrs.data(), globalCounts.data(),
        reinterpret_cast<const int32_t(*)[3]>(ratios.data()), scene.pieces.data(),
        static_cast<int>(capacity), counts.data());
    if (status == AMRVR_OK) break;
    if (capacity > (std::size_t(1) << 24)) amrvr_b200::check(nullptr, status, "amrvr_convexify_levels");
    capacity *= 4;
  }
  std::vector<amrvr_level> convex(levels);
  std::size_t base = 0;
  for (int level = 0; level < nlevels; ++level) {
    convex[level].nfabs = counts[level];
    convex[level].fabs = scene.pieces.data() + base;
    base += static_cast<std::size_t>(counts[level]);
  }

  // BuildSceneGeometry (:113-124)
  scene.boxes.assign(std::max<std::size_t>(base, 1), amrvr_box{});
  const int status = amrvr_scene_build(scene.ctx, nlevels, convex.data(), minLevel, maxLevel, component,
                                       logScaleInput ? 1 : 0, normalizeToDataRange ? 1 : 0,
                                       scene.boxes.data(), static_cast<int>(scene.boxes.size()),
                                       &scene.summary);
  if (status != AMRVR_OK) {
    const std::string why = amrvr_last_error(scene.ctx);
    if (why == "no data")
      throw std::runtime_error("Failed to locate any volumetric data within the AMReX inputs.");
    if (why == "invalid scalar range")
      throw std::runtime_error("Failed to compute a valid scalar range from the AMReX inputs.");
    amrvr_b200::check(scene.ctx, status, "amrvr_scene_build");
  }

  if (scene.rank == 0) {  // :126-146
    std::cout << "Loaded AMReX data component " << component << " across " << (maxLevel - minLevel + 1)
              << " level(s)";
    if (minLevel > 0 || maxLevel < finestLevel) std::cout << " (levels " << minLevel << "-" << maxLevel << ")";
    if (normalizeToDataRange)
      std::cout << "; normalized scalar range [0, 1]";
    else
      std::cout << "; scalar range [" << scene.summary.processed_range[0] << ", "
                << scene.summary.processed_range[1] << "]";
    if (logScaleInput) std::cout << " (log scaled)";
    std::cout << std::endl;
  }
  return scene;
}

// ---- image files: SavePPM.cpp:17-36 byte for byte; PNG as a zlib stream of stored
// blocks (libpng is not needed to WRITE one: the pixel bytes are the path's product)

bool writePPM(const std::string& filename, int width, int height, const std::vector<uint8_t>& rgb8) {
  std::ofstream file(filename.c_str(), std::ios_base::binary | std::ios_base::out);
  file << "P6" << std::endl;
  file << width << " " << height << std::endl;
  file << 255 << std::endl;
  file.write(reinterpret_cast<const char*>(rgb8.data()), static_cast<std::streamsize>(rgb8.size()));
  file.close();
  return !file.fail();
}

uint32_t crc32Of(const uint8_t* data, std::size_t n, uint32_t crc) {
  static uint32_t table[256];
  static bool ready = false;
  if (!ready) {
    for (uint32_t i = 0; i < 256; ++i) {
      uint32_t c = i;
      for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : (c >> 1);
      table[i] = c;
    }
    ready = true;
  }
  crc = ~crc;
  for (std::size_t i = 0; i < n; ++i) crc = table[(crc ^ data[i]) & 0xffu] ^ (crc >> 8);
  return ~crc;
}

bool writePNG(const std::string& filename, int width, int height, const std::vector<uint8_t>& rgb8) {
  std::vector<uint8_t> raw;  // filter byte 0 + RGB per row (rows are already top first)
  raw.reserve(static_cast<std::size_t>(height) * (1 + 3 * static_cast<std::size_t>(width)));
  for (int y = 0; y < height; ++y) {
    raw.push_back(0);
    const uint8_t* row = rgb8.data() + static_cast<std::size_t>(y) * width * 3;
    raw.insert(raw.end(), row, row + static_cast<std::size_t>(width) * 3);
  }
  std::vector<uint8_t> z = {0x78, 0x01};
  uint32_t a = 1, b = 0;
  for (uint8_t v : raw) {
    a = (a + v) % 65521u;
    b = (b + a) % 65521u;
  }
  for (std::size_t off = 0; off < raw.size() || off == 0; off += 65535) {
    const std::size_t len = std::min<std::size_t>(65535, raw.size() - off);
    z.push_back(off + len >= raw.size() ? 1 : 0);
    z.push_back(static_cast<uint8_t>(len & 0xff));
    z.push_back(static_cast<uint8_t>(len >> 8));
    z.push_back(static_cast<uint8_t>(~len & 0xff));
    z.push_back(static_cast<uint8_t>((~len >> 8) & 0xff));
    z.insert(z.end(), raw.begin() + static_cast<std::ptrdiff_t>(off),
             raw.begin() + static_cast<std::ptrdiff_t>(off + len));
    if (raw.empty()) break;
  }
  const uint32_t adler = (b << 16) | a;
  for (int s = 24; s >= 0; s -= 8) z.push_back(static_cast<uint8_t>(adler >> s));
  std::ofstream file(filename.c_str(), std::ios_base::binary | std::ios_base::out);
  const auto be32 = [&](uint32_t v) {
    const char bytes[4] = {static_cast<char>(v >> 24), static_cast<char>(v >> 16), static_cast<char>(v >> 8),
                           static_cast<char>(v)};
    file.write(bytes, 4);
  };
  const auto chunk = [&](const char* tag, const std::vector<uint8_t>& body) {
    be32(static_cast<uint32_t>(body.size()));
    file.write(tag, 4);
    file.write(reinterpret_cast<const char*>(body.data()), static_cast<std::streamsize>(body.size()));
    uint32_t crc = crc32Of(reinterpret_cast<const uint8_t*>(tag), 4, 0);
    crc = crc32Of(body.data(), body.size(), crc);
    be32(crc);
  };
  const unsigned char magic[8] = {0x89, 'P', 'N', 'G', 0x0d, 0x0a, 0x1a, 0x0a};
  file.write(reinterpret_cast<const char*>(magic), 8);
  std::vector<uint8_t> header(13, 0);
  for (int s = 0; s < 4; ++s) {
    header[s] = static_cast<uint8_t>(static_cast<uint32_t>(width) >> (24 - 8 * s));
    header[4 + s] = static_cast<uint8_t>(static_cast<uint32_t>(height) >> (24 - 8 * s));
  }
  header[8] = 8;  // bit depth
  header[9] = 2;  // colour type RGB
  chunk("IHDR", header);
  chunk("IDAT", z);
  chunk("IEND", {});
  file.close();
  return !file.fail();
}

std::string lowercaseExtension(const std::string& filename) {
  const std::size_t dot = filename.find_last_of('.');
  if (dot == std::string::npos) return {};
  std::string ext = filename.substr(dot);
  std::transform(ext.begin(), ext.end(), ext.begin(),
                 [](unsigned char c) { return static_cast<char>(std::tolower(c)); });
  return ext;
}

}  // namespace

int Render(const AmrData& data, const RenderOptions& options) {
  // VolumeRendererApi.cpp:257-276
  if (options.outputFilename.empty()) throw std::invalid_argument("output filename must not be empty");
  if (options.minLevel < 0) throw std::invalid_argument("min level must be non-negative");
  if (options.maxLevel < -1)
    throw std::invalid_argument("max level must be non-negative or -1 for all levels");
  if (options.maxLevel >= 0 && options.minLevel > options.maxLevel)
    throw std::runtime_error("min level must not exceed max level");
  validateUpVector(options.upVector);
  validateScalarRange(options.scalarRange);
  validateColorMap(options.colorMap);
  validateCamera(options.camera);

  const bool hasScalarOverride = options.scalarRange.has_value();
  LoadedScene scene = loadMultiFabScene(data, options.minLevel, options.maxLevel, options.component,
                                        options.logScaleInput, /*normalizeToDataRange=*/!hasScalarOverride);
  amrvr_scene_summary& summary = scene.summary;

  // :301-379 processed range, scalar-range override, colour map -> normalised values
  const float processedMin = summary.processed_range[0], processedMax = summary.processed_range[1];
  const float processedSpan = processedMax - processedMin;
  if (!(processedSpan > 0.0f) || !std::isfinite(processedSpan))
    throw std::runtime_error("Failed to establish a finite scalar range for color mapping.");
  const auto toProcessed = [&](float physicalValue) -> float {
    if (!std::isfinite(physicalValue))
      throw std::invalid_argument("color_map scalar values must be finite.");
    if (options.logScaleInput) {
      if (!(physicalValue > 0.0f))
        throw std::invalid_argument(
            "color_map scalar values must be positive when log scaling is enabled.");
      return std::log(physicalValue);
    }
    return physicalValue;
  };
  float normalizationMin = processedMin, normalizationMax = processedMax;
  if (hasScalarOverride) {
    normalizationMin = toProcessed(options.scalarRange->first);
    normalizationMax = toProcessed(options.scalarRange->second);
    if (!(normalizationMin < normalizationMax))
      throw std::invalid_argument("scalar_range must contain two values with min < max.");
  }
  const float normalizationSpan = normalizationMax - normalizationMin;
  if (!(normalizationSpan > 0.0f) || !std::isfinite(normalizationSpan))
    throw std::runtime_error("Failed to establish a finite scalar range for color mapping.");
  float rangeMin = summary.scalar_range[0], rangeMax = summary.scalar_range[1];
  if (hasScalarOverride) {  // SetSceneNormalizationRange, SceneBuilder.cpp:427-443
    const double lo = normalizationMin, hi = normalizationMax;
    summary.transform.normalize_to_unit_range = 1;
    summary.transform.normalization_min = lo;
    summary.transform.normalization_max = hi;
    summary.transform.inverse_normalization_span = 1.0 / (hi - lo);
    rangeMin = 0.0f;
    rangeMax = 1.0f;
  }
  std::vector<amrvr_colormap_pt> points;
  if (options.colorMap) {
    for (const auto& controlPoint : *options.colorMap) {
      const float normalized = (toProcessed(controlPoint.value) - normalizationMin) / normalizationSpan;
      if (!std::isfinite(normalized))
        throw std::invalid_argument("color_map produced a non-finite normalized scalar value.");
      points.push_back(amrvr_colormap_pt{std::clamp(normalized, 0.0f, 1.0f), controlPoint.red,
                                         controlPoint.green, controlPoint.blue, controlPoint.alpha});
    }
  }

  // renderScene's own entry checks and banner (VolumeRenderer.cpp:562-586, 952-953)
  if (options.width <= 0 || options.height <= 0)
    throw std::invalid_argument("image dimensions must be positive");
  if (options.boxTransparency < 0.0f || options.boxTransparency > 1.0f)
    throw std::invalid_argument("box transparency must be between 0 and 1");
  if (options.antialiasing <= 0) throw std::invalid_argument("antialiasing must be positive");
  {
    const int root = static_cast<int>(std::lround(std::sqrt(options.antialiasing)));
    if (root * root != options.antialiasing)
      throw std::invalid_argument("antialiasing must be a perfect square (1, 4, 9, ...)");
  }
  if (scene.rank == 0) {
    int numProcs = 1;
    MPI_Comm_size(MPI_COMM_WORLD, &numProcs);
    std::cout << "volume_renderer: Using AMReX volume mapper on " << numProcs << " ranks" << std::endl;
  }

  // the camera: the caller's (up re-normalised, :381-384) or renderScene's automatic one
  amrvr_camera camera{};
  if (options.camera) {
    VolumeRenderer::CameraParameters given = *options.camera;
    given.up = camera::safeNormalize(given.up);
    camera = amrvr_b200::toAbi(given);
  } else {
    double up[3] = {0.0, 1.0, 0.0};
    if (options.upVector)
      for (int d = 0; d < 3; ++d) up[d] = (*options.upVector)[d];
    amrvr_b200::check(nullptr,
                      amrvr_auto_camera(summary.bounds_min, summary.bounds_max, 91021u,
                                        options.upVector ? up : nullptr, &camera),
                      "amrvr_auto_camera");
  }

  amrvr_ctx* ctx = scene.ctx;
  amrvr_scene* uploaded = nullptr;
  // The statistics pass of amrvr_scene_build left these FABs in the staging buffer: host MultiFabs
  // cross PCIe once, not twice.  The staged copy is recognised by pointers, extents and strides only,
  // so NOTHING may touch the MultiFabs between buildScene (amrvr_scene_build) above and this call --
  // no FillBoundary, no reallocation at the same address: the doubles quantised here would be the
  // stale staged ones.  AMRVR_DEBUG_STAGED=1 prints which path the upload took.
  amrvr_b200::check(ctx,
                    amrvr_scene_upload_staged(ctx, summary.nboxes, scene.boxes.data(), &summary.transform,
                                              rangeMin, rangeMax, &uploaded),
                    "amrvr_scene_upload_staged");
  if (const char* debugStaged = std::getenv("AMRVR_DEBUG_STAGED"); debugStaged != nullptr && debugStaged[0] == '1')
    std::fprintf(stderr, "[amrvr] api::Render rank %d: bricks quantised from %s\n", scene.rank,
                 amrvr_scene_upload_was_staged(uploaded) ? "the doubles amrvr_scene_build staged (no second PCIe pass)"
                                                         : "a fresh copy of the FABs");
  amrvr_frame_params params{};
  params.width = options.width;
  params.height = options.height;
  params.box_transparency = options.boxTransparency;
  params.antialiasing = options.antialiasing;
  params.use_visibility_graph = options.visibilityGraph ? 1 : 0;
  params.count_samples = 0;
  params.draw_bounding_box = 1;
  if (options.writeVisibilityGraph && options.visibilityGraph && scene.rank == 0) {
    // exportGraph on rank 0 (VisibilityOrdering.cpp:319-352); file counter as the reference's static
    static int graphCounter = 0;
    int nglobal = 0, nranks = 1;
    const amrvr_box_record* records = nullptr;
    MPI_Comm_size(MPI_COMM_WORLD, &nranks);
    amrvr_scene_global_boxes(uploaded, &nglobal, &records);
    const float aspect = static_cast<float>(options.width) / static_cast<float>(options.height);
    for (int round = 0;; ++round) {
      int64_t needed = 0;
      int rounds = 0;
      if (amrvr_order_graph_dot_host(nranks, nglobal, records, &camera, aspect, round, nullptr, 0, &needed,
                                     &rounds) != AMRVR_OK || round >= rounds)
        break;
      std::string text(static_cast<std::size_t>(needed), '\0');
      amrvr_order_graph_dot_host(nranks, nglobal, records, &camera, aspect, round, text.data(), needed,
                                 nullptr, nullptr);
      const std::string name = "visibility_graph_" + std::to_string(graphCounter++) + ".dot";
      std::ofstream dot(name, std::ios_base::binary);
      dot.write(text.data(), static_cast<std::streamsize>(text.size()));
      if (dot.good()) std::cout << "Wrote visibility graph to '" << name << "'" << std::endl;
      else std::cerr << "Failed to write visibility graph to '" << name << "'" << std::endl;
    }
  }
  std::vector<uint8_t> rgb8;
  if (scene.rank == 0)
    rgb8.resize(static_cast<std::size_t>(std::max(options.width, 0)) * std::max(options.height, 0) * 3);
  amrvr_frame_stats stats{};
  const int status = amrvr_render_frame(ctx, uploaded, &camera, &params, points.empty() ? nullptr : points.data(),
                                        static_cast<int>(points.size()), 0, nullptr,
                                        scene.rank == 0 ? rgb8.data() : nullptr, &stats);
  const std::string why = (status == AMRVR_OK) ? std::string() : std::string(amrvr_last_error(ctx));
  amrvr_scene_free(ctx, uploaded);
  if (status != AMRVR_OK) {
    if (status == AMRVR_ERR_INVALID) throw std::invalid_argument("amrvr_render_frame: " + why);
    throw std::runtime_error("amrvr_render_frame: " + why);
  }

  if (scene.rank == 0) {
    // the reference's stage lines (VolumeRenderer.cpp:1121-1136, 1223, 1245, 1253) with this
    // rank's CUDA-event times, the pixel count (:1296-1299) and the image file (:1316-1336)
    const std::pair<const char*, float> stages[3] = {{"per-box rendering", stats.march_ms},
                                                     {"visibility graph computation", stats.order_ms},
                                                     {"compositing", stats.composite_ms}};
    for (const auto& stage : stages) {
      std::ostringstream line;
      line << std::fixed << std::setprecision(4);
      line << "Render: " << stage.first << " took " << static_cast<double>(stage.second) * 1e-3 << " s";
      std::cout << line.str() << std::endl;
    }
    const int sqrtAA = std::max(static_cast<int>(std::lround(std::sqrt(static_cast<double>(options.antialiasing)))), 1);
    std::cout << "Render: composed " << options.width * sqrtAA * options.height * sqrtAA
              << " pixels on rank 0" << std::endl;
    const std::string extension = lowercaseExtension(options.outputFilename);
    if (!extension.empty() && extension != ".ppm" && extension != ".png")
      std::cerr << "Render: unrecognized image extension '" << extension << "', defaulting to PPM output."
                << std::endl;
    const bool saved = (extension == ".png") ? writePNG(options.outputFilename, options.width, options.height, rgb8)
                                             : writePPM(options.outputFilename, options.width, options.height, rgb8);
    if (saved)
      std::cout << "Saved volume composited image to '" << options.outputFilename << "'" << std::endl;
    else
      std::cerr << "Failed to save composited image to '" << options.outputFilename << "'" << std::endl;
  }
  return 0;
}

VolumeRenderer::ScalarHistogram ComputeHistogram(const AmrData& data, const HistogramOptions& options) {
  if (options.binCount <= 0) throw std::invalid_argument("binCount must be positive");  // :399-401
  LoadedScene scene = loadMultiFabScene(data, options.minLevel, options.maxLevel, options.component,
                                        options.logScaleInput, /*normalizeToDataRange=*/true);
  const amrvr_scene_summary& summary = scene.summary;
  // ComputeSceneHistogram, SceneBuilder.cpp:445-577
  VolumeRenderer::ScalarHistogram histogram;
  histogram.binCounts.assign(static_cast<std::size_t>(options.binCount), 0);
  histogram.normalizedRange = {summary.scalar_range[0], summary.scalar_range[1]};
  histogram.processedRange = {summary.processed_range[0], summary.processed_range[1]};
  histogram.hasProcessedRange = true;
  histogram.originalRange = {summary.original_range[0], summary.original_range[1]};
  histogram.hasOriginalRange = summary.has_original_range != 0;
  uint64_t total = 0;
  amrvr_b200::check(scene.ctx,
                    amrvr_scene_histogram(scene.ctx, summary.nboxes, scene.boxes.data(), &summary.transform,
                                          summary.scalar_range[0], summary.scalar_range[1], options.binCount,
                                          histogram.binCounts.data(), &total),
                    "amrvr_scene_histogram");
  histogram.sampleCount = total;
  return histogram;
}

}  // namespace api
}  // namespace amrVolumeRenderer
