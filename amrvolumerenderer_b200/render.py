"""render() / compute_histogram(): the two calls of the reference's nanobind
module (python/amrVolumeRenderer/module.cpp:131-262, :304-356) with the same
keywords, defaults and error behaviour (ValueError = std::invalid_argument,
RuntimeError = std::runtime_error), over libamrvr_b200.so.

What runs where: plotfile parsing, convexify and the argument/colour-map
logic of VolumeRenderer::run (VolumeRenderer/VolumeRenderer.cpp:1341-1576) on
the host; every per-cell and per-pixel step on the GPU (scene statistics,
quantise, ray-march, composite, overlay, bytes).  There is no CPU render path:
without the CUDA library or a B200 these calls raise.

One process per GPU: pass context=Context(device, rank, nranks, nccl_id) when
running under torchrun; the default is a single-rank context on cuda:0.
"""
from __future__ import annotations

import ctypes as C
import math
import struct
import sys
import zlib
from pathlib import Path

import numpy as np

from . import ctypes_api as api
from .engine import Context
from .plotfile import PlotFile

_P = C.POINTER
DEFAULT_OUTPUT = "volume-renderer.ppm"  # RunOptions::outputFilename, VolumeRenderer.hpp:62
DEFAULT_CAMERA_SEED = 91021             # RenderParameters::cameraSeed, VolumeRenderer.hpp:40


# ------------------------------------------------------------------ image files
def write_ppm(path, rgb8: np.ndarray) -> None:
    """Binary PPM as Common/SavePPM.cpp writes it (rows already top first)."""
    h, w, _ = rgb8.shape
    with open(path, "wb") as f:
        f.write(f"P6\n{w} {h}\n255\n".encode())
        f.write(np.ascontiguousarray(rgb8).tobytes())


def write_png(path, rgb8: np.ndarray) -> None:
    """8-bit RGB PNG (the container Common/SavePNG.cpp asks libpng for)."""
    h, w, _ = rgb8.shape
    raw = np.zeros((h, 1 + 3 * w), np.uint8)  # filter byte 0 per scanline
    raw[:, 1:] = rgb8.reshape(h, 3 * w)

    def chunk(tag: bytes, body: bytes) -> bytes:
        return struct.pack(">I", len(body)) + tag + body + struct.pack(">I", zlib.crc32(tag + body))

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(raw.tobytes(), 6)))
        f.write(chunk(b"IEND", b""))


def save_image(path, rgb8: np.ndarray) -> None:
    """Extension rule of renderSingleTrial (VolumeRenderer.cpp:1316-1328)."""
    ext = Path(str(path)).suffix.lower()
    if ext not in ("", ".ppm", ".png"):
        print(f"Render: unrecognized image extension '{ext}', defaulting to PPM output.",
              file=sys.stderr)
    (write_png if ext == ".png" else write_ppm)(path, rgb8)


# ----------------------------------------------------------------- validation
def _finite(*values) -> bool:
    return all(math.isfinite(float(v)) for v in values)


def _f32(v) -> float:
    return float(np.float32(v))


def _validate_parameters(width, height, box_transparency, antialiasing):
    # VolumeRenderer::validateRenderParameters, VolumeRenderer.cpp:562-579
    if width <= 0 or height <= 0:
        raise ValueError("image dimensions must be positive")
    if box_transparency < 0.0 or box_transparency > 1.0:
        raise ValueError("box transparency must be between 0 and 1")
    if antialiasing <= 0:
        raise ValueError("antialiasing must be positive")
    root = int(round(math.sqrt(antialiasing)))
    if root * root != antialiasing:
        raise ValueError("antialiasing must be a perfect square (1, 4, 9, ...)")


def _validate_color_map(color_map):
    # VolumeRenderer.cpp:1382-1416
    if len(color_map) < 2:
        raise ValueError("color map must provide at least two control points")
    previous = -math.inf
    for point in color_map:
        if len(point) != 5:
            raise ValueError("color map entries are (value, red, green, blue, alpha)")
        value = _f32(point[0])
        if not math.isfinite(value):
            raise ValueError("color map control point values must be finite")
        if value <= previous:
            raise ValueError("color map control point values must be strictly increasing")
        previous = value
        for comp, name in zip(point[1:], ("red", "green", "blue", "alpha")):
            c = _f32(comp)
            if not math.isfinite(c) or c < 0.0 or c > 1.0:
                raise ValueError(f"color map {name} components must be finite and within [0, 1]")


def _explicit_camera(eye, look_at, up, fov_y, near, far) -> api.Camera:
    # module.cpp:203-231, then VolumeRenderer.cpp:1418-1462 and :1558-1561
    if eye is None or look_at is None:
        raise ValueError("camera_eye and camera_look_at must be provided when specifying a camera")
    up = np.array(up if up is not None else (0.0, 1.0, 0.0), np.float32).astype(np.float64)
    eye = np.array(eye, np.float32).astype(np.float64)
    look_at = np.array(look_at, np.float32).astype(np.float64)
    up_len = _f32(np.sqrt((up * up).sum()))
    if not (up_len > 0.0) or not math.isfinite(up_len):
        raise ValueError("camera_up must contain finite, non-zero components")
    up = up / up_len
    fov_y = _f32(45.0 if fov_y is None else fov_y)
    near = _f32(0.1 if near is None else near)
    far = _f32(1000.0 if far is None else far)
    if not (_finite(*eye) and _finite(*look_at) and _finite(*up)):
        raise ValueError("camera vectors must have finite components")
    forward = look_at - eye
    flen = _f32(np.sqrt((forward * forward).sum()))
    if not (flen > 0.0) or not math.isfinite(flen):
        raise ValueError("camera eye and look-at must be distinct")
    ulen = _f32(np.sqrt((up * up).sum()))
    if not (ulen > 0.0) or not math.isfinite(ulen):
        raise ValueError("camera up vector must be non-zero")
    cross = np.cross(forward, up)
    if not (_f32(np.sqrt((cross * cross).sum())) > 1e-6):
        raise ValueError("camera up vector must not be parallel to the view direction")
    if not math.isfinite(fov_y) or not (0.0 < fov_y < 180.0):
        raise ValueError("camera fov must be in (0, 180) degrees")
    if not math.isfinite(near) or not (near > 0.0):
        raise ValueError("camera near plane must be > 0")
    if not math.isfinite(far) or not (far > near):
        raise ValueError("camera far plane must exceed the near plane")
    n = float(np.sqrt((up * up).sum()))  # safeNormalize
    up = up / n if (n > 0.0 and math.isfinite(n)) else np.array([0.0, 0.0, -1.0])
    cam = api.Camera()
    for d in range(3):
        cam.eye[d], cam.look_at[d], cam.up[d] = float(eye[d]), float(look_at[d]), float(up[d])
    cam.fov_y_degrees, cam.near_plane, cam.far_plane = fov_y, near, far
    return cam


# -------------------------------------------------------------- scene loading
class LoadedScene:
    """What VolumeRenderer::loadPlotFileGeometry returns (a SceneGeometry), with
    the FAB doubles resident in HBM until release()."""

    def __init__(self, ctx, boxes, summary, device_blocks, finest_level, min_level, max_level,
                 component_name):
        self.ctx, self.boxes, self.summary = ctx, boxes, summary
        self._blocks = device_blocks
        self.finest_level, self.min_level, self.max_level = finest_level, min_level, max_level
        self.component_name = component_name

    def release(self):
        for ptr in self._blocks:
            self.ctx.check(self.ctx.lib.amrvr_device_free(self.ctx._h, ptr))
        self._blocks = []


def load_plotfile_geometry(ctx: Context, plotfile_path, variable="", min_level=0, max_level=-1,
                           log_scale=False, normalize_to_data_range=True, reader_threads=8,
                           distribution=0) -> LoadedScene:
    """VolumeRenderer::loadPlotFileGeometry (VolumeRenderer.cpp:588-714):
    PlotFileData -> get(level, name) -> convexify -> BuildSceneGeometry."""
    if not plotfile_path:
        raise ValueError("Plotfile path must not be empty.")
    lib = ctx.lib
    with PlotFile(plotfile_path) as pf:
        if pf.space_dim != 3:
            raise RuntimeError(f"Plotfile '{plotfile_path}' has space dimension {pf.space_dim}. "
                               "The volume renderer currently expects 3D data.")
        if not pf.var_names:
            raise RuntimeError("Plotfile contains no cell variables to render.")
        name = variable or pf.var_names[0]
        component = pf.var_index(name)
        if component < 0:
            raise RuntimeError(f"Variable '{name}' not found in plotfile '{plotfile_path}'.")
        finest = pf.finest_level
        lo = min(max(int(min_level), 0), finest)
        hi = finest if (max_level < 0 or max_level > finest) else int(max_level)
        if lo > hi:
            raise RuntimeError(f"Minimum AMR level {lo} exceeds available maximum level {hi}.")

        nlev = hi + 1
        levels = (api.Level * nlev)()
        ratios = (C.c_int32 * 3 * max(nlev - 1, 1))()
        global_boxes, keep, blocks = [], [], []
        for l in range(nlev):
            meta = pf.level(l)
            all_boxes = np.ascontiguousarray(pf.boxes(l))
            global_boxes.append(all_boxes)
            idx, off = pf.level_share(l, ctx.rank, ctx.nranks, distribution)
            total = int(off[-1])
            fabs = (api.Fab * max(len(idx), 1))()
            if total > 0:
                # disk -> pinned host -> HBM, once; statistics and quantise read it in place
                host = ctx._pinned((total,), np.float64)
                pf.read_level(l, component, host, ctx.rank, ctx.nranks, distribution, reader_threads)
                dev = C.c_void_p()
                ctx.check(lib.amrvr_device_alloc(ctx._h, total * 8, C.byref(dev)))
                blocks.append(dev)
                ctx.check(lib.amrvr_copy_to_device(ctx._h, dev, host.ctypes.data_as(C.c_void_p),
                                                   total * 8))
                for k, b in enumerate(idx):
                    blo, bhi = all_boxes[b, :3], all_boxes[b, 3:]
                    nx, ny, nz = (int(bhi[d] - blo[d] + 1) for d in range(3))
                    f = fabs[k]
                    for d in range(3):
                        f.lo[d], f.hi[d], f.begin[d] = int(blo[d]), int(bhi[d]), int(blo[d])
                    f.data = dev.value + int(off[k]) * 8
                    f.jstride, f.kstride, f.nstride = nx, nx * ny, nx * ny * nz
            keep.append(fabs)
            for d in range(3):
                levels[l].prob_lo[d] = pf.prob_lo[d]
                levels[l].cell_size[d] = meta.cell_size[d]
                if l < nlev - 1:
                    ratios[l][d] = meta.ref_ratio[d]
            levels[l].nfabs = len(idx)
            levels[l].fabs = fabs

    # amrex::convexify (VolumeRenderer.cpp:669-670): views, no data motion
    counts = (C.c_int32 * nlev)()
    ptrs = (_P(C.c_int32) * nlev)(*[g.ctypes.data_as(_P(C.c_int32)) for g in global_boxes])
    nglobal = (C.c_int32 * nlev)(*[len(g) for g in global_boxes])
    capacity = 64
    while True:
        pieces = (api.Fab * capacity)()
        status = lib.amrvr_convexify_levels(nlev, levels, ptrs, nglobal, ratios, pieces, capacity,
                                            counts)
        if status == 0:
            break
        if capacity > (1 << 24):
            raise api.AmrvrError(status, lib.amrvr_last_error(None).decode())
        capacity *= 4
    convex = (api.Level * nlev)()
    base = 0
    for l in range(nlev):
        convex[l].prob_lo, convex[l].cell_size = levels[l].prob_lo, levels[l].cell_size
        convex[l].nfabs = counts[l]
        convex[l].fabs = C.cast(C.byref(pieces, base * C.sizeof(api.Fab)), _P(api.Fab))
        base += counts[l]

    try:
        boxes, summary = ctx.scene_build(convex, nlev, lo, hi, 0, log_scale, normalize_to_data_range,
                                         capacity=max(base, 1))
    except api.AmrvrError as e:
        for ptr in blocks:
            lib.amrvr_device_free(ctx._h, ptr)
        text = {"no data": "Failed to locate any volumetric data within the plotfile.",
                "invalid scalar range": "Failed to compute a valid scalar range from the plotfile."}
        raise RuntimeError(text.get(e.message, e.message)) from e
    scene = LoadedScene(ctx, boxes, summary, blocks, finest, lo, hi, name)
    scene._keep = (keep, pieces, global_boxes)
    if ctx.rank == 0:
        msg = (f"Loaded plotfile '{plotfile_path}' with variable '{name}' across "
               f"{hi - lo + 1} level(s)")
        if lo > 0 or hi < finest:
            msg += f" (levels {lo}-{hi})"
        if normalize_to_data_range:
            msg += "; normalized scalar range [0, 1]"
        else:
            msg += (f"; scalar range [{summary.processed_range[0]:g}, "
                    f"{summary.processed_range[1]:g}]")
        if log_scale:
            msg += " (log scaled)"
        print(msg)
    return scene


def resolve_scalar_options(summary, log_scale, scalar_range, color_map):
    """The scalar part of VolumeRenderer::run after the scene is loaded
    (VolumeRenderer.cpp:1478-1555): processed range check, scalar-range override
    (SetSceneNormalizationRange, SceneBuilder.cpp:427-443) and the colour map
    converted from physical to normalised scalar values.  Pure host arithmetic in
    the reference's float/double types.  Returns (transform, (range_min,
    range_max), colour-map points or None); `summary.transform` is updated in
    place when the override applies."""
    has_override = scalar_range is not None
    pmin, pmax = _f32(summary.processed_range[0]), _f32(summary.processed_range[1])
    span = _f32(np.float32(pmax) - np.float32(pmin))
    if not (span > 0.0) or not math.isfinite(span):
        raise RuntimeError("Failed to establish a finite scalar range for color mapping.")

    def to_processed(value: float) -> np.float32:  # VolumeRenderer.cpp:1490-1504
        value = np.float32(value)
        if not np.isfinite(value):
            raise ValueError("color_map scalar values must be finite.")
        if log_scale:
            if not (value > 0.0):
                raise ValueError("color_map scalar values must be positive when log scaling "
                                 "is enabled.")
            return np.float32(math.log(float(value)))
        return value

    norm_min, norm_max = np.float32(pmin), np.float32(pmax)
    transform = summary.transform
    value_range = (float(summary.scalar_range[0]), float(summary.scalar_range[1]))
    if has_override:
        norm_min, norm_max = to_processed(scalar_range[0]), to_processed(scalar_range[1])
        if not (norm_min < norm_max):
            raise ValueError("scalar_range must contain two values with min < max.")
    norm_span = np.float32(norm_max - norm_min)
    if not (norm_span > 0.0) or not np.isfinite(norm_span):
        raise RuntimeError("Failed to establish a finite scalar range for color mapping.")
    if has_override:  # SetSceneNormalizationRange, SceneBuilder.cpp:427-443
        dmin, dmax = float(norm_min), float(norm_max)
        transform.normalize_to_unit_range = 1
        transform.normalization_min = dmin
        transform.normalization_max = dmax
        transform.inverse_normalization_span = 1.0 / (dmax - dmin)
        value_range = (0.0, 1.0)

    points = None
    if color_map is not None:  # VolumeRenderer.cpp:1543-1555
        points = []
        for p in color_map:
            t = np.float32((to_processed(p[0]) - norm_min) / norm_span)
            if not np.isfinite(t):
                raise ValueError("color_map produced a non-finite normalized scalar value.")
            points.append((float(min(max(t, np.float32(0.0)), np.float32(1.0))),
                           *[float(c) for c in p[1:]]))
    return transform, value_range, points


# ------------------------------------------------------------------ runtime
# initialize_runtime / finalize_runtime (module.cpp:104-119): in the reference
# they keep MPI and AMReX up across several render() calls.  Here the thing
# worth keeping is the GPU context (CUDA context, streams, pinned staging,
# brick pools), so a held runtime means render() and compute_histogram() reuse
# one Context instead of creating and destroying one per call.
_runtime = {"refs": 0, "context": None}


def initialize_runtime() -> None:
    _runtime["refs"] += 1


def finalize_runtime() -> None:
    if _runtime["refs"] == 0:
        raise RuntimeError("amrVolumeRenderer.finalize_runtime requires a matching "
                           "initialize_runtime call")
    _runtime["refs"] -= 1
    if _runtime["refs"] == 0 and _runtime["context"] is not None:
        _runtime["context"].close()
        _runtime["context"] = None


def _acquire_context(context):
    """(context to use, whether the caller must close it)."""
    if context is not None:
        return context, False
    if _runtime["refs"] > 0:
        if _runtime["context"] is None:
            _runtime["context"] = Context.from_env()
        return _runtime["context"], False
    return Context.from_env(), True


# ------------------------------------------------------------------- render()
def render(plotfile, width=512, height=512, box_transparency=0.0, antialiasing=1,
           visibility_graph=True, write_visibility_graph=False, variable=None, min_level=0,
           max_level=-1, log_scale=False, up_vector=None, output=None, scalar_range=None,
           camera_eye=None, camera_look_at=None, camera_up=None, camera_fov_y=None,
           camera_near=None, camera_far=None, color_map=None, *, context: Context | None = None,
           return_frame: bool = False):
    """Render a plotfile with the DirectSend compositor; returns 0 like the
    reference (or the frame dict when return_frame=True, an addition for tests:
    'rgb8', 'final5', 'stats', 'camera', 'scene')."""
    if output is not None and output == "":
        raise ValueError("output filename must not be empty")
    output = DEFAULT_OUTPUT if output is None else output
    _validate_parameters(width, height, box_transparency, antialiasing)
    if not plotfile:
        raise RuntimeError("plotfile path is required")
    if min_level < 0:
        raise ValueError("min level must be non-negative")
    if max_level < -1:
        raise ValueError("max level must be non-negative or -1 for all levels")
    if max_level >= 0 and min_level > max_level:
        raise RuntimeError("min level must not exceed max level")

    custom_up = None
    if up_vector is not None:
        v = np.array(up_vector, np.float32).astype(np.float64)
        length = _f32(np.sqrt((v * v).sum()))
        if not (length > 0.0) or not math.isfinite(length):
            raise ValueError("up_vector must contain finite, non-zero components")
        custom_up = v / length
    if scalar_range is not None:
        rmin, rmax = _f32(scalar_range[0]), _f32(scalar_range[1])
        if not _finite(rmin, rmax) or not (rmin < rmax):
            raise ValueError("scalar_range must contain two finite values with min < max")
    if color_map is not None:
        _validate_color_map(color_map)
    camera = None
    if any(v is not None for v in (camera_eye, camera_look_at, camera_up, camera_fov_y,
                                   camera_near, camera_far)):
        camera = _explicit_camera(camera_eye, camera_look_at, camera_up, camera_fov_y,
                                  camera_near, camera_far)
    if not Path(str(plotfile)).exists():
        raise RuntimeError(f"plotfile path '{plotfile}' does not exist")

    ctx, own_ctx = _acquire_context(context)
    scene = None
    dscene = None
    try:
        has_override = scalar_range is not None
        scene = load_plotfile_geometry(ctx, str(plotfile), variable or "", min_level, max_level,
                                       log_scale, normalize_to_data_range=not has_override)
        summary = scene.summary
        transform, value_range, points = resolve_scalar_options(summary, log_scale, scalar_range,
                                                                color_map)

        if camera is None:  # renderScene without a camera, VolumeRenderer.cpp:975-1030
            camera = api.Camera()
            up = (C.c_double * 3)(*custom_up) if custom_up is not None else None
            status = ctx.lib.amrvr_auto_camera(summary.bounds_min, summary.bounds_max,
                                               DEFAULT_CAMERA_SEED, up, C.byref(camera))
            if status != 0:
                raise api.AmrvrError(status, ctx.lib.amrvr_last_error(None).decode())
        dscene = ctx.upload(scene.boxes, summary.nboxes, transform, value_range)
        if write_visibility_graph and visibility_graph and ctx.rank == 0:
            # exportGraph on rank 0 (VisibilityOrdering.cpp:319-352); the aspect is the
            # render size's (VolumeRenderer.cpp:1116-1119, 1221-1224)
            from .engine import write_visibility_graphs

            records, nglobal = dscene.global_boxes()
            write_visibility_graphs(ctx.nranks, records, nglobal, camera,
                                    float(np.float32(width) / np.float32(height)))
        scene.release()  # the doubles are no longer needed: bricks are bytes now
        params = api.FrameParams(width=int(width), height=int(height),
                                 box_transparency=float(box_transparency),
                                 antialiasing=int(antialiasing),
                                 use_visibility_graph=int(bool(visibility_graph)),
                                 count_samples=0, draw_bounding_box=1)
        frame = ctx.render_frame(dscene, camera, params, colormap=points, root=0,
                                 want_rgba=return_frame, want_rgb8=True)
        if ctx.rank == 0:
            # the reference's three stage lines and pixel count (VolumeRenderer.cpp:1121-1136,
            # 1223, 1245, 1253, 1296-1299); times are this rank's CUDA-event stage times
            stats = frame["stats"]
            for label, ms in (("per-box rendering", stats.march_ms),
                              ("visibility graph computation", stats.order_ms),
                              ("compositing", stats.composite_ms)):
                print(f"Render: {label} took {ms * 1e-3:.4f} s")
            root_aa = max(int(round(math.sqrt(int(antialiasing)))), 1)
            print(f"Render: composed {int(width) * root_aa * int(height) * root_aa} pixels on rank 0")
            save_image(output, frame["rgb8"])
            print(f"Saved volume composited image to '{output}'")
        if return_frame:
            out = dict(frame)
            if out["rgb8"] is not None:
                out["rgb8"] = out["rgb8"].copy()
            if out.get("final5") is not None:
                out["final5"] = out["final5"].copy()
            out["camera"], out["transform"], out["scalar_range"] = camera, transform, value_range
            out["summary"], out["color_map"] = summary, points
            out["boxes"] = [_box_tuple(scene.boxes[i]) for i in range(summary.nboxes)]
            return out
        return 0
    finally:
        if dscene is not None:
            dscene.free()
        if scene is not None:
            scene.release()
        if own_ctx:
            ctx.close()


def _box_tuple(b):
    return (tuple(b.min_corner), tuple(b.max_corner), tuple(b.dims), tuple(b.lo))


def compute_histogram(plotfile, variable=None, min_level=0, max_level=-1, log_scale=False,
                      bins=256, *, context: Context | None = None) -> dict:
    """module.cpp:304-356 over VolumeRenderer::computeScalarHistogram
    (VolumeRenderer.cpp:880-897)."""
    if bins <= 0:
        raise ValueError("binCount must be positive")
    ctx, own_ctx = _acquire_context(context)
    scene = None
    try:
        scene = load_plotfile_geometry(ctx, str(plotfile), variable or "", min_level, max_level,
                                       log_scale)
        s = scene.summary
        counts, total = ctx.scene_histogram(scene.boxes, s.nboxes, s.transform,
                                            (s.scalar_range[0], s.scalar_range[1]), bins)
        return {
            "counts": [int(c) for c in counts],
            "normalized_range": (float(s.scalar_range[0]), float(s.scalar_range[1])),
            "processed_range": (float(s.processed_range[0]), float(s.processed_range[1])),
            "original_range": ((float(s.original_range[0]), float(s.original_range[1]))
                               if s.has_original_range else None),
            "samples": int(total),
        }
    finally:
        if scene is not None:
            scene.release()
        if own_ctx:
            ctx.close()
