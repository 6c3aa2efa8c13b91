"""ctypes view of include/amrvr.h: POD structs, prototypes and the loader of
``libamrvr_b200.so``.

The product path has NO CPU fallback: :func:`load_library` raises
:class:`LibraryMissingError` when the CUDA extension has not been built, and
every compute entry point of the library itself fails with ``AMRVR_ERR_CUDA``
when no sm_100 device is usable.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

AMRVR_OK = 0
AMRVR_ERR_INVALID = 1
AMRVR_ERR_RUNTIME = 2
AMRVR_ERR_CUDA = 3
AMRVR_ERR_NCCL = 4
COLOR_TABLE_SIZE = 256
NCCL_ID_BYTES = 128
BRICK_STATS_STATIC_MAX_BOXES = 4096  # AMRVR_BRICK_STATS_STATIC_MAX_BOXES

PKG_DIR = Path(__file__).resolve().parent
REPO_ROOT = PKG_DIR.parent
LIB_PATH = PKG_DIR / "csrc" / "libamrvr_b200.so"


class Box(C.Structure):
    """amrvr_box == volume::AmrBox + Array4 (Common/VolumeTypes.hpp:69-76)."""

    _fields_ = [
        ("min_corner", C.c_double * 3),
        ("max_corner", C.c_double * 3),
        ("dims", C.c_int32 * 3),
        ("lo", C.c_int32 * 3),
        ("data", C.c_void_p),
        ("jstride", C.c_int64),
        ("kstride", C.c_int64),
        ("nstride", C.c_int64),
        ("begin", C.c_int32 * 3),
        ("component", C.c_int32),
    ]


class ScalarTransform(C.Structure):
    """amrvr_scalar_transform == volume::ScalarTransform (VolumeTypes.hpp:21-31)."""

    _fields_ = [
        ("log_scale_input", C.c_int32),
        ("normalize_to_unit_range", C.c_int32),
        ("positive_floor", C.c_double),
        ("processed_min", C.c_double),
        ("processed_max", C.c_double),
        ("inverse_processed_span", C.c_double),
        ("normalization_min", C.c_double),
        ("normalization_max", C.c_double),
        ("inverse_normalization_span", C.c_double),
    ]


class Camera(C.Structure):
    """amrvr_camera == volume::CameraParameters (VolumeTypes.hpp:83-90)."""

    _fields_ = [
        ("eye", C.c_double * 3),
        ("look_at", C.c_double * 3),
        ("up", C.c_double * 3),
        ("fov_y_degrees", C.c_float),
        ("near_plane", C.c_float),
        ("far_plane", C.c_float),
    ]


class ColorMapPoint(C.Structure):
    """amrvr_colormap_pt == volume::ColorMapControlPoint (VolumeTypes.hpp:92-98)."""

    _fields_ = [
        ("value", C.c_float),
        ("red", C.c_float),
        ("green", C.c_float),
        ("blue", C.c_float),
        ("alpha", C.c_float),
    ]


class BoxRecord(C.Structure):
    _fields_ = [
        ("min_corner", C.c_double * 3),
        ("max_corner", C.c_double * 3),
        ("dims", C.c_int32 * 3),
        ("owner", C.c_int32),
        ("local_index", C.c_int32),
        ("reserved", C.c_int32),
    ]


class MarchStats(C.Structure):
    _fields_ = [
        ("samples", C.c_int64),
        ("rays", C.c_int64),
        ("fragment_pixels", C.c_int64),
        ("kernel_ms", C.c_float),
    ]


class FrameParams(C.Structure):
    _fields_ = [
        ("width", C.c_int32),
        ("height", C.c_int32),
        ("box_transparency", C.c_float),
        ("antialiasing", C.c_int32),
        ("use_visibility_graph", C.c_int32),
        ("count_samples", C.c_int32),
        ("draw_bounding_box", C.c_int32),
        ("rgb8_on_device", C.c_int32),
    ]


class FrameStats(C.Structure):
    _fields_ = [
        ("march_ms", C.c_float),
        ("order_ms", C.c_float),
        ("composite_ms", C.c_float),
        ("gather_ms", C.c_float),
        ("post_ms", C.c_float),
        ("fold_ms", C.c_float),
        ("samples", C.c_int64),
        ("exchanged_bytes", C.c_int64),
        ("stolen_tiles", C.c_int64),
        ("tiles", C.c_int64),
        ("fragment_pixels", C.c_int64),
        ("march_span_ns", C.c_int64),
        ("march_cta_busy_ns", C.c_int64),
        ("march_table_ns", C.c_int64),
        ("march_ctas", C.c_int64),
        ("pre_ms", C.c_float),
        ("pad_", C.c_float),
    ]


class Fab(C.Structure):
    """amrvr_fab: one FAB as MFIter hands it out (valid box + Array4 view)."""

    _fields_ = [
        ("lo", C.c_int32 * 3),
        ("hi", C.c_int32 * 3),
        ("data", C.c_void_p),
        ("jstride", C.c_int64),
        ("kstride", C.c_int64),
        ("nstride", C.c_int64),
        ("begin", C.c_int32 * 3),
        ("reserved", C.c_int32),
    ]


class Level(C.Structure):
    """amrvr_level: one AMR level of api::AmrData (this rank's FABs + geometry)."""

    _fields_ = [
        ("prob_lo", C.c_double * 3),
        ("cell_size", C.c_double * 3),
        ("nfabs", C.c_int32),
        ("reserved", C.c_int32),
        ("fabs", C.POINTER(Fab)),
    ]


class SceneSummary(C.Structure):
    """amrvr_scene_summary: SceneGeometry minus the boxes (VolumeRenderer.hpp:46-62)."""

    _fields_ = [
        ("transform", ScalarTransform),
        ("scalar_range", C.c_float * 2),
        ("processed_range", C.c_float * 2),
        ("original_range", C.c_float * 2),
        ("has_original_range", C.c_int32),
        ("nboxes", C.c_int32),
        ("bounds_min", C.c_double * 3),
        ("bounds_max", C.c_double * 3),
        ("global_scale", C.c_double),
    ]


class PlotfileInfo(C.Structure):
    """amrvr_plotfile_info: what PlotFileData reports about the whole file."""

    _fields_ = [
        ("space_dim", C.c_int32),
        ("ncomp", C.c_int32),
        ("finest_level", C.c_int32),
        ("reserved", C.c_int32),
        ("time", C.c_double),
        ("prob_lo", C.c_double * 3),
        ("prob_hi", C.c_double * 3),
    ]


class PlotfileLevel(C.Structure):
    """amrvr_plotfile_level: per-level geometry + BoxArray size."""

    _fields_ = [
        ("cell_size", C.c_double * 3),
        ("ref_ratio", C.c_int32 * 3),
        ("domain_lo", C.c_int32 * 3),
        ("domain_hi", C.c_int32 * 3),
        ("nboxes", C.c_int32),
        ("real_bytes", C.c_int32),
    ]


class LibraryMissingError(RuntimeError):
    """libamrvr_b200.so has not been built (run __graft_entry__.build())."""


class AmrvrError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"amrvr status {status}: {message}")
        self.status = status
        self.message = message


def colormap_array(points):
    """[(value, r, g, b, a), ...] or None -> (ctypes array or None, count)."""
    if points is None or len(points) == 0:
        return None, 0
    arr = (ColorMapPoint * len(points))()
    for i, p in enumerate(points):
        arr[i] = ColorMapPoint(*[float(x) for x in p])
    return arr, len(points)


_P = C.POINTER
_lib = None


def _declare(lib):
    vp, i32, i64, f32 = C.c_void_p, C.c_int, C.c_int64, C.c_float
    protos = {
        "amrvr_abi_version": (C.c_int, []),
        "amrvr_device_count": (C.c_int, [_P(i32)]),
        "amrvr_nccl_unique_id": (C.c_int, [vp]),
        "amrvr_ctx_create": (C.c_int, [i32, i32, i32, vp, _P(vp)]),
        "amrvr_ctx_destroy": (None, [vp]),
        "amrvr_last_error": (C.c_char_p, [vp]),
        "amrvr_ctx_stream": (vp, [vp]),
        "amrvr_ctx_kernel_launches": (i64, [vp]),
        "amrvr_ctx_set_order_device_min_boxes": (C.c_int, [vp, i32]),
        "amrvr_ctx_frame_rgb8_device": (vp, [vp]),
        "amrvr_ctx_last_rank_order": (C.c_int, [vp, _P(i32), _P(i32)]),
        "amrvr_scene_is_replicated": (C.c_int, [vp]),
        "amrvr_scene_upload_stats": (C.c_int, [vp, _P(f32), _P(i64), _P(i64)]),
        "amrvr_peer_bandwidth": (C.c_int, [vp, C.c_size_t, i32, i32, i32, _P(f32)]),
        "amrvr_brick_stats": (C.c_int, [vp, i32, _P(Box), _P(C.c_double), _P(C.c_double),
                                        _P(C.c_double), _P(i64)]),
        "amrvr_brick_stats_static": (C.c_int, [i32, vp, i32, _P(Box), _P(C.c_double), _P(C.c_double),
                                               _P(C.c_double), _P(i64)]),
        "amrvr_ctx_last_composite_ms": (C.c_int, [vp, _P(f32)]),
        "amrvr_scene_upload": (C.c_int, [vp, i32, _P(Box), _P(ScalarTransform), f32, f32, _P(vp)]),
        "amrvr_stripe_plan_host": (C.c_int, [i32, i32, _P(C.c_int32), _P(C.c_int32), _P(C.c_int32),
                                             _P(C.c_uint64), _P(C.c_int32), _P(C.c_uint64)]),
        "amrvr_stripe_layer_row_host": (C.c_int64, [i32, i32, i32, i32]),
        "amrvr_scene_upload_async": (C.c_int, [vp, i32, _P(Box), _P(ScalarTransform), f32, f32, _P(vp)]),
        "amrvr_scene_wait": (C.c_int, [vp, vp]),
        "amrvr_scene_upload_staged": (C.c_int, [vp, i32, _P(Box), _P(ScalarTransform), f32, f32, _P(vp)]),
        "amrvr_scene_upload_was_staged": (C.c_int, [vp]),
        "amrvr_scene_free": (C.c_int, [vp, vp]),
        "amrvr_scene_build": (C.c_int, [vp, i32, _P(Level), i32, i32, i32, i32, i32, _P(Box), i32,
                                        _P(SceneSummary)]),
        "amrvr_scene_histogram": (C.c_int, [vp, i32, _P(Box), _P(ScalarTransform), f32, f32, i32,
                                            _P(C.c_uint64), _P(C.c_uint64)]),
        "amrvr_scene_global_boxes": (C.c_int, [vp, _P(i32), _P(_P(BoxRecord))]),
        "amrvr_scene_read_brick": (C.c_int, [vp, vp, i32, vp]),
        "amrvr_plotfile_open": (C.c_int, [C.c_char_p, _P(vp)]),
        "amrvr_plotfile_close": (None, [vp]),
        "amrvr_plotfile_get_info": (C.c_int, [vp, _P(PlotfileInfo)]),
        "amrvr_plotfile_var_name": (C.c_char_p, [vp, i32]),
        "amrvr_plotfile_var_index": (C.c_int, [vp, C.c_char_p]),
        "amrvr_plotfile_get_level": (C.c_int, [vp, i32, _P(PlotfileLevel)]),
        "amrvr_plotfile_level_boxes": (C.c_int, [vp, i32, _P(C.c_int32), i32]),
        "amrvr_plotfile_read_fab": (C.c_int, [vp, i32, i32, i32, vp, i64]),
        "amrvr_plotfile_read_level": (C.c_int, [vp, i32, i32, i32, i32, i32, i32, vp, i64,
                                                _P(C.c_int32), _P(i64), i32, _P(C.c_int32)]),
        "amrvr_convexify_boxes": (C.c_int, [i32, _P(C.c_int32), i32, _P(C.c_int32), _P(C.c_int32),
                                            _P(C.c_int32), _P(C.c_int32), i32, _P(C.c_int32)]),
        "amrvr_convexify_levels": (C.c_int, [i32, _P(Level), _P(_P(C.c_int32)), _P(C.c_int32),
                                             _P(C.c_int32 * 3), _P(Fab), i32, _P(C.c_int32)]),
        "amrvr_auto_camera": (C.c_int, [_P(C.c_double), _P(C.c_double), C.c_uint, _P(C.c_double),
                                        _P(Camera)]),
        "amrvr_order_boxes": (C.c_int, [vp, vp, _P(Camera), f32, i32, _P(i32), _P(i32)]),
        "amrvr_order_records": (C.c_int, [vp, i32, i32, _P(BoxRecord), _P(Camera), f32, i32,
                                          _P(i32), _P(i32)]),
        "amrvr_composite_virtual": (C.c_int, [vp, vp, vp, _P(Camera), i32, _P(i32), _P(i32), vp,
                                              _P(i32)]),
        "amrvr_order_boxes_host": (C.c_int, [i32, i32, _P(BoxRecord), _P(Camera), f32, i32,
                                             _P(i32), _P(i32)]),
        "amrvr_order_graph_dot_host": (C.c_int, [i32, i32, _P(BoxRecord), _P(Camera), f32, i32,
                                                 C.c_char_p, C.c_int64, _P(C.c_int64), _P(i32)]),
        "amrvr_layer_order_host": (C.c_int, [i32, _P(BoxRecord), _P(Camera), _P(f32), _P(i32)]),
        "amrvr_piece_range": (C.c_int, [i32, i32, i32, _P(i32), _P(i32)]),
        "amrvr_reference_sample_distance": (C.c_int, [i32, _P(BoxRecord), _P(C.c_double),
                                                      _P(C.c_double), _P(f32)]),
        "amrvr_color_table": (C.c_int, [f32, f32, f32, f32, _P(ColorMapPoint), i32, _P(f32)]),
        "amrvr_raymarch": (C.c_int, [vp, vp, _P(Camera), i32, i32, f32, f32, _P(ColorMapPoint),
                                     i32, i32, _P(vp)]),
        "amrvr_layers_free": (C.c_int, [vp, vp]),
        "amrvr_layers_count": (C.c_int, [vp]),
        "amrvr_layers_stats": (C.c_int, [vp, _P(MarchStats)]),
        "amrvr_layers_rect": (C.c_int, [vp, i32, _P(i32), _P(i32), _P(i32), _P(i32)]),
        "amrvr_layers_read": (C.c_int, [vp, vp, i32, vp]),
        "amrvr_composite": (C.c_int, [vp, vp, vp, _P(Camera), _P(i32), _P(i32), _P(i32), _P(vp)]),
        "amrvr_exchange_plan_host": (C.c_int, [i32, i32, i32, _P(BoxRecord), _P(Camera), i32, i32,
                                               _P(i32), _P(i64), _P(i64), _P(i32), _P(i32)]),
        "amrvr_composite_dense": (C.c_int, [vp, i32, i32, i32, _P(_P(f32)), _P(f32), _P(i32), _P(i32),
                                            _P(i32), vp]),
        "amrvr_gather": (C.c_int, [vp, vp, i32, i32, i32, i32, vp]),
        "amrvr_downsample": (C.c_int, [vp, vp, i32, i32, i32, vp]),
        "amrvr_quantize_rgb8": (C.c_int, [vp, vp, i32, i32, vp]),
        "amrvr_device_alloc": (C.c_int, [vp, C.c_size_t, _P(vp)]),
        "amrvr_device_free": (C.c_int, [vp, vp]),
        "amrvr_copy_to_host": (C.c_int, [vp, vp, vp, C.c_size_t]),
        "amrvr_copy_to_device": (C.c_int, [vp, vp, vp, C.c_size_t]),
        "amrvr_host_alloc_pinned": (C.c_int, [C.c_size_t, _P(vp)]),
        "amrvr_host_free_pinned": (C.c_int, [vp]),
        "amrvr_synchronize": (C.c_int, [vp]),
        "amrvr_render_frame": (C.c_int, [vp, vp, _P(Camera), _P(FrameParams), _P(ColorMapPoint),
                                         i32, i32, vp, vp, _P(FrameStats)]),
    }
    for name, (res, args) in protos.items():
        fn = getattr(lib, name)  # AttributeError here == header/library mismatch
        fn.restype = res
        fn.argtypes = args
    return protos


def declared_symbols():
    """Names include/amrvr.h declares (used by the CPU symbol-export test)."""
    import re

    text = (REPO_ROOT / "include" / "amrvr.h").read_text()
    return sorted(set(re.findall(r"\b(amrvr_[a-z0-9_]+)\s*\(", text)))


def load_library(path: os.PathLike | None = None):
    """Load libamrvr_b200.so (once).  Never falls back to anything else."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    # AMRVR_LIB: another build of the same library (tools/march_sweep.py compares variants)
    p = Path(path) if path is not None else Path(os.environ.get("AMRVR_LIB") or LIB_PATH)
    if not p.exists():
        raise LibraryMissingError(
            f"{p} not found: build the CUDA extension first "
            "(python -c 'import __graft_entry__ as g; g.build()'). "
            "There is no CPU fallback for the render path.")
    lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
    _declare(lib)
    if path is None:
        _lib = lib
    return lib


def as_f64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float64)
