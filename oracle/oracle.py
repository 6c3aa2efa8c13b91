"""oracle/oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes wrappers over
  * oracle/liboracle.so          the plain-C restatement (oracle/restate/*.c)   -> class Oracle
  * oracle/_ref/libamrvr_ref.so  the reference's own sources built against the
                                 AMReX/MPI shim (oracle/Makefile)               -> class Ref
Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import
this module.  Nothing under amrvolumerenderer_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

from amrvolumerenderer_b200 import ctypes_api as api

HERE = Path(__file__).resolve().parent
RESTATE_LIB = HERE / "liboracle.so"
REF_LIB = HERE / "_ref" / "libamrvr_ref.so"
DROPIN_LIB = HERE / "_ref" / "libamrvr_dropin.so"
REFERENCE_ROOT = Path("/root/reference")

_P = C.POINTER
_f32p = _P(C.c_float)


def build(ref: bool | None = None) -> None:
    """Compile the restatement and, when the reference tree is present, _ref."""
    subprocess.run(["make", "-s", "-C", str(HERE), "restate"], check=True)
    if ref is None:
        ref = REFERENCE_ROOT.exists()
    if ref:
        subprocess.run(["make", "-s", "-j8", "-C", str(HERE), "ref"], check=True)
        # the reference's driver with its three hot-path files swapped for the
        # adapters over libamrvr_b200.so (needs that library built first)
        if (HERE.parent / "amrvolumerenderer_b200" / "csrc" / "libamrvr_b200.so").exists():
            subprocess.run(["make", "-s", "-j8", "-C", str(HERE), "dropin"], check=True)


def _fptr(a: np.ndarray):
    return a.ctypes.data_as(_f32p)


def _d3(a):
    return (C.c_double * 3)(*[float(x) for x in a])


class _Base:
    def _cmap(self, colormap):
        return api.colormap_array(colormap)


class Oracle(_Base):
    """Plain-C restatement."""

    def __init__(self):
        if not RESTATE_LIB.exists():
            build(ref=False)
        self.lib = C.CDLL(str(RESTATE_LIB))
        L = self.lib
        L.orc_depth_hint.restype = C.c_float
        L.orc_reference_sample_distance.restype = C.c_float
        L.orc_cell_index.restype = C.c_int
        L.orc_cell_index.argtypes = [C.c_double, _P(api.ScalarTransform), C.c_float, C.c_float]
        L.orc_order.restype = C.c_int

    def color_table(self, alpha_scale, norm, rmin, rmax, colormap=None) -> np.ndarray:
        out = np.zeros((256, 4), np.float32)
        pts, n = self._cmap(colormap)
        self.lib.orc_color_table(C.c_float(alpha_scale), C.c_float(norm), C.c_float(rmin),
                                 C.c_float(rmax), pts, n, _fptr(out))
        return out

    def quantize_box(self, cbox, transform, rmin, rmax) -> np.ndarray:
        n = cbox.dims[0] * cbox.dims[1] * cbox.dims[2]
        out = np.zeros(n, np.uint8)
        self.lib.orc_quantize_box(C.byref(cbox), C.byref(transform), C.c_float(rmin),
                                  C.c_float(rmax), out.ctypes.data_as(C.c_void_p))
        return out

    def cell_index(self, raw, transform, rmin, rmax) -> int:
        return self.lib.orc_cell_index(float(raw), C.byref(transform), rmin, rmax)

    def paint(self, cbox, transform, rmin, rmax, box_transparency, ref_dist, width, height,
              camera, colormap=None):
        out = np.zeros((height, width, 5), np.float32)
        samples = C.c_longlong(0)
        pts, n = self._cmap(colormap)
        self.lib.orc_paint(C.byref(cbox), C.byref(transform), C.c_float(rmin), C.c_float(rmax),
                           C.c_float(box_transparency), C.c_float(ref_dist), width, height,
                           C.byref(camera), pts, n, _fptr(out), C.byref(samples))
        return out, samples.value

    def depth_hint(self, record, camera) -> float:
        return float(self.lib.orc_depth_hint(C.byref(record), C.byref(camera)))

    def layer_order(self, records, n, camera):
        hints = np.zeros(max(n, 1), np.float32)
        order = np.zeros(max(n, 1), np.int32)
        self.lib.orc_layer_order(n, records, C.byref(camera), _fptr(hints),
                                 order.ctypes.data_as(_P(C.c_int)))
        return hints[:n], order[:n]

    def reference_sample_distance(self, records, n, bmin, bmax) -> float:
        return float(self.lib.orc_reference_sample_distance(n, records, _d3(bmin), _d3(bmax)))

    def order(self, nranks, records, n, camera, aspect, use_graph=True):
        out = np.zeros(nranks, np.int32)
        ok = self.lib.orc_order(nranks, n, records, C.byref(camera), C.c_float(aspect),
                                int(use_graph), out.ctypes.data_as(_P(C.c_int)))
        return out, bool(ok)

    def piece_range(self, npix, piece, npieces):
        b, e = C.c_int(0), C.c_int(0)
        self.lib.orc_piece_range(npix, piece, npieces, C.byref(b), C.byref(e))
        return b.value, e.value

    def blend(self, top, bottom) -> np.ndarray:
        t = np.ascontiguousarray(top, np.float32)
        b = np.ascontiguousarray(bottom, np.float32)
        out = np.zeros(5, np.float32)
        self.lib.orc_blend(_fptr(t), _fptr(b), _fptr(out))
        return out

    def compose(self, nranks, width, height, layers, owners, hints, rank_order):
        n = len(layers)
        layers = [np.ascontiguousarray(l, np.float32) for l in layers]
        ptrs = (_f32p * max(n, 1))(*[_fptr(l) for l in layers])
        own = np.ascontiguousarray(owners, np.int32)
        hin = np.ascontiguousarray(hints, np.float32)
        ro = np.ascontiguousarray(rank_order, np.int32)
        full = np.zeros((height, width, 5), np.float32)
        piece_owner = np.zeros(height * width, np.int32)
        self.lib.orc_compose(nranks, width, height, n, own.ctypes.data_as(_P(C.c_int)), ptrs,
                             _fptr(hin), ro.ctypes.data_as(_P(C.c_int)), _fptr(full),
                             piece_owner.ctypes.data_as(_P(C.c_int)))
        return full, piece_owner

    def downsample(self, src5, width, height, sqrt_aa):
        src = np.ascontiguousarray(src5, np.float32)
        out = np.zeros((height, width, 5), np.float32)
        self.lib.orc_downsample(_fptr(src), width, height, sqrt_aa, _fptr(out))
        return out

    def overlay(self, image5, tight_min, tight_max, camera, sqrt_aa=1):
        img = np.ascontiguousarray(image5, np.float32).copy()
        h, w = img.shape[:2]
        self.lib.orc_overlay(_d3(tight_min), _d3(tight_max), C.byref(camera), sqrt_aa, w, h,
                             _fptr(img))
        return img

    def quantize_rgb8(self, image5):
        img = np.ascontiguousarray(image5, np.float32)
        h, w = img.shape[:2]
        out = np.zeros((h, w, 3), np.uint8)
        self.lib.orc_quantize_rgb8(_fptr(img), w, h, out.ctypes.data_as(C.c_void_p))
        return out

    def auto_camera(self, bmin, bmax, azimuth, altitude, up=None) -> api.Camera:
        cam = api.Camera()
        self.lib.orc_auto_camera(_d3(bmin), _d3(bmax), C.c_float(azimuth), C.c_float(altitude),
                                 _d3(up) if up is not None else None, C.byref(cam))
        return cam

    def scene_build(self, levels, nlevels, min_level=0, max_level=None, component=0,
                    log_scale=False, normalize_to_data_range=True):
        """orc_scene_build over levels holding the FABs of all ranks ->
        (Box array, SceneSummary, status)."""
        if max_level is None:
            max_level = nlevels - 1
        cap = max(sum(int(levels[i].nfabs) for i in range(nlevels)), 1)
        boxes = (api.Box * cap)()
        summary = api.SceneSummary()
        self.lib.orc_scene_build.restype = C.c_int
        st = self.lib.orc_scene_build(nlevels, levels, int(min_level), int(max_level), int(component),
                                      int(bool(log_scale)), int(bool(normalize_to_data_range)), boxes,
                                      cap, C.byref(summary))
        return boxes, summary, st

    def scene_histogram(self, cboxes, n, transform, scalar_range, bin_count):
        counts = np.zeros(max(int(bin_count), 1), np.uint64)
        total = C.c_uint64(0)
        self.lib.orc_scene_histogram(n, cboxes, C.byref(transform), C.c_float(scalar_range[0]),
                                     C.c_float(scalar_range[1]), int(bin_count),
                                     counts.ctypes.data_as(_P(C.c_uint64)), C.byref(total))
        return counts, int(total.value)

    def render_frame(self, scene, camera, params, colormap=None):
        """scene: amrvolumerenderer_b200.scenes.Scene.  Returns dict."""
        cboxes, n, _keep = scene.c_boxes()
        owners = scene.owners()
        final = np.zeros((params.height, params.width, 5), np.float32)
        rgb = np.zeros((params.height, params.width, 3), np.uint8)
        order = np.zeros(scene.nranks, np.int32)
        samples = C.c_longlong(0)
        pts, npts = self._cmap(colormap)
        self.lib.orc_render_frame(
            scene.nranks, n, cboxes, owners.ctypes.data_as(_P(C.c_int32)),
            C.byref(scene.transform), C.c_float(scene.scalar_range[0]),
            C.c_float(scene.scalar_range[1]), _d3(scene.bounds_min), _d3(scene.bounds_max),
            C.byref(camera), C.byref(params), pts, npts, _fptr(final),
            rgb.ctypes.data_as(C.c_void_p), order.ctypes.data_as(_P(C.c_int)),
            C.byref(samples))
        return {"final5": final, "rgb8": rgb, "rank_order": order, "samples": samples.value}


class RefLevel(C.Structure):
    _fields_ = [
        ("prob_lo", C.c_double * 3),
        ("cell_size", C.c_double * 3),
        ("nboxes", C.c_int32),
        ("ncomp", C.c_int32),
        ("lo", _P(C.c_int32)),
        ("hi", _P(C.c_int32)),
        ("owner", _P(C.c_int32)),
        ("data", _P(C.c_void_p)),
    ]


class RefUnavailable(RuntimeError):
    pass


class Ref(_Base):
    """The reference's own code (oracle/_ref)."""

    LIB = REF_LIB

    def __init__(self):
        if not self.LIB.exists():
            if REFERENCE_ROOT.exists():
                build(ref=True)
            else:
                raise RefUnavailable(f"{self.LIB} not built and {REFERENCE_ROOT} is absent")
        if not self.LIB.exists():
            raise RefUnavailable(f"{self.LIB} could not be built")
        self.lib = C.CDLL(str(self.LIB))
        self.lib.ref_last_error.restype = C.c_char_p

    def _check(self, status):
        if status != 0:
            raise api.AmrvrError(status, self.lib.ref_last_error().decode())

    def paint(self, cbox, transform, rmin, rmax, box_transparency, ref_dist, width, height,
              camera, colormap=None):
        out = np.zeros((height, width, 5), np.float32)
        samples = C.c_longlong(0)
        pts, n = self._cmap(colormap)
        self._check(self.lib.ref_paint(
            C.byref(cbox), C.byref(transform), C.c_float(rmin), C.c_float(rmax),
            C.c_float(box_transparency), C.c_float(ref_dist), width, height, C.byref(camera),
            pts, n, _fptr(out), C.byref(samples)))
        return out, samples.value

    def order(self, nranks, records, n, camera, aspect, use_graph=True):
        out = np.zeros(nranks, np.int32)
        log = C.create_string_buffer(1 << 16)
        self._check(self.lib.ref_order(nranks, n, records, C.byref(camera), C.c_float(aspect),
                                       int(use_graph), out.ctypes.data_as(_P(C.c_int)), log,
                                       len(log)))
        return out, log.value.decode()

    def order_graphs(self, nranks, records, n, camera, aspect, workdir):
        """BuildVisibilityOrderedGroup with writeVisibilityGraph = true, run with
        `workdir` as the current directory: (rank order, [bytes of every
        visibility_graph_<n>.dot the call wrote, in counter order], log)."""
        import os
        import re

        workdir = os.fspath(workdir)
        before = set(os.listdir(workdir))
        out = np.zeros(nranks, np.int32)
        log = C.create_string_buffer(1 << 16)
        here = os.getcwd()
        os.chdir(workdir)
        try:
            self._check(self.lib.ref_order_write_graph(nranks, n, records, C.byref(camera),
                                                       C.c_float(aspect),
                                                       out.ctypes.data_as(_P(C.c_int)), log, len(log)))
        finally:
            os.chdir(here)
        new = [f for f in os.listdir(workdir)
               if f not in before and re.fullmatch(r"visibility_graph_\d+\.dot", f)]
        new.sort(key=lambda f: int(re.findall(r"\d+", f)[0]))
        docs = []
        for f in new:
            with open(os.path.join(workdir, f), "rb") as fh:
                docs.append(fh.read())
        return out, docs, log.value.decode()

    def compose(self, nranks, width, height, layers, owners, hints, rank_order):
        n = len(layers)
        layers = [np.ascontiguousarray(l, np.float32) for l in layers]
        ptrs = (_f32p * max(n, 1))(*[_fptr(l) for l in layers])
        own = np.ascontiguousarray(owners, np.int32)
        hin = np.ascontiguousarray(hints, np.float32)
        ro = np.ascontiguousarray(rank_order, np.int32)
        full = np.zeros((height, width, 5), np.float32)
        piece_owner = np.full(height * width, -1, np.int32)
        seconds = C.c_double(0)
        self._check(self.lib.ref_compose(
            nranks, width, height, n, own.ctypes.data_as(_P(C.c_int)), ptrs, _fptr(hin),
            ro.ctypes.data_as(_P(C.c_int)), _fptr(full),
            piece_owner.ctypes.data_as(_P(C.c_int)), C.byref(seconds)))
        return full, piece_owner, seconds.value

    def render_scene(self, scene, camera, params, out_path, colormap=None, pin_cores=False):
        """Real VolumeRenderer::renderScene -> PPM at out_path (rank 0)."""
        cboxes, n, _keep = scene.c_boxes()
        owners = scene.owners()
        samples = C.c_longlong(0)
        wall = C.c_double(0)
        log = C.create_string_buffer(1 << 16)
        pts, npts = self._cmap(colormap)
        self._check(self.lib.ref_render_scene(
            scene.nranks, n, cboxes, owners.ctypes.data_as(_P(C.c_int32)),
            C.byref(scene.transform), C.c_float(scene.scalar_range[0]),
            C.c_float(scene.scalar_range[1]), 1, _d3(scene.bounds_min), _d3(scene.bounds_max),
            C.byref(camera) if camera is not None else None, C.byref(params), pts, npts,
            str(out_path).encode(), int(pin_cores), C.byref(samples), C.byref(wall), log,
            len(log)))
        return {"samples": samples.value, "wall_s": wall.value, "log": log.value.decode()}

    def build_scene(self, scene, min_level=0, max_level=None, component=0, log_scale=False,
                    normalize_to_data_range=True, ncomp=1):
        """The real detail::BuildSceneGeometry on scene.nranks thread-ranks from the
        scene's levels -> dict (boxes, owners, transform, ranges, bounds, handle)."""
        nlevels = len(scene.levels)
        if max_level is None:
            max_level = nlevels - 1
        per = [[] for _ in range(nlevels)]
        for b in sorted(scene.boxes, key=lambda b: b.gen_index):
            per[b.level].append(b)
        levels = (RefLevel * max(nlevels, 1))()
        keep = []
        for li in range(nlevels):
            n = len(per[li])
            lo = (C.c_int32 * max(3 * n, 1))()
            hi = (C.c_int32 * max(3 * n, 1))()
            owner = (C.c_int32 * max(n, 1))()
            data = (C.c_void_p * max(n, 1))()
            for i, b in enumerate(per[li]):
                for d in range(3):
                    lo[3 * i + d] = int(b.lo[d])
                    hi[3 * i + d] = int(b.lo[d] + b.dims[d] - 1)
                owner[i] = int(b.owner)
                if ncomp == 1:
                    arr = b.data
                else:
                    arr = np.full((ncomp,) + b.data.shape, 1e30, np.float64)
                    arr[component] = b.data
                keep.append(arr)
                data[i] = arr.ctypes.data
            prob_lo, cs = scene.levels[li]
            for d in range(3):
                levels[li].prob_lo[d] = float(prob_lo[d])
                levels[li].cell_size[d] = float(cs[d])
            levels[li].nboxes, levels[li].ncomp = n, ncomp
            levels[li].lo, levels[li].hi, levels[li].owner = lo, hi, owner
            levels[li].data = data
            keep += [lo, hi, owner, data]
        handle = C.c_void_p()
        self._check(self.lib.ref_build_scene(scene.nranks, nlevels, levels, int(min_level),
                                             int(max_level), int(component), int(bool(log_scale)),
                                             int(bool(normalize_to_data_range)), C.byref(handle)))
        nb = C.c_int(0)
        boxes = _P(api.Box)()
        owners = _P(C.c_int32)()
        t = api.ScalarTransform()
        rmin, rmax = C.c_float(0), C.c_float(0)
        bmin, bmax = (C.c_double * 3)(), (C.c_double * 3)()
        proc, orig = (C.c_float * 2)(), (C.c_float * 2)()
        self.lib.ref_scene_info(handle, C.byref(nb), C.byref(boxes), C.byref(owners), C.byref(t),
                                C.byref(rmin), C.byref(rmax), bmin, bmax, proc, orig)
        return {"handle": handle, "keep": keep, "nboxes": nb.value, "boxes": boxes,
                "owners": [owners[i] for i in range(nb.value)], "transform": t,
                "scalar_range": (rmin.value, rmax.value), "bounds_min": list(bmin),
                "bounds_max": list(bmax), "processed_range": tuple(proc), "original_range": tuple(orig)}

    def _single_level(self, scene, ncomp=1, component=0):
        """ref_level of a one-level scene (boxes in generation order) + keepalive."""
        boxes = sorted(scene.boxes, key=lambda b: b.gen_index)
        n = len(boxes)
        level = RefLevel()
        lo = (C.c_int32 * max(3 * n, 1))()
        hi = (C.c_int32 * max(3 * n, 1))()
        owner = (C.c_int32 * max(n, 1))()
        data = (C.c_void_p * max(n, 1))()
        keep = [lo, hi, owner, data]
        for i, b in enumerate(boxes):
            for d in range(3):
                lo[3 * i + d] = int(b.lo[d])
                hi[3 * i + d] = int(b.lo[d] + b.dims[d] - 1)
            owner[i] = int(b.owner)
            if ncomp == 1:
                arr = b.data
            else:
                arr = np.full((ncomp,) + b.data.shape, 1e30, np.float64)
                arr[component] = b.data
            keep.append(arr)
            data[i] = arr.ctypes.data
        prob_lo, cs = scene.levels[0]
        for d in range(3):
            level.prob_lo[d] = float(prob_lo[d])
            level.cell_size[d] = float(cs[d])
        level.nboxes, level.ncomp = n, ncomp
        level.lo, level.hi, level.owner, level.data = lo, hi, owner, data
        return level, keep

    def api_render(self, scene, params, out_path, component=0, ncomp=1, log_scale=False,
                   scalar_range=None, up_vector=None, camera=None, colormap=None):
        """The real api::Render(AmrData, RenderOptions) (VolumeRendererApi.cpp:257-395) on a
        one-level scene, scene.nranks thread-ranks -> image at out_path (rank 0)."""
        level, _keep = self._single_level(scene, ncomp, component)
        samples = C.c_longlong(0)
        log = C.create_string_buffer(1 << 16)
        pts, npts = self._cmap(colormap)
        rng = (C.c_float * 2)(*[float(v) for v in scalar_range]) if scalar_range is not None else None
        up = (C.c_double * 3)(*[float(v) for v in up_vector]) if up_vector is not None else None
        self._check(self.lib.ref_api_render(
            scene.nranks, C.byref(level), C.byref(params), int(component), int(bool(log_scale)), rng, up,
            C.byref(camera) if camera is not None else None, pts, npts, str(out_path).encode(),
            C.byref(samples), log, len(log)))
        return {"samples": samples.value, "log": log.value.decode()}

    def api_histogram(self, scene, bin_count, component=0, ncomp=1, log_scale=False):
        """The real api::ComputeHistogram on a one-level scene -> dict."""
        level, _keep = self._single_level(scene, ncomp, component)
        counts = np.zeros(max(int(bin_count), 1), np.uint64)
        total = C.c_uint64(0)
        ranges = (C.c_float * 6)()
        flags = (C.c_int * 2)()
        self._check(self.lib.ref_api_histogram(scene.nranks, C.byref(level), int(component),
                                               int(bool(log_scale)), int(bin_count),
                                               counts.ctypes.data_as(_P(C.c_uint64)), C.byref(total),
                                               ranges, flags))
        return {"counts": counts, "samples": int(total.value), "normalized_range": tuple(ranges[0:2]),
                "processed_range": tuple(ranges[2:4]), "original_range": tuple(ranges[4:6]),
                "has_processed": bool(flags[0]), "has_original": bool(flags[1])}

    def register_plotfile(self, path, levels, names):
        """Hand the CONTENT of a plotfile (list of amrvolumerenderer_b200.plotfile.
        LevelSpec + variable names; prob_lo = 0) to the shim's PlotFileData under
        `path`, so the reference's own loadPlotFileGeometry / run() can read it."""
        ncomp = len(names)
        clev = (RefLevel * max(len(levels), 1))()
        keep = []
        for li, L in enumerate(levels):
            n = len(L.boxes)
            lo = (C.c_int32 * max(3 * n, 1))()
            hi = (C.c_int32 * max(3 * n, 1))()
            owner = (C.c_int32 * max(n, 1))()
            data = (C.c_void_p * max(n, 1))()
            for i, (blo, arr) in enumerate(L.boxes):
                arr = np.ascontiguousarray(arr if arr.ndim == 4 else arr[None], np.float64)
                nz, ny, nx = arr.shape[1:]
                for d, ext in enumerate((nx, ny, nz)):
                    lo[3 * i + d] = int(blo[d])
                    hi[3 * i + d] = int(blo[d] + ext - 1)
                keep.append(arr)
                data[i] = arr.ctypes.data
            for d in range(3):
                clev[li].prob_lo[d] = 0.0
                clev[li].cell_size[d] = float(L.cell_size[d])
            clev[li].nboxes, clev[li].ncomp = n, ncomp
            clev[li].lo, clev[li].hi, clev[li].owner, clev[li].data = lo, hi, owner, data
            keep += [lo, hi, owner, data]
        cnames = (C.c_char_p * ncomp)(*[n.encode() for n in names])
        self._check(self.lib.ref_register_plotfile(str(path).encode(), len(levels), clev, ncomp, cnames))

    def set_convexify_override(self, pieces_per_level):
        """Box cuts the shim's amrex::convexify returns (list of int32[n, 6] = lo, hi
        per level); None or [] restores the default (single level only)."""
        levels = [np.ascontiguousarray(p, np.int32).reshape(-1, 6) for p in (pieces_per_level or [])]
        counts = (C.c_int32 * max(len(levels), 1))(*[len(p) for p in levels])
        ptrs = (_P(C.c_int32) * max(len(levels), 1))(*[p.ctypes.data_as(_P(C.c_int32)) for p in levels])
        self._check(self.lib.ref_set_convexify_override(len(levels), counts, ptrs))

    def run(self, plotfile, output, params, nranks=1, variable=None, min_level=0, max_level=-1,
            log_scale=False, up_vector=None, scalar_range=None, camera=None, colormap=None):
        """The reference's own VolumeRenderer::run(RunOptions) over a registered
        plotfile -> image at `output`; returns its captured log.  Raises
        AmrvrError with the reference's message (status 1 = std::invalid_argument)."""
        log = C.create_string_buffer(1 << 16)
        pts, npts = self._cmap(colormap)
        sr = (C.c_float * 2)(*scalar_range) if scalar_range is not None else None
        self._check(self.lib.ref_run(
            int(nranks), str(plotfile).encode(), C.byref(params), str(output).encode(),
            variable.encode() if variable else None, int(min_level), int(max_level),
            int(bool(log_scale)), _d3(up_vector) if up_vector is not None else None, sr,
            C.byref(camera) if camera is not None else None, pts, npts, log, len(log)))
        return log.value.decode()

    def scene_histogram(self, built, bin_count):
        """The real detail::ComputeSceneHistogram on the scene build_scene returned."""
        counts = np.zeros(max(int(bin_count), 1), np.uint64)
        total = C.c_uint64(0)
        self._check(self.lib.ref_scene_histogram(built["handle"], int(bin_count),
                                                 counts.ctypes.data_as(_P(C.c_uint64)), C.byref(total)))
        return counts, int(total.value)

    def free_scene(self, built):
        self.lib.ref_scene_free(built["handle"])
        built["handle"] = None

    @staticmethod
    def stage_seconds(log: str) -> dict:
        """Parse the three 'Render: <stage> took X s' lines (VolumeRenderer.cpp:1121-1136)."""
        import re

        out = {}
        for label, key in (("per-box rendering", "march_s"),
                           ("visibility graph computation", "order_s"),
                           ("compositing", "composite_s")):
            m = re.search(rf"Render: {label} took ([0-9.eE+-]+) s", log)
            if m:
                out[key] = float(m.group(1))
        return out


class DropIn(Ref):
    """The reference's driver, scene builder and image classes with
    VolumePainter / BuildVisibilityOrderedGroup / DirectSendBase replaced by the
    adapters of amrvolumerenderer_b200/adapters (-> libamrvr_b200.so).  Same
    entry points as Ref; needs a B200 to run anything."""

    LIB = DROPIN_LIB


def read_ppm(path) -> np.ndarray:
    """P6 reader for the files SavePPM writes (top row first)."""
    raw = Path(path).read_bytes()
    parts = raw.split(b"\n", 3)
    assert parts[0] == b"P6", parts[0]
    w, h = [int(x) for x in parts[1].split()]
    assert parts[2] == b"255"
    return np.frombuffer(parts[3], np.uint8, count=w * h * 3).reshape(h, w, 3).copy()
