"""Pins the host logic of render() -- the restatement of VolumeRenderer::run
(VolumeRenderer/VolumeRenderer.cpp:1341-1576) in amrvolumerenderer_b200/render.py
plus amrvr_auto_camera -- against the reference's OWN run(), executed on the CPU
from oracle/_ref over a plotfile whose content is handed to the shim's
PlotFileData (single level: the shim's convexify is the identity there).

Our side never touches the GPU here: scene summary from the oracle restatement
of BuildSceneGeometry (itself pinned to the reference in test_cpu_scene.py),
then resolve_scalar_options / _explicit_camera / amrvr_auto_camera, then the
oracle frame.  The two PPM payloads must be identical byte for byte."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes
from amrvolumerenderer_b200.plotfile import synthetic_hierarchy
from amrvolumerenderer_b200.render import (DEFAULT_CAMERA_SEED, _explicit_camera, render,
                                           resolve_scalar_options)
from oracle.oracle import read_ppm


def _levels_struct(levels, component):
    clev = (api.Level * len(levels))()
    keep = []
    for l, L in enumerate(levels):
        fabs = (api.Fab * len(L.boxes))()
        for f, (lo, data) in zip(fabs, L.boxes):
            arr = np.ascontiguousarray(data[component])
            keep.append(arr)
            nz, ny, nx = arr.shape
            for d, ext in enumerate((nx, ny, nz)):
                f.lo[d], f.begin[d], f.hi[d] = lo[d], lo[d], lo[d] + ext - 1
            f.data, f.jstride, f.kstride, f.nstride = arr.ctypes.data, nx, nx * ny, nx * ny * nz
        keep.append(fabs)
        clev[l].nfabs, clev[l].fabs = len(L.boxes), fabs
        for d in range(3):
            clev[l].prob_lo[d], clev[l].cell_size[d] = 0.0, L.cell_size[d]
    return clev, keep


def _our_frame(lib, oracle, levels, component, nranks, params, log_scale=False, scalar_range=None,
               color_map=None, up_vector=None, camera_kw=None):
    clev, keep = _levels_struct(levels, component)
    boxes, summary, status = oracle.scene_build(clev, len(levels), log_scale=log_scale,
                                                normalize_to_data_range=scalar_range is None)
    assert status == 0
    transform, value_range, points = resolve_scalar_options(summary, log_scale, scalar_range, color_map)
    if camera_kw:
        camera = _explicit_camera(camera_kw.get("eye"), camera_kw.get("look_at"), camera_kw.get("up"),
                                  camera_kw.get("fov_y"), camera_kw.get("near"), camera_kw.get("far"))
    else:
        camera = api.Camera()
        up = None
        if up_vector is not None:
            v = np.array(up_vector, np.float32).astype(np.float64)
            up = (C.c_double * 3)(*(v / float(np.float32(np.sqrt((v * v).sum())))))
        assert lib.amrvr_auto_camera(summary.bounds_min, summary.bounds_max, DEFAULT_CAMERA_SEED, up,
                                     C.byref(camera)) == 0
    specs = []
    for i in range(summary.nboxes):
        b = boxes[i]
        nx, ny, nz = b.dims
        data = np.ctypeslib.as_array((C.c_double * (nx * ny * nz)).from_address(b.data)).reshape(nz, ny, nx)
        specs.append(scenes.BoxSpec(0, tuple(b.lo), tuple(b.dims), np.array(b.min_corner),
                                    np.array(b.max_corner), np.ascontiguousarray(data), i % nranks, i))
    specs.sort(key=lambda s: s.owner)  # owner-major, as an allgather delivers them
    scene = scenes.Scene("run", specs, transform, value_range, np.array(summary.bounds_min),
                         np.array(summary.bounds_max), nranks, [])
    return oracle.render_frame(scene, camera, params, colormap=points)["rgb8"], keep


CASES = {
    "defaults_auto_camera": dict(),
    "explicit_camera": dict(camera_kw=dict(eye=(1.9, 1.3, 2.4), look_at=(0.5, 0.5, 0.5), far=10.0)),
    "override_and_colormap": dict(scalar_range=(0.1, 0.55),
                                  color_map=[(0.05, 0, 0, .4, 0), (0.3, .1, .8, .3, .2), (0.7, 1, .9, .1, .9)]),
    "log_override_colormap_up": dict(log_scale=True, scalar_range=(0.04, 0.5), up_vector=(0.0, 0.0, 2.0),
                                     color_map=[(0.05, 0, 0, .4, 0), (0.2, .1, .8, .3, .2), (0.45, 1, .9, .1, .9)]),
    "log_scale_data_range": dict(log_scale=True),
    "two_ranks_aa4": dict(nranks=2, aa=4, transparency=0.15),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_run_matches_the_reference_byte_for_byte(lib, oracle, ref, tmp_path, case):
    kw = dict(CASES[case])
    nranks, aa, transparency = kw.pop("nranks", 1), kw.pop("aa", 1), kw.pop("transparency", 0.0)
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=1, seed=31, ncomp=2,
                                           component=1, extent=(1.0, 1.25, 0.8))
    key = f"registered://{case}"
    ref.register_plotfile(key, levels, ("junk", "rho"))
    params = api.FrameParams(width=72, height=56, box_transparency=transparency, antialiasing=aa,
                             use_visibility_graph=1, count_samples=0, draw_bounding_box=1)
    out = tmp_path / "ref.ppm"
    cam = None
    if kw.get("camera_kw"):
        c = kw["camera_kw"]
        cam = _explicit_camera(c.get("eye"), c.get("look_at"), c.get("up"), c.get("fov_y"), c.get("near"),
                               c.get("far"))
    log = ref.run(key, out, params, nranks=nranks, variable="rho", log_scale=kw.get("log_scale", False),
                  up_vector=(np.array(kw["up_vector"]) / np.linalg.norm(kw["up_vector"])
                             if kw.get("up_vector") else None),
                  scalar_range=kw.get("scalar_range"), camera=cam, colormap=kw.get("color_map"))
    assert "Loaded plotfile" in log
    want = read_ppm(out)
    got, keep = _our_frame(lib, oracle, levels, 1, nranks, params, **kw)
    assert want.any()
    assert np.array_equal(got, want), int(np.abs(got.astype(int) - want.astype(int)).max())


def test_run_error_messages_match_the_reference(lib, ref, tmp_path):
    """Same exception class and text from the reference's run() and from render()
    for the checks that fire before any data is touched."""
    levels, plo, phi = synthetic_hierarchy(base_cells=8, box_cells=8, nlevels=1, seed=1)
    key = str(tmp_path / "exists")  # render() looks for the path on disk, the shim in its registry
    (tmp_path / "exists").mkdir()
    ref.register_plotfile(key, levels, ("rho",))
    ok = dict(width=32, height=32, box_transparency=0.0, antialiasing=1)
    cases = [
        (dict(ok, width=0), {}, {}),
        (dict(ok, box_transparency=2.0), {}, {}),
        (dict(ok, antialiasing=2), {}, {}),
        (ok, dict(min_level=-1), dict(min_level=-1)),
        (ok, dict(max_level=-3), dict(max_level=-3)),
        (ok, dict(min_level=3, max_level=1), dict(min_level=3, max_level=1)),
        (ok, dict(scalar_range=(2.0, 1.0)), dict(scalar_range=(2.0, 1.0))),
        (ok, dict(colormap=[(0.0, 1, 0, 0, 1)]), dict(color_map=[(0.0, 1, 0, 0, 1)])),
        (ok, dict(colormap=[(0.5, 1, 0, 0, 1), (0.25, 0, 1, 0, 1)]),
         dict(color_map=[(0.5, 1, 0, 0, 1), (0.25, 0, 1, 0, 1)])),
        (ok, dict(colormap=[(0.0, 1, 0, 0, 1), (1.0, 0, 0, 1.5, 1)]),
         dict(color_map=[(0.0, 1, 0, 0, 1), (1.0, 0, 0, 1.5, 1)])),
    ]
    for frame, ref_kw, our_kw in cases:
        params = api.FrameParams(use_visibility_graph=1, draw_bounding_box=1, **frame)
        with pytest.raises(api.AmrvrError) as theirs:
            ref.run(key, tmp_path / "x.ppm", params, **ref_kw)
        expected = ValueError if theirs.value.status == api.AMRVR_ERR_INVALID else RuntimeError
        with pytest.raises(expected) as ours:
            render(key, output=str(tmp_path / "y.ppm"), **frame, **our_kw)
        if "scalar_range" in our_kw:
            # the nanobind wrapper checks this one itself and words it with the keyword
            # (module.cpp:190-198 "scalar_range must ..."; run() says "scalar range must ...")
            assert str(ours.value) == theirs.value.message.replace("scalar range", "scalar_range")
        else:
            assert str(ours.value) == theirs.value.message, (frame, ref_kw)


def _convexified(lib, levels, component, upto):
    """amrvr_convexify_levels over levels[0..upto] -> (convex Level array, pieces per level, keep)."""
    nlev = upto + 1
    clev, keep = _levels_struct(levels[:nlev], component)
    ratios = (C.c_int32 * 3 * max(nlev - 1, 1))()
    for l in range(nlev - 1):
        for d in range(3):
            ratios[l][d] = 2
    out = (api.Fab * 1024)()
    counts = (C.c_int32 * nlev)()
    assert lib.amrvr_convexify_levels(nlev, clev, None, None, ratios, out, 1024, counts) == 0
    convex = (api.Level * nlev)()
    pieces, base = [], 0
    for l in range(nlev):
        convex[l].prob_lo, convex[l].cell_size = clev[l].prob_lo, clev[l].cell_size
        convex[l].nfabs = counts[l]
        convex[l].fabs = C.cast(C.byref(out, base * C.sizeof(api.Fab)), C.POINTER(api.Fab))
        pieces.append(np.array([[out[base + k].lo[d] for d in range(3)] + [out[base + k].hi[d] for d in range(3)]
                                for k in range(counts[l])], np.int32).reshape(-1, 6))
        base += counts[l]
    return convex, pieces, (keep, out, clev)


def _view(b):
    """The cells of an amrvr_box (a strided view into its source FAB) as a dense (nz, ny, nx) array."""
    nx, ny, nz = b.dims
    first = b.data + 8 * ((b.lo[0] - b.begin[0]) + (b.lo[1] - b.begin[1]) * b.jstride
                          + (b.lo[2] - b.begin[2]) * b.kstride)
    span = (nz - 1) * b.kstride + (ny - 1) * b.jstride + nx
    flat = np.ctypeslib.as_array((C.c_double * span).from_address(first))
    return np.ascontiguousarray(np.lib.stride_tricks.as_strided(
        flat, shape=(nz, ny, nx), strides=(8 * b.kstride, 8 * b.jstride, 8)))


@pytest.mark.parametrize("min_level,max_level,nranks", [(0, -1, 1), (1, 2, 1), (0, 1, 1), (0, -1, 2)])
def test_multi_level_run_with_our_cuts_matches_the_reference(lib, oracle, ref, tmp_path, min_level,
                                                             max_level, nranks):
    """Three-level, NON-convex plotfile: the reference's run() is handed the box cuts
    OUR convexify produced (through the shim's convexify override) and must then
    write the same bytes as our chain -- views into the source FABs, scene build,
    level window, automatic camera, frame.  What stays unpinned is only which
    cuts AMReX itself would choose."""
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=41)
    key = f"registered://multi/{min_level}/{max_level}/{nranks}"
    ref.register_plotfile(key, levels, ("rho",))
    finest = 2
    hi = finest if (max_level < 0 or max_level > finest) else max_level
    convex, pieces, keep = _convexified(lib, levels, 0, hi)
    params = api.FrameParams(width=64, height=48, box_transparency=0.0, antialiasing=1,
                             use_visibility_graph=1, count_samples=0, draw_bounding_box=1)
    out = tmp_path / "ref.ppm"
    ref.set_convexify_override(pieces)
    try:
        ref.run(key, out, params, nranks=nranks, min_level=min_level, max_level=max_level)
    finally:
        ref.set_convexify_override(None)
    want = read_ppm(out)

    boxes, summary, status = oracle.scene_build(convex, hi + 1, min_level=min_level, max_level=hi)
    assert status == 0
    transform, value_range, points = resolve_scalar_options(summary, False, None, None)
    camera = api.Camera()
    assert lib.amrvr_auto_camera(summary.bounds_min, summary.bounds_max, DEFAULT_CAMERA_SEED, None,
                                 C.byref(camera)) == 0
    # the shim deals SOURCE box b of every level to rank b % nranks and a piece stays with its source
    owner_of = {}
    for l in range(hi + 1):
        for b, (lo, data) in enumerate(levels[l].boxes):
            owner_of[(l, tuple(lo))] = b % nranks
    specs = []
    for i in range(summary.nboxes):
        b = boxes[i]
        level = int(np.argmin([abs(levels[l].cell_size[0] * summary.global_scale
                                   - (b.max_corner[0] - b.min_corner[0]) / b.dims[0]) for l in range(3)]))
        src_lo = tuple((b.lo[d] // 8) * 8 for d in range(3))  # 8^3 source boxes on an 8-aligned lattice
        specs.append(scenes.BoxSpec(level, tuple(b.lo), tuple(b.dims), np.array(b.min_corner),
                                    np.array(b.max_corner), _view(b), owner_of[(level, src_lo)], i))
    specs.sort(key=lambda s: s.owner)
    scene = scenes.Scene("multi", specs, transform, value_range, np.array(summary.bounds_min),
                         np.array(summary.bounds_max), nranks, [])
    got = oracle.render_frame(scene, camera, params)["rgb8"]
    assert want.any()
    assert np.array_equal(got, want), int(np.abs(got.astype(int) - want.astype(int)).max())
