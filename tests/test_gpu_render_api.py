"""GPU parity for the reference's Python entry points (module.cpp:131-356) over
the full input side: plotfile -> convexify -> BuildSceneGeometry -> frame ->
image file.  The plotfiles hold NON-convex hierarchies (coarse data under the
fine levels, as a simulation dumps them); the oracle then renders the box list
our convexify produced, so the comparison pins everything but convexify's
exact cuts (unpinned, see oracle/amr_io_oracle.py): float RGBA <= 1e-4, bytes
<= 1 LSB, histogram counts bit-exact."""
import struct
import zlib

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes
from amrvolumerenderer_b200.plotfile import synthetic_hierarchy, write_plotfile
from amrvolumerenderer_b200.render import compute_histogram, render
from oracle.oracle import read_ppm

pytestmark = pytest.mark.gpu


def _scene_from_frame(out, levels, extent, component=0, precision="double"):
    """scenes.Scene over the convex boxes render() reports, data cut from the hierarchy."""
    scale = out["summary"].global_scale
    boxes = []
    for gi, (bmin, bmax, dims, lo) in enumerate(out["boxes"]):
        dx = (bmax[0] - bmin[0]) / dims[0]
        level = int(np.argmin([abs(L.cell_size[0] * scale - dx) for L in levels]))
        data = None
        for slo, sdata in levels[level].boxes:
            sd = sdata[component] if sdata.ndim == 4 else sdata
            nz, ny, nx = sd.shape
            if all(slo[d] <= lo[d] and lo[d] + dims[d] <= slo[d] + (nx, ny, nz)[d] for d in range(3)):
                data = sd[lo[2] - slo[2]:lo[2] - slo[2] + dims[2], lo[1] - slo[1]:lo[1] - slo[1] + dims[1],
                          lo[0] - slo[0]:lo[0] - slo[0] + dims[0]]
                break
        assert data is not None
        if precision == "single":
            data = data.astype(np.float32)
        data = np.ascontiguousarray(data, dtype=np.float64)
        boxes.append(scenes.BoxSpec(level, lo, dims, np.array(bmin), np.array(bmax), data, 0, gi))
    s = out["summary"]
    return scenes.Scene("plotfile", boxes, out["transform"], out["scalar_range"], np.array(s.bounds_min),
                        np.array(s.bounds_max), 1, [])


def _params(width, height, aa=1, transparency=0.0, graph=True):
    return api.FrameParams(width=width, height=height, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=int(graph), count_samples=0, draw_bounding_box=1)


def _check(out, want):
    err = float(np.abs(out["final5"][..., :4] - want["final5"][..., :4]).max())
    lsb = int(np.abs(out["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
    assert err <= 1e-4, err
    assert lsb <= 1, lsb
    assert out["rgb8"].any()


def _read_png(path):
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, size = 8, b"", None
    while pos < len(raw):
        n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body)
        if tag == b"IHDR":
            size = struct.unpack(">II", body[:8])
            assert body[8:] == bytes([8, 2, 0, 0, 0])
        elif tag == b"IDAT":
            idat += body
        pos += 12 + n
    w, h = size
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 1 + 3 * w)
    assert not rows[:, 0].any()
    return rows[:, 1:].reshape(h, w, 3)


def test_render_three_level_plotfile_explicit_camera(ctx, oracle, tmp_path):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=21)
    path = write_plotfile(tmp_path / "plt00010", levels, ("density",), plo, phi)
    ppm = tmp_path / "out.ppm"
    out = render(str(path), width=96, height=64, output=str(ppm), camera_eye=(1.9, 1.3, 2.4),
                 camera_look_at=(0.5, 0.5, 0.5), camera_far=10.0, context=ctx, return_frame=True)
    assert len(out["boxes"]) == 24 + 24 + 8  # each coarse box lost one corner to the finer level
    scene = _scene_from_frame(out, levels, phi)
    assert scene.total_cells() == 2 * (16 ** 3 - 8 ** 3) + 16 ** 3
    want = oracle.render_frame(scene, out["camera"], _params(96, 64))
    _check(out, want)
    assert np.array_equal(read_ppm(ppm), out["rgb8"])


def test_render_auto_camera_override_colormap_png(ctx, oracle, tmp_path):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=2, seed=22, ncomp=3,
                                           component=1, extent=(1.0, 1.37, 0.9))
    path = write_plotfile(tmp_path / "plt", levels, ("a", "rho", "c"), plo, phi, precision="single", nghost=1)
    png = tmp_path / "img.PNG"
    cmap = [(0.05, 0.0, 0.0, 0.4, 0.0), (0.2, 0.1, 0.8, 0.3, 0.2), (0.45, 1.0, 0.9, 0.1, 0.9)]
    out = render(str(path), width=80, height=60, box_transparency=0.15, antialiasing=4, variable="rho",
                 log_scale=True, up_vector=(0.0, 0.0, 2.0), output=str(png), scalar_range=(0.04, 0.5),
                 color_map=cmap, context=ctx, return_frame=True)
    t = out["transform"]
    assert t.log_scale_input == 1 and t.normalize_to_unit_range == 1
    assert t.normalization_min == float(np.float32(np.log(float(np.float32(0.04)))))
    assert out["scalar_range"] == (0.0, 1.0)
    assert [p[0] for p in out["color_map"]] == sorted(p[0] for p in out["color_map"])
    assert list(out["camera"].up) == [0.0, 0.0, 1.0]
    scene = _scene_from_frame(out, levels, phi, component=1, precision="single")
    want = oracle.render_frame(scene, out["camera"], _params(80, 60, aa=4, transparency=0.15),
                               colormap=out["color_map"])
    _check(out, want)
    assert np.array_equal(_read_png(png), out["rgb8"])


def test_render_level_window_and_no_graph(ctx, oracle, tmp_path):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=23)
    path = write_plotfile(tmp_path / "plt", levels, ("density",), plo, phi, big_endian=True)
    out = render(str(path), width=64, height=64, min_level=1, max_level=1, visibility_graph=False,
                 output=str(tmp_path / "w.ppm"), context=ctx, return_frame=True)
    # level 1 alone, minus what level 2 covers... max_level=1 means level 2 is never loaded
    assert len(out["boxes"]) == 8
    scene = _scene_from_frame(out, levels, phi)
    assert {b.level for b in scene.boxes} == {1}
    want = oracle.render_frame(scene, out["camera"], _params(64, 64, graph=False))
    _check(out, want)
    with pytest.raises(RuntimeError, match="not found in plotfile"):
        render(str(path), variable="pressure", context=ctx)


def test_compute_histogram_matches_oracle(ctx, oracle, tmp_path):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=24)
    path = write_plotfile(tmp_path / "plt", levels, ("density",), plo, phi)
    got = compute_histogram(str(path), bins=64, context=ctx)
    out = render(str(path), width=16, height=16, output=str(tmp_path / "t.ppm"), context=ctx, return_frame=True)
    scene = _scene_from_frame(out, levels, phi)
    cboxes, n, keep = scene.c_boxes()
    want, total = oracle.scene_histogram(cboxes, n, out["transform"], out["scalar_range"], 64)
    assert got["samples"] == total == scene.total_cells()
    assert got["counts"] == [int(c) for c in want]
    assert got["normalized_range"] == (0.0, 1.0)
    assert got["original_range"][0] < got["original_range"][1]
    with pytest.raises(ValueError, match="binCount"):
        compute_histogram(str(path), bins=0, context=ctx)
