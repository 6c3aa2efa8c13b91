"""CPU tests for the data formats either side of the path (SURVEY.md 8f rows f2,
f3): the plotfile reader and convexify behind the C ABI against the numpy
restatement in oracle/amr_io_oracle.py, the automatic camera against the
oracle, and the argument checks of render().  PARITY UNPINNED for f2/f3 (AMReX
is not vendored and the reference holds no fixture for them): the checks are
byte-level for FAB payloads and set-level for convexify."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes
from amrvolumerenderer_b200.plotfile import LevelSpec, PlotFile, synthetic_hierarchy, write_plotfile
from oracle import amr_io_oracle as aio

_P = C.POINTER


def _convexify(lib, coarse, fine, ratio):
    coarse = np.ascontiguousarray(coarse, np.int32).reshape(-1, 6)
    fine = np.ascontiguousarray(fine, np.int32).reshape(-1, 6)
    r = (C.c_int32 * 3)(*ratio)
    n = C.c_int32(0)
    cap = 16
    while True:
        out = np.zeros((cap, 6), np.int32)
        src = np.zeros(cap, np.int32)
        st = lib.amrvr_convexify_boxes(len(coarse), coarse.ctypes.data_as(_P(C.c_int32)), len(fine),
                                       fine.ctypes.data_as(_P(C.c_int32)), r,
                                       out.ctypes.data_as(_P(C.c_int32)),
                                       src.ctypes.data_as(_P(C.c_int32)), cap, C.byref(n))
        if st == 0:
            return out[:n.value], src[:n.value]
        assert n.value > cap, lib.amrvr_last_error(None)
        cap = n.value


@pytest.mark.parametrize("precision,big_endian,nghost,ncomp", [
    ("double", False, 0, 1), ("double", False, 2, 3), ("single", False, 0, 2),
    ("double", True, 1, 1), ("single", True, 0, 1)])
def test_plotfile_reader_matches_payload(lib, tmp_path, precision, big_endian, nghost, ncomp):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=7, ncomp=ncomp,
                                           component=ncomp - 1, extent=(1.0, 1.37, 0.9))
    names = [f"var{c}" for c in range(ncomp)]
    path = write_plotfile(tmp_path / "plt00000", levels, names, plo, phi, precision=precision,
                          big_endian=big_endian, nghost=nghost, fabs_per_file=3)
    want = aio.read_plotfile(path)
    with PlotFile(path) as pf:
        assert pf.space_dim == 3 and pf.finest_level == 2 and pf.var_names == names
        assert list(pf.prob_lo) == want["prob_lo"] and list(pf.prob_hi) == want["prob_hi"]
        assert pf.var_index(names[-1]) == ncomp - 1 and pf.var_index("nope") == -1
        for l, L in enumerate(want["levels"]):
            meta = pf.level(l)
            assert list(meta.cell_size) == L["cell_size"]
            assert list(meta.domain_lo) + list(meta.domain_hi) == list(L["domain"])
            assert list(meta.ref_ratio) == ([2, 2, 2] if l < 2 else [0, 0, 0])
            assert meta.real_bytes == (8 if precision == "double" else 4)
            boxes = pf.boxes(l)
            assert [tuple(b) for b in boxes] == L["boxes"]
            for b in (0, len(boxes) // 2, len(boxes) - 1):
                for c in range(ncomp):
                    got = pf.read_fab(l, b, c)
                    assert np.array_equal(got, L["fabs"][b][c])  # bit-exact incl. float widening
                    src = levels[l].boxes[b][1][c]
                    if precision == "double":
                        assert np.array_equal(got, src)
                    else:
                        assert np.array_equal(got, src.astype(np.float32).astype(np.float64))


@pytest.mark.parametrize("nranks,rule", [(1, 0), (3, 0), (3, 1), (5, 1)])
def test_plotfile_read_level_deals_every_box_once(lib, tmp_path, nranks, rule):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=4, nlevels=2, seed=3)
    path = write_plotfile(tmp_path / "plt", levels, ("rho",), plo, phi, fabs_per_file=5)
    with PlotFile(path) as pf:
        for l in range(2):
            seen = []
            for r in range(nranks):
                idx, off = pf.level_share(l, r, nranks, rule)
                dst = np.full(int(off[-1]) + 3, np.nan)
                idx2, off2 = pf.read_level(l, 0, dst, r, nranks, rule, threads=4)
                assert np.array_equal(idx, idx2) and np.array_equal(off, off2)
                for k, b in enumerate(idx):
                    want = levels[l].boxes[b][1][0]
                    assert np.array_equal(dst[off[k]:off[k + 1]].reshape(want.shape), want)
                assert np.isnan(dst[int(off[-1]):]).all()
                seen += list(idx)
                if rule == 0:
                    assert all(b % nranks == r for b in idx)
            assert sorted(seen) == list(range(len(levels[l].boxes)))


def test_plotfile_errors(lib, tmp_path):
    with pytest.raises(ValueError, match="must not be empty"):
        PlotFile("")
    with pytest.raises(RuntimeError, match="does not exist"):
        PlotFile(tmp_path / "missing")
    bad = tmp_path / "bad"
    bad.mkdir()
    (bad / "Header").write_text("HyperCLaw-V1.1\n1\nrho\n3\n0.0\n0\n")
    with pytest.raises(RuntimeError, match="Header"):
        PlotFile(bad)
    levels, plo, phi = synthetic_hierarchy(base_cells=8, box_cells=8, nlevels=1)
    path = write_plotfile(tmp_path / "v2", levels, ("rho",), plo, phi)
    h = (path / "Level_0" / "Cell_H").read_text().split("\n")
    h[0] = "3"
    (path / "Level_0" / "Cell_H").write_text("\n".join(h))
    with pytest.raises(RuntimeError, match="version 3 is not supported"):
        PlotFile(path)
    path = write_plotfile(tmp_path / "ok", levels, ("rho",), plo, phi)
    with PlotFile(path) as pf:
        with pytest.raises(RuntimeError, match="component out of range"):
            pf.read_fab(0, 0, 4)
        with pytest.raises(RuntimeError, match="bad level"):
            pf.level(3)
    # a BoxArray entry with lo > hi, an index that does not fit, a FAB record that
    # does not cover its entry: refused when read, never turned into offsets
    good = (path / "Level_0" / "Cell_H").read_text()
    for old, new, msg in (("((0,0,0) (7,7,7)", "((9,0,0) (7,7,7)", "empty or implausibly large"),
                          ("((0,0,0) (7,7,7)", "((0,0,0) (7,7,99999999999)", "index out of range"),
                          ("((0,0,0) (7,7,7)", "((0,0,0) (7,7,3000000)", "empty or implausibly large")):
        assert old in good
        (path / "Level_0" / "Cell_H").write_text(good.replace(old, new, 1))
        with pytest.raises(RuntimeError, match=msg):
            PlotFile(path)
    (path / "Level_0" / "Cell_H").write_text(good.replace("((0,0,0) (7,7,7)", "((0,0,0) (7,7,8)", 1))
    with PlotFile(path) as pf:
        with pytest.raises(RuntimeError, match="does not cover its BoxArray entry"):
            pf.read_fab(0, 0, 0)
    (path / "Level_0" / "Cell_H").write_text(good)
    data = path / "Level_0" / "Cell_D_00000"
    data.write_bytes(data.read_bytes()[:-100])
    with PlotFile(path) as pf:
        with pytest.raises(RuntimeError, match="short read"):
            pf.read_fab(0, 0, 0)


def test_plotfile_reader_survives_corrupted_files(tmp_path):
    """200 seeded corruptions (truncation, byte flips, dropped / doubled / junk
    header lines) of a two-level plotfile: every one either reads or raises the
    ABI's error -- run in a child process so that a crash would be seen as one."""
    import os
    import subprocess
    import sys

    worker = os.path.join(os.path.dirname(__file__), "fuzz_plotfile_worker.py")
    out = subprocess.run([sys.executable, worker, str(tmp_path), "7", "200"], capture_output=True,
                         text=True, timeout=600)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    opened, refused = (int(v) for v in out.stdout.split()[-2:])
    assert opened + refused == 200 and opened > 20 and refused > 20


def test_convexify_known_answer_cut_order(lib):
    """(0..7)^3 minus the coarsened (4..11)^3 = (2..5)^3: six slabs, z first, low before high."""
    pieces, src = _convexify(lib, [[0, 0, 0, 7, 7, 7]], [[4, 4, 4, 11, 11, 11]], (2, 2, 2))
    assert pieces.tolist() == [[0, 0, 0, 7, 7, 1], [0, 0, 6, 7, 7, 7], [0, 0, 2, 7, 1, 5],
                               [0, 6, 2, 7, 7, 5], [0, 2, 2, 1, 5, 5], [6, 2, 2, 7, 5, 5]]
    assert src.tolist() == [0] * 6
    # nothing to remove / everything removed / negative indices use floor division
    p, s = _convexify(lib, [[0, 0, 0, 3, 3, 3], [4, 0, 0, 7, 3, 3]], np.zeros((0, 6)), (2, 2, 2))
    assert p.tolist() == [[0, 0, 0, 3, 3, 3], [4, 0, 0, 7, 3, 3]] and s.tolist() == [0, 1]
    # a fine box spanning the whole int32 range coarsens without overflow (INT_MIN / 2)
    p, s = _convexify(lib, [[0, 0, 0, 7, 7, 7]], [[-2 ** 31, 0, 0, 2 ** 31 - 1, 3, 3]], (2, 2, 2))
    assert p.tolist() == [[0, 0, 2, 7, 7, 7], [0, 2, 0, 7, 7, 1]]
    p, s = _convexify(lib, [[0, 0, 0, 3, 3, 3]], [[0, 0, 0, 7, 7, 7]], (2, 2, 2))
    assert len(p) == 0
    p, s = _convexify(lib, [[-4, -4, -4, -1, -1, -1]], [[-3, -8, -8, -1, -1, -1]], (2, 2, 2))
    assert p.tolist() == [[-4, -4, -4, -3, -1, -1]]  # floor(-3/2) = -2: x in [-2,-1] is covered


@pytest.mark.parametrize("seed", range(8))
def test_convexify_is_an_exact_disjoint_cover(lib, seed):
    rng = np.random.default_rng(100 + seed)
    ratio = [(2, 2, 2), (4, 4, 4), (2, 4, 2), (3, 3, 3)][seed % 4]
    n = 24
    domain = (0, 0, 0, n - 1, n - 1, n - 1)
    cb = [4, 6, 8, 12][seed % 4]
    coarse = [(i, j, k, i + cb - 1, j + cb - 1, k + cb - 1) for k in range(0, n, cb)
              for j in range(0, n, cb) for i in range(0, n, cb) if rng.random() > 0.15]
    fine = []
    for _ in range(int(rng.integers(1, 9))):
        lo = rng.integers(0, n - 2, 3)
        hi = np.minimum(lo + rng.integers(1, 10, 3), n - 1)
        fine.append([int(lo[d] * ratio[d]) for d in range(3)] +
                    [int((hi[d] + 1) * ratio[d] - 1) for d in range(3)])
    pieces, src = _convexify(lib, coarse, fine, ratio)
    want = aio.uncovered_mask(domain, coarse, fine, ratio)
    got, most = aio.pieces_mask(domain, pieces)
    assert most <= 1, "pieces overlap"
    assert np.array_equal(got, want)
    for p, s in zip(pieces, src):  # every piece inside the coarse box it names
        c = coarse[s]
        assert all(c[d] <= p[d] and p[3 + d] <= c[3 + d] for d in range(3))
    assert list(src) == sorted(src)


def test_convexify_levels_returns_views(lib):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=11)
    nlev = len(levels)
    clev = (api.Level * nlev)()
    keep = []
    for l, L in enumerate(levels):
        fabs = (api.Fab * len(L.boxes))()
        for f, (lo, data) in zip(fabs, L.boxes):
            arr = np.ascontiguousarray(data[0])
            keep.append(arr)
            nz, ny, nx = arr.shape
            for d in range(3):
                f.lo[d], f.begin[d] = lo[d], lo[d]
                f.hi[d] = lo[d] + (nx, ny, nz)[d] - 1
            f.data, f.jstride, f.kstride, f.nstride = arr.ctypes.data, nx, nx * ny, nx * ny * nz
        keep.append(fabs)
        clev[l].nfabs, clev[l].fabs = len(L.boxes), fabs
        for d in range(3):
            clev[l].prob_lo[d], clev[l].cell_size[d] = plo[d], L.cell_size[d]
    ratios = (C.c_int32 * 3 * (nlev - 1))()
    for l in range(nlev - 1):
        for d in range(3):
            ratios[l][d] = 2
    out = (api.Fab * 256)()
    counts = (C.c_int32 * nlev)()
    assert lib.amrvr_convexify_levels(nlev, clev, None, None, ratios, out, 256, counts) == 0
    # 8 boxes of 8^3 per level; the central 8^3 of each of the two coarser levels is refined:
    # every coarse box loses one 4^3 corner -> 3 pieces each
    assert list(counts) == [24, 24, 8]
    base = 0
    for l in range(nlev):
        cells = 0
        for k in range(counts[l]):
            v = out[base + k]
            owner = [f for f in clev[l].fabs[:clev[l].nfabs]
                     if all(f.lo[d] <= v.lo[d] and v.hi[d] <= f.hi[d] for d in range(3))]
            assert len(owner) == 1
            f = owner[0]
            assert (v.data, v.jstride, v.kstride, v.nstride, list(v.begin)) == \
                   (f.data, f.jstride, f.kstride, f.nstride, list(f.begin))
            cells += int(np.prod([v.hi[d] - v.lo[d] + 1 for d in range(3)]))
        assert cells == (16 ** 3 - 8 ** 3 if l < nlev - 1 else 16 ** 3)
        base += counts[l]
    assert lib.amrvr_convexify_levels(nlev, clev, None, None, ratios, out, 10, counts) == api_invalid()


def api_invalid():
    return 1  # AMRVR_ERR_INVALID


@pytest.mark.parametrize("bmin,bmax,up", [((-0.05, -0.05, -0.05), (1.05, 1.05, 1.05), None),
                                          ((-0.07, -0.07, -0.07), (1.07, 1.44, 0.97), (0.0, 0.0, 1.0)),
                                          ((0.0, 0.0, 0.0), (0.0, 0.0, 0.0), None)])
def test_auto_camera_bit_exact(lib, oracle, bmin, bmax, up):
    az, alt = scenes.auto_camera_draws(91021)
    want = oracle.auto_camera(bmin, bmax, az, alt, up)
    got = api.Camera()
    d3 = lambda v: (C.c_double * 3)(*v)
    assert lib.amrvr_auto_camera(d3(bmin), d3(bmax), 91021, d3(up) if up else None, C.byref(got)) == 0
    for name in ("eye", "look_at", "up"):
        assert list(getattr(got, name)) == list(getattr(want, name)), name
    assert (got.fov_y_degrees, got.near_plane, got.far_plane) == \
           (want.fov_y_degrees, want.near_plane, want.far_plane)


def test_render_argument_errors(tmp_path):
    """The checks of module.cpp:131-231 and VolumeRenderer::run (:1341-1467) fire before any GPU work."""
    from amrvolumerenderer_b200.render import render

    p = str(tmp_path / "nothing")
    cases = [
        (dict(width=0), ValueError, "image dimensions"),
        (dict(box_transparency=1.5), ValueError, "box transparency"),
        (dict(antialiasing=3), ValueError, "perfect square"),
        (dict(antialiasing=0), ValueError, "antialiasing must be positive"),
        (dict(output=""), ValueError, "output filename"),
        (dict(min_level=-1), ValueError, "min level"),
        (dict(max_level=-2), ValueError, "max level"),
        (dict(min_level=2, max_level=1), RuntimeError, "must not exceed"),
        (dict(up_vector=(0, 0, 0)), ValueError, "up_vector"),
        (dict(scalar_range=(1.0, 1.0)), ValueError, "scalar_range"),
        (dict(color_map=[(0.0, 1, 0, 0, 1)]), ValueError, "at least two"),
        (dict(color_map=[(0.0, 1, 0, 0, 1), (0.0, 0, 1, 0, 1)]), ValueError, "strictly increasing"),
        (dict(color_map=[(0.0, 1, 0, 0, 1), (1.0, 0, 2, 0, 1)]), ValueError, "green"),
        (dict(camera_fov_y=30.0), ValueError, "camera_eye and camera_look_at"),
        (dict(camera_eye=(1, 1, 1), camera_look_at=(1, 1, 1)), ValueError, "distinct"),
        (dict(camera_eye=(0, 0, 2), camera_look_at=(0, 0, 0), camera_up=(0, 0, 1)), ValueError, "parallel"),
        (dict(camera_eye=(0, 0, 2), camera_look_at=(0, 0, 0), camera_fov_y=180.0), ValueError, "fov"),
        (dict(camera_eye=(0, 0, 2), camera_look_at=(0, 0, 0), camera_near=2.0, camera_far=1.0),
         ValueError, "far plane"),
        (dict(), RuntimeError, "does not exist"),
    ]
    for kwargs, exc, text in cases:
        with pytest.raises(exc, match=text):
            render(p, **kwargs)
    with pytest.raises(RuntimeError, match="plotfile path is required"):
        render("")


def test_image_writers_round_trip(tmp_path, capsys):
    """PPM as Common/SavePPM.cpp lays it out, PNG readable by a plain zlib decoder,
    unknown extensions fall back to PPM with the reference's warning (VolumeRenderer.cpp:1316-1328)."""
    import struct
    import zlib

    from amrvolumerenderer_b200.render import save_image
    from oracle.oracle import read_ppm

    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    save_image(tmp_path / "a.ppm", img)
    assert np.array_equal(read_ppm(tmp_path / "a.ppm"), img)
    save_image(tmp_path / "b.PNG", img)
    raw = (tmp_path / "b.PNG").read_bytes()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat = 8, b""
    while pos < len(raw):
        n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
        body = raw[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body)
        if tag == b"IHDR":
            assert struct.unpack(">IIBBBBB", body) == (53, 37, 8, 2, 0, 0, 0)
        if tag == b"IDAT":
            idat += body
        pos += 12 + n
    rows = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(37, 1 + 3 * 53)
    assert not rows[:, 0].any() and np.array_equal(rows[:, 1:].reshape(37, 53, 3), img)
    save_image(tmp_path / "c.jpeg", img)
    assert "unrecognized image extension '.jpeg'" in capsys.readouterr().err
    assert np.array_equal(read_ppm(tmp_path / "c.jpeg"), img)


def test_convexify_three_levels_is_a_partition_of_the_domain(lib):
    """After convexify every point of the coarsest domain is covered by exactly one level's pieces."""
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=4, nlevels=3, seed=2)
    boxes = [[tuple(lo) + tuple(lo[d] + 3 for d in range(3)) for lo, _ in L.boxes] for L in levels]
    fine_domain = (0, 0, 0, 63, 63, 63)
    covered = np.zeros((64, 64, 64), np.int32)
    for l in range(3):
        if l < 2:
            pieces, _src = _convexify(lib, boxes[l], boxes[l + 1], (2, 2, 2))
        else:
            pieces = np.array(boxes[l], np.int32)
        scale = 4 >> l  # level-l cell = scale^3 finest cells
        for p in pieces:
            covered[p[2] * scale:(p[5] + 1) * scale, p[1] * scale:(p[4] + 1) * scale,
                    p[0] * scale:(p[3] + 1) * scale] += 1
    assert covered.min() == 1 and covered.max() == 1, fine_domain


def test_runtime_refcount_like_the_reference(monkeypatch):
    """initialize_runtime / finalize_runtime (module.cpp:104-119): nested holds,
    an unmatched finalize raises with the reference's text, and while a hold is
    open render()/compute_histogram() share ONE context that the last finalize
    closes (contexts are faked here: no GPU in this test)."""
    from amrvolumerenderer_b200 import render as R

    made = []

    class FakeContext:
        def __init__(self, device=0):
            self.closed = False
            made.append(self)

        @classmethod
        def from_env(cls):
            return cls()

        def close(self):
            self.closed = True

    monkeypatch.setattr(R, "Context", FakeContext)
    assert R._runtime == {"refs": 0, "context": None}
    with pytest.raises(RuntimeError, match="requires a matching initialize_runtime call"):
        R.finalize_runtime()

    ctx, owned = R._acquire_context(None)  # no hold: one context per call, closed by the caller
    assert owned and len(made) == 1
    mine = object()
    assert R._acquire_context(mine) == (mine, False)  # an explicit context is never taken over

    R.initialize_runtime()
    R.initialize_runtime()
    a, owned_a = R._acquire_context(None)
    b, owned_b = R._acquire_context(None)
    assert a is b and not owned_a and not owned_b and len(made) == 2
    R.finalize_runtime()
    assert not a.closed and R._acquire_context(None)[0] is a
    R.finalize_runtime()
    assert a.closed and R._runtime == {"refs": 0, "context": None}
    with pytest.raises(RuntimeError, match="requires a matching initialize_runtime call"):
        R.finalize_runtime()

    import amrvolumerenderer_b200 as pkg

    assert pkg.initialize_runtime is R.initialize_runtime and pkg.compute_histogram is R.compute_histogram
