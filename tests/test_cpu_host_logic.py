"""CPU suite, part 2: the host logic of libamrvr_b200.so ([host-only] entry
points of include/amrvr.h) against the oracle, and the shape of the ABI itself.
No compute call is made -- there is no GPU here."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, engine, scenes


def test_library_exports_every_declared_symbol(lib):
    names = api.declared_symbols()
    assert len(names) >= 40
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing
    assert lib.amrvr_abi_version() == 2


def test_every_declared_symbol_has_a_ctypes_prototype(lib):
    protos = api._declare(lib)
    assert sorted(protos) == api.declared_symbols()


def test_no_cpu_fallback(lib):
    """Without a usable sm_100 device the context cannot be created."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(api.AmrvrError) as e:
        engine.Context(device=0)
    assert e.value.status == api.AMRVR_ERR_CUDA
    assert "no CPU fallback" in e.value.message


def test_product_package_never_imports_the_oracle():
    import re
    from pathlib import Path

    pkg = Path(api.__file__).resolve().parent
    for path in list(pkg.rglob("*.py")) + list(pkg.rglob("*.cpp")) + list(pkg.rglob("*.cu")) + \
            list(pkg.rglob("*.hpp")):
        text = path.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle", text, re.M), path
        assert "liboracle" not in text and "oracle/restate" not in text, path


def test_struct_sizes_match_the_header(lib):
    # sizes the C compiler gives the PODs of include/amrvr.h (LP64)
    assert C.sizeof(api.Box) == 120
    assert C.sizeof(api.BoxRecord) == 72
    assert C.sizeof(api.Camera) == 88
    assert C.sizeof(api.ScalarTransform) == 64
    assert C.sizeof(api.FrameParams) == 32


@pytest.mark.parametrize("cmap", [None, [(0.0, 0.1, 0.1, 0.6, 0.0), (0.3, 0.2, 0.8, 0.3, 0.05),
                                         (0.7, 0.9, 0.9, 0.1, 0.4), (1.0, 1.0, 0.2, 0.1, 0.9)]])
def test_color_table_matches_oracle(oracle, cmap):
    """buildColorTable (Common/VolumePainter.cpp:442-516), bit for bit."""
    for alpha_scale, norm, rmin, rmax in ((1.0, 1.0, 0.0, 1.0), (0.85, 0.5, 0.0, 1.0),
                                          (0.4, 0.25, 0.1, 0.9), (1.0, 2.0, 0.0, 0.5)):
        got = engine.color_table(alpha_scale, norm, rmin, rmax, cmap)
        want = oracle.color_table(alpha_scale, norm, rmin, rmax, cmap)
        assert np.array_equal(got, want)


@pytest.mark.parametrize("nranks,rule", [(1, "chunks"), (2, "chunks"), (3, "round_robin"), (8, "round_robin")])
def test_host_order_matches_oracle(oracle, nranks, rule):
    scene = scenes.scene_c4(box_cells=4, nranks=nranks, rule=rule)
    recs, n = scene.records()
    rng = np.random.default_rng(17)
    for _ in range(12):
        eye = rng.normal(size=3)
        eye = 0.5 + eye / np.linalg.norm(eye) * rng.uniform(1.2, 3.0)
        cam = scenes.make_camera(eye=eye)
        for graph in (True, False):
            want, ok = oracle.order(nranks, recs, n, cam, 1.25, graph)
            got, failed = engine.order_boxes_host(nranks, recs, n, cam, 1.25, graph)
            assert np.array_equal(got, want)
            assert failed == (not ok)


def test_host_order_random_boxes_with_cycles(oracle):
    rng = np.random.default_rng(23)
    for trial in range(40):
        n = int(rng.integers(1, 48))
        recs = (api.BoxRecord * n)()
        local = {}
        for i in range(n):
            lo = rng.integers(0, 4, size=3) / 4.0
            hi = lo + rng.integers(1, 3, size=3) / 4.0
            for d in range(3):
                recs[i].min_corner[d], recs[i].max_corner[d] = lo[d], hi[d]
                recs[i].dims[d] = 4
            recs[i].owner = int(i * 4 // n)
            recs[i].local_index = local.get(recs[i].owner, 0)
            local[recs[i].owner] = recs[i].local_index + 1
        cam = scenes.make_camera(eye=tuple(rng.uniform(-2, 3, size=3)))
        want, ok = oracle.order(4, recs, n, cam, 1.0)
        got, failed = engine.order_boxes_host(4, recs, n, cam, 1.0)
        assert np.array_equal(got, want), trial
        assert failed == (not ok)


@pytest.mark.parametrize("scale", [1.0, 37.5, 4096.0])
def test_host_order_at_the_touching_tolerance(oracle, scale):
    """Faces within, at and just beyond the 1e-5 * scale tolerance of
    VisibilityOrdering.cpp:211-228, zero-thickness boxes, axis-parallel views and
    coordinates far from the origin: the SIMD pair filter must keep every pair the
    exact test would accept (same rank order as the oracle, which test_cpu_oracle.py
    pins to the reference)."""
    rng = np.random.default_rng(int(scale) + 5)
    tol = 1e-5 * max(1.0, scale)
    for trial in range(60):
        n = int(rng.integers(2, 40))
        recs = (api.BoxRecord * n)()
        local = {}
        for i in range(n):
            lo = rng.integers(0, 3, size=3) * (scale / 3.0)
            ext = rng.integers(0 if trial % 7 == 0 else 1, 3, size=3) * (scale / 3.0)  # 0: a flat box
            jitter = rng.choice([0.0, 0.4 * tol, 0.99 * tol, 1.01 * tol, 2.5 * tol, -0.99 * tol, -3 * tol], size=6)
            for d in range(3):
                recs[i].min_corner[d] = lo[d] + jitter[d]
                recs[i].max_corner[d] = lo[d] + ext[d] + jitter[3 + d]
                recs[i].dims[d] = 4
            recs[i].owner = int(i * 5 // n)
            recs[i].local_index = local.get(recs[i].owner, 0)
            local[recs[i].owner] = recs[i].local_index + 1
        eye = tuple(rng.uniform(-2, 3, size=3) * scale)
        if trial % 5 == 0:  # straight down an axis: |dir| components at the 1e-6 cut-off
            eye = (0.5 * scale, 0.5 * scale, 3.0 * scale)
        cam = scenes.make_camera(eye=eye, look_at=(0.5 * scale,) * 3, far=20.0 * scale)
        for graph in (True, False):
            want, ok = oracle.order(5, recs, n, cam, 1.0, graph)
            got, failed = engine.order_boxes_host(5, recs, n, cam, 1.0, graph)
            assert np.array_equal(got, want), (trial, graph)
            assert failed == (not ok)


def test_layer_order_and_hints_match_oracle(oracle):
    """computeBoxDepthHint + the (depth, owner, local index) sort, bit-exact."""
    for P, rule in ((1, "chunks"), (4, "round_robin"), (8, "chunks")):
        scene = scenes.scene_c4(box_cells=4, nranks=P, rule=rule)
        recs, n = scene.records()
        for eye in ((1.9, 1.3, 2.4), (0.5, 0.5, 3.0), (-1.0, 0.2, 0.4)):
            cam = scenes.make_camera(eye=eye)
            hints, order = engine.layer_order_host(recs, n, cam)
            want_hints, want_order = oracle.layer_order(recs, n, cam)
            assert np.array_equal(hints, want_hints)
            assert np.array_equal(order, want_order)


def test_piece_range_matches_oracle(oracle):
    for npix, P in ((10800, 8), (7, 3), (5, 8), (0, 2), (4096 * 4096, 7), (2048 * 2048, 8)):
        for i in range(P):
            assert engine.piece_range(npix, i, P) == oracle.piece_range(npix, i, P)
    lib = api.load_library()
    b, e = C.c_int(), C.c_int()
    assert lib.amrvr_piece_range(10, 3, 3, C.byref(b), C.byref(e)) == api.AMRVR_ERR_INVALID


def test_reference_sample_distance_matches_oracle(oracle):
    for scene in (scenes.scene_c1(), scenes.scene_c2prime(box_cells=10), scenes.scene_c4(box_cells=4)):
        recs, n = scene.records()
        got = engine.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
        assert got == oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    # no boxes: falls back to 1 % of the shortest bounds edge
    empty = (api.BoxRecord * 1)()
    assert engine.reference_sample_distance(empty, 0, (0, 0, 0), (1, 2, 3)) == \
        oracle.reference_sample_distance(empty, 0, (0, 0, 0), (1, 2, 3))


def test_null_arguments_are_refused():
    """NULL pointers, negative counts and NULL handles through the C ABI come
    back as AMRVR_ERR_INVALID with a message (the adapters rethrow it as the
    reference's std::invalid_argument); nothing is dereferenced.  In a child
    process, so that a crash shows up as a failure rather than as a dead suite."""
    import os
    import subprocess
    import sys

    worker = os.path.join(os.path.dirname(__file__), "abi_null_worker.py")
    out = subprocess.run([sys.executable, worker], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert out.stdout.split()[-2] == "ok"


def test_visibility_graph_dot_matches_the_reference(ref, tmp_path):
    """writeVisibilityGraph: the DOT documents of every round (cycle-breaking
    rounds included) against the files the reference's own
    BuildVisibilityOrderedGroup writes (VisibilityOrdering.cpp:319-352)."""
    cases = []
    scene = scenes.scene_c4(box_cells=4, nranks=4, rule="round_robin")
    recs, n = scene.records()
    for eye in ((1.9, 1.3, 2.4), (-1.2, 0.3, 2.0), (0.5, 0.5, 3.0)):
        cases.append((4, recs, n, scenes.make_camera(eye=eye), 1.25))
    rng = np.random.default_rng(23)
    for _ in range(30):  # random overlapping boxes: most need several rounds
        n = int(rng.integers(1, 40))
        recs = (api.BoxRecord * n)()
        for i in range(n):
            lo = rng.integers(0, 4, size=3) / 4.0
            hi = lo + rng.integers(1, 3, size=3) / 4.0
            for d in range(3):
                recs[i].min_corner[d], recs[i].max_corner[d] = lo[d], hi[d]
                recs[i].dims[d] = 4
            recs[i].owner = int(i * 3 // n)
        cases.append((3, recs, n, scenes.make_camera(eye=tuple(rng.uniform(-2, 3, size=3))), 1.0))
    rounds = []
    for k, (nranks, recs, n, cam, aspect) in enumerate(cases):
        work = tmp_path / f"case{k}"
        work.mkdir()
        want_order, want_docs, _log = ref.order_graphs(nranks, recs, n, cam, aspect, work)
        got_docs = engine.order_graph_dots_host(nranks, recs, n, cam, aspect)
        assert len(got_docs) == len(want_docs) >= 1, k
        for r, (got, want) in enumerate(zip(got_docs, want_docs)):
            assert got == want, (k, r)
        got_order, _failed = engine.order_boxes_host(nranks, recs, n, cam, aspect)
        assert np.array_equal(got_order, want_order)
        rounds.append(len(want_docs))
    assert max(rounds) > 1  # the multi-round path was exercised


def test_write_visibility_graphs_names_files_like_the_reference(tmp_path, capsys):
    scene = scenes.scene_c3(box_cells=4)
    recs, n = scene.records()
    cam = scenes.make_camera()
    start = engine._graph_file_counter
    names = engine.write_visibility_graphs(1, recs, n, cam, 1.0, directory=tmp_path)
    assert names == [f"visibility_graph_{start + k}.dot" for k in range(len(names))] and names
    out = capsys.readouterr().out
    assert out.splitlines() == [f"Wrote visibility graph to '{name}'" for name in names]
    text = (tmp_path / names[0]).read_text()
    assert text.startswith("digraph VisibilityGraph {\n  rankdir=LR;\n  box0 [label=\"box 0\\nrank 0\\nminDepth ")
    assert text.endswith("}\n")


def test_documents_name_only_entry_points_the_header_declares():
    """Every `amrvr_...` name in DESIGN.md / INTEGRATION.md / README.md is a
    symbol, type or family prefix of include/amrvr.h (no stale names in the docs)."""
    import pathlib
    import re

    root = pathlib.Path(__file__).resolve().parents[1]
    declared = set(re.findall(r"\b(amrvr_[a-z0-9_]+)\b", (root / "include" / "amrvr.h").read_text()))
    for doc in ("DESIGN.md", "INTEGRATION.md", "README.md"):
        for name in set(re.findall(r"`(amrvr_[a-z0-9_]+)", (root / doc).read_text())):
            if name.endswith("_"):  # a family, e.g. `amrvr_plotfile_*`
                assert any(d.startswith(name) for d in declared), (doc, name)
            else:
                assert name in declared, (doc, name)


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 8])
def test_stripe_plan_against_a_row_by_row_model(nranks):
    """The plan of a frame striped over the GPUs of a box (amrvr_stripe_plan_host, the
    arithmetic frame_fused and the fold kernel share): GPU q folds the 16-row strips
    q, q + P, ... -- every footprint row belongs to exactly one (queue, tile row), the
    march's store row equals the fold's read row, layers of one pool do not overlap."""
    lib = api.load_library()
    rng = np.random.default_rng(700 + nranks)
    height, nb = 333, 41
    rects = np.zeros((nb, 4), np.int32)
    for b in range(nb):
        h = int(rng.integers(1, 120))
        y0 = int(rng.integers(0, height - h + 1))
        w = 4 * int(rng.integers(1, 40))
        rects[b] = (4 * int(rng.integers(0, 50)), y0, w, h)
    rects[0] = (0, 16, 16, 16)   # exactly one strip
    rects[1] = (0, 15, 4, 2)     # two rows in two strips
    rects[2] = (0, 0, 8, height)  # the whole frame height
    n = nranks * nb
    tile_base, ry_first = np.zeros(n, np.int32), np.zeros(n, np.int32)
    layer_off, pool = np.zeros(n, np.uint64), np.zeros(nranks, np.uint64)
    qbegin = np.zeros(nranks + 1, np.int32)
    P32, P64 = C.POINTER(C.c_int32), C.POINTER(C.c_uint64)
    st = lib.amrvr_stripe_plan_host(nranks, nb, rects.ctypes.data_as(P32), tile_base.ctypes.data_as(P32),
                                    ry_first.ctypes.data_as(P32), layer_off.ctypes.data_as(P64),
                                    qbegin.ctypes.data_as(P32), pool.ctypes.data_as(P64))
    assert st == 0
    tiles, covered = 0, np.zeros((nb, height), np.int32)
    for q in range(nranks):
        assert qbegin[q] == tiles
        floats = 0
        for b in range(nb):
            _, y0, w, h = (int(v) for v in rects[b])
            ks = [k for k in range(y0 >> 4, ((y0 + h - 1) >> 4) + 1) if k % nranks == q]
            i = q * nb + b
            assert tile_base[i] == tiles and layer_off[i] == floats
            k = y0 >> 4
            while k % nranks != q:
                k += 1
            assert ry_first[i] == 16 * k
            for j, ks_j in enumerate(ks):  # tile row j of (queue, box), as the march walks it
                top = int(ry_first[i]) + j * 16 * nranks
                assert top == 16 * ks_j
                for y_in in range(16):
                    y = top + y_in
                    if y0 <= y < y0 + h:
                        covered[b, y] += 1
                        # the march stores this row at layer row 16 j + y_in; the fold looks it up here
                        assert lib.amrvr_stripe_layer_row_host(nranks, q, y0, y) == 16 * j + y_in
            for y in range(y0, y0 + h):  # rows of other ranks are not this rank's
                if (y >> 4) % nranks != q:
                    assert lib.amrvr_stripe_layer_row_host(nranks, q, y0, y) == -1
            tiles += len(ks) * ((w + 15) // 16)
            floats += len(ks) * 16 * w * 5
        assert pool[q] == floats
    assert qbegin[nranks] == tiles
    for b in range(nb):
        y0, h = int(rects[b, 1]), int(rects[b, 3])
        assert (covered[b, y0:y0 + h] == 1).all() and covered[b].sum() == h
    # the rows a rank folds partition the frame
    owner = np.array([(y >> 4) % nranks for y in range(height)])
    assert sorted(np.unique(owner)) == list(range(min(nranks, (height + 15) // 16)))
    assert lib.amrvr_stripe_plan_host(0, nb, rects.ctypes.data_as(P32), tile_base.ctypes.data_as(P32),
                                      ry_first.ctypes.data_as(P32), layer_off.ctypes.data_as(P64),
                                      qbegin.ctypes.data_as(P32), pool.ctypes.data_as(P64)) != 0
