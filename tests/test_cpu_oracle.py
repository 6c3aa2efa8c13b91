"""CPU suite (`-m "not gpu"`): pins the oracle.

1. The plain-C restatement (oracle/restate, class Oracle) against the
   REFERENCE'S OWN SOURCES compiled here (oracle/_ref, class Ref) -- skipped
   where /root/reference and the prebuilt _ref are both absent.
2. The restatement against golden fixtures the reference produced
   (tests/golden/*.npz, made by tests/golden/make_golden.py) -- runs anywhere.
3. The one known-answer value the reference's own tests hold for this path
   (Common/Testing/ImageFullTest.cpp:253-255).

Tolerances: integer / index / byte results bit-exact; floats bit-exact too,
because both sides are the same ISO C arithmetic without contraction.
"""
import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes
from pathlib import Path

GOLDEN = Path(__file__).resolve().parent / "golden"


def _params(w, h, aa=1, transparency=0.0, graph=1, bbox=1):
    return api.FrameParams(width=w, height=h, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=graph, count_samples=1, draw_bounding_box=bbox)


def _same_image(a, b):
    fin = np.isfinite(a[..., 4])
    assert np.array_equal(fin, np.isfinite(b[..., 4]))
    assert np.array_equal(a[..., :4], b[..., :4])
    assert np.array_equal(a[..., 4][fin], b[..., 4][fin])


# ------------------------------------------------- restatement vs the reference

@pytest.mark.parametrize("case", ["c1", "c2", "c2prime", "inside"])
def test_paint_matches_reference(oracle, ref, case):
    """VolumePainter::paint (Common/VolumePainter.cpp:548-961), every box."""
    cmap = None
    if case == "c1":
        scene, cam, w, h, tr = scenes.scene_c1(), scenes.make_camera(), 64, 48, 0.0
    elif case == "c2":
        scene, cam, w, h, tr = scenes.scene_c2(box_cells=8), scenes.make_camera(), 72, 72, 0.15
    elif case == "c2prime":
        scene = scenes.scene_c2prime(box_cells=10)
        cam, w, h, tr = scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45)), 80, 60, 0.0
    else:
        scene = scenes.scene_c2(box_cells=4)
        cam = scenes.make_camera(eye=(0.52, 0.47, 0.55), look_at=(0.9, 0.6, 0.1), fov_y=70.0)
        w, h, tr = 48, 32, 0.3
        cmap = [(0.0, 0.1, 0.1, 0.6, 0.0), (0.3, 0.2, 0.8, 0.3, 0.05), (0.7, 0.9, 0.9, 0.1, 0.4),
                (1.0, 1.0, 0.2, 0.1, 0.9)]
    recs, n = scene.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    cboxes, nb, _ = scene.c_boxes()
    for b in range(nb):
        want, ws = ref.paint(cboxes[b], scene.transform, *scene.scalar_range, tr, ref_d, w, h, cam, cmap)
        got, gs = oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, tr, ref_d, w, h, cam, cmap)
        assert gs == ws, f"box {b}: sample count"
        _same_image(got, want)


@pytest.mark.parametrize("nranks,rule", [(1, "chunks"), (2, "chunks"), (3, "round_robin"), (8, "chunks")])
def test_order_matches_reference(oracle, ref, nranks, rule):
    """BuildVisibilityOrderedGroup (Common/VisibilityOrdering.cpp:63-632)."""
    scene = scenes.scene_c4(box_cells=4, nranks=nranks, rule=rule)
    recs, n = scene.records()
    rng = np.random.default_rng(3)
    for _ in range(10):
        eye = rng.normal(size=3)
        eye = 0.5 + eye / np.linalg.norm(eye) * rng.uniform(1.2, 3.0)
        cam = scenes.make_camera(eye=eye)
        for graph in (True, False):
            want, _log = ref.order(nranks, recs, n, cam, 1.25, graph)
            got, _ok = oracle.order(nranks, recs, n, cam, 1.25, graph)
            assert np.array_equal(got, want)


def test_order_random_boxes_with_cycles_matches_reference(oracle, ref):
    rng = np.random.default_rng(11)
    for trial in range(30):
        n = int(rng.integers(1, 40))
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
        want, _ = ref.order(4, recs, n, cam, 1.0)
        got, _ = oracle.order(4, recs, n, cam, 1.0)
        assert np.array_equal(got, want), trial


@pytest.mark.parametrize("scale", [1.0, 37.5, 4096.0])
def test_order_at_the_touching_tolerance_matches_reference(oracle, ref, lib, scale):
    """Faces within, at and just beyond the 1e-5 * scale tolerance
    (VisibilityOrdering.cpp:211-228), flat boxes, axis-parallel views, large
    coordinates: restatement AND the product's host ordering against the
    reference's own BuildVisibilityOrderedGroup."""
    from amrvolumerenderer_b200 import engine

    rng = np.random.default_rng(int(scale) + 5)
    tol = 1e-5 * max(1.0, scale)
    for trial in range(40):
        n = int(rng.integers(2, 36))
        recs = (api.BoxRecord * n)()
        local = {}
        for i in range(n):
            lo = rng.integers(0, 3, size=3) * (scale / 3.0)
            ext = rng.integers(0 if trial % 7 == 0 else 1, 3, size=3) * (scale / 3.0)
            jitter = rng.choice([0.0, 0.4 * tol, 0.99 * tol, 1.01 * tol, 2.5 * tol, -0.99 * tol, -3 * tol], size=6)
            for d in range(3):
                recs[i].min_corner[d] = lo[d] + jitter[d]
                recs[i].max_corner[d] = lo[d] + ext[d] + jitter[3 + d]
                recs[i].dims[d] = 4
            recs[i].owner = int(i * 5 // n)
            recs[i].local_index = local.get(recs[i].owner, 0)
            local[recs[i].owner] = recs[i].local_index + 1
        eye = tuple(rng.uniform(-2, 3, size=3) * scale)
        if trial % 5 == 0:
            eye = (0.5 * scale, 0.5 * scale, 3.0 * scale)
        cam = scenes.make_camera(eye=eye, look_at=(0.5 * scale,) * 3, far=20.0 * scale)
        want, _ = ref.order(5, recs, n, cam, 1.0)
        got, _ = oracle.order(5, recs, n, cam, 1.0)
        assert np.array_equal(got, want), trial
        ours, _ = engine.order_boxes_host(5, recs, n, cam, 1.0)
        assert np.array_equal(ours, want), trial


@pytest.mark.parametrize("nranks,rule", [(1, "chunks"), (2, "chunks"), (3, "round_robin"), (5, "round_robin")])
def test_compose_matches_reference(oracle, ref, nranks, rule):
    """DirectSendBase::composeLayered (DirectSend/Base/DirectSendBase.cpp:316-458)
    over threads-as-ranks, incl. pieces that do not end on row boundaries."""
    scene = scenes.scene_c3(box_cells=4, nranks=nranks, rule=rule)
    cam = scenes.make_camera()
    w, h = 50, 37
    recs, n = scene.records()
    cboxes, n, _ = scene.c_boxes()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    dense = [oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15, ref_d, w, h, cam)[0]
             for b in range(n)]
    hints = np.array([oracle.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    order, _ = oracle.order(nranks, recs, n, cam, w / h)
    want, want_owner, _s = ref.compose(nranks, w, h, dense, scene.owners(), hints, order)
    got, got_owner = oracle.compose(nranks, w, h, dense, scene.owners(), hints, order)
    assert np.array_equal(got_owner, want_owner)
    _same_image(got, want)


def test_compose_empty_rank_and_reversed_order(oracle, ref):
    """A rank with no boxes still owns a piece; a non-identity rank order moves pieces."""
    scene = scenes.scene_c1(nranks=2)
    cam = scenes.make_camera()
    w, h = 33, 21
    recs, n = scene.records()
    cboxes, n, _ = scene.c_boxes()
    dense = [oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.0, 1 / 64, w, h, cam)[0]
             for b in range(n)]
    hints = np.array([oracle.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    owners = scene.owners()
    for nranks, order in ((3, [2, 0, 1]), (2, [1, 0]), (4, [3, 2, 1, 0])):
        want, want_owner, _ = ref.compose(nranks, w, h, dense, owners, hints, order)
        got, got_owner = oracle.compose(nranks, w, h, dense, owners, hints, order)
        assert np.array_equal(got_owner, want_owner)
        _same_image(got, want)


@pytest.mark.parametrize("case", ["c1_auto", "c2prime_p3", "c3_aa4_p2", "c4_p4_nobbox"])
def test_frame_matches_reference(oracle, ref, case, tmp_path):
    """VolumeRenderer::renderScene end to end (VolumeRenderer.cpp:947-1339) -> PPM bytes."""
    from oracle.oracle import read_ppm

    if case == "c1_auto":
        scene, p, cam = scenes.scene_c1(), _params(64, 64), None
    elif case == "c2prime_p3":
        scene = scenes.scene_c2prime(box_cells=10, nranks=3, rule="round_robin")
        p, cam = _params(80, 60), scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45))
    elif case == "c3_aa4_p2":
        scene, p, cam = scenes.scene_c3(box_cells=4, nranks=2), _params(48, 48, aa=4, transparency=0.15), scenes.make_camera()
    else:
        scene, p, cam = scenes.scene_c4(box_cells=4, nranks=4), _params(64, 64, bbox=1), scenes.make_camera()
    out = ref.render_scene(scene, cam, p, tmp_path / "f.ppm")
    want = read_ppm(tmp_path / "f.ppm")
    if cam is None:
        az, alt = scenes.auto_camera_draws()
        cam = oracle.auto_camera(scene.bounds_min, scene.bounds_max, az, alt)
    got = oracle.render_frame(scene, cam, p)
    assert got["samples"] == out["samples"]
    assert np.array_equal(got["rgb8"], want)


# ------------------------------------------------ restatement vs golden fixtures

def test_golden_c1_frame(oracle):
    g = np.load(GOLDEN / "c1_frame_64.npz")
    scene = scenes.scene_c1()
    az, alt = scenes.auto_camera_draws()
    cam = oracle.auto_camera(scene.bounds_min, scene.bounds_max, az, alt)
    got = oracle.render_frame(scene, cam, _params(64, 64))
    assert got["samples"] == int(g["samples"])
    assert np.array_equal(got["rgb8"], g["rgb8"])


def test_golden_c1_layers(oracle):
    g = np.load(GOLDEN / "c1_layers_48x40.npz")
    scene = scenes.scene_c1()
    cboxes, _n, _ = scene.c_boxes()
    for slot, b in enumerate(g["boxes"]):
        got, s = oracle.paint(cboxes[int(b)], scene.transform, *scene.scalar_range, 0.0,
                              float(g["ref_distance"]), 48, 40, scenes.make_camera())
        assert s == int(g["samples"][slot])
        _same_image(got, g["layers"][slot])


def test_golden_c2prime_frame(oracle):
    g = np.load(GOLDEN / "c2prime_frame_80x60_p3.npz")
    scene = scenes.scene_c2prime(box_cells=10, nranks=3, rule="round_robin")
    cam = scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45))
    got = oracle.render_frame(scene, cam, _params(80, 60))
    assert got["samples"] == int(g["samples"])
    assert np.array_equal(got["rgb8"], g["rgb8"])


def test_golden_c3_frame(oracle):
    g = np.load(GOLDEN / "c3_frame_48_aa4_p2.npz")
    got = oracle.render_frame(scenes.scene_c3(box_cells=4, nranks=2), scenes.make_camera(),
                              _params(48, 48, aa=4, transparency=0.15))
    assert got["samples"] == int(g["samples"])
    assert np.array_equal(got["rgb8"], g["rgb8"])


def test_golden_rank_order(oracle):
    g = np.load(GOLDEN / "c4_rank_order.npz")
    for P, rule in ((2, "chunks"), (4, "round_robin"), (8, "chunks")):
        scene = scenes.scene_c4(box_cells=4, nranks=P, rule=rule)
        recs, n = scene.records()
        for row, eye in zip(g[f"p{P}_{rule}"], g["cameras"]):
            got, _ = oracle.order(P, recs, n, scenes.make_camera(eye=eye), 1.25)
            assert np.array_equal(got, row)


def test_golden_compose(oracle):
    g = np.load(GOLDEN / "c3_compose_50x37_p3.npz")
    scene = scenes.scene_c3(box_cells=4, nranks=3, rule="round_robin")
    cam = scenes.make_camera()
    cboxes, n, _ = scene.c_boxes()
    dense = [oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15,
                          float(g["ref_distance"]), 50, 37, cam)[0] for b in range(n)]
    recs, n = scene.records()
    hints = np.array([oracle.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    assert np.array_equal(hints, g["hints"])
    order, _ = oracle.order(3, recs, n, cam, 50 / 37)
    assert np.array_equal(order, g["rank_order"])
    got, owner = oracle.compose(3, 50, 37, dense, scene.owners(), hints, order)
    assert np.array_equal(owner, g["piece_owner"])
    _same_image(got, g["full5"])


# -------------------------------------------------------- known answers / edges

def test_blend_known_answer(oracle):
    """ImageFullTest.cpp:253-255: over((.5,0,0,.5),(0,0,.5,.5)) = (.5,0,.25,.75)."""
    out = oracle.blend([0.5, 0, 0, 0.5, 1.0], [0, 0, 0.5, 0.5, 2.0])
    assert np.array_equal(out, np.array([0.5, 0.0, 0.25, 0.75, 1.0], np.float32))
    # depth decides who is in front, not argument order (ImageRGBAFloatColorDepthSort.hpp:13-27)
    swapped = oracle.blend([0, 0, 0.5, 0.5, 2.0], [0.5, 0, 0, 0.5, 1.0])
    assert np.array_equal(swapped, out)
    # a transparent pixel at +inf is the identity on either side
    clear = [0, 0, 0, 0, np.inf]
    x = [0.3, 0.2, 0.1, 0.35, 1.7]
    assert np.array_equal(oracle.blend(x, clear), np.array(x, np.float32))
    assert np.array_equal(oracle.blend(clear, x), np.array(x, np.float32))


def test_piece_range_covers_frame(oracle):
    """getPieceRange (DirectSendBase.cpp:59-74): N/P each, last takes the remainder."""
    for npix, P in ((10800, 8), (7, 3), (5, 8), (0, 2), (4096 * 4096, 7)):
        spans = [oracle.piece_range(npix, i, P) for i in range(P)]
        assert spans[0][0] == 0 and spans[-1][1] == npix
        for (b0, e0), (b1, _e1) in zip(spans, spans[1:]):
            assert e0 == b1 and e0 - b0 == npix // P


def test_quantise_byte_rule(oracle):
    """Color.hpp:86-90: byte = clamp(int(c*256), 0, 255); rows written top = y_max."""
    img = np.zeros((2, 3, 5), np.float32)
    img[0, :, 0] = [0.0, 0.00390625, 0.999]
    img[1, :, 0] = [1.0, 1.5, -0.2]
    out = oracle.quantize_rgb8(img)
    assert out[1, :, 0].tolist() == [0, 1, 255]      # image row 0 is the LAST output row
    assert out[0, :, 0].tolist() == [255, 255, 0]


def test_cell_index_edges(oracle):
    t, _ = scenes.data_range_transform(0.0, 2.0)
    idx = [oracle.cell_index(v, t, 0.0, 1.0) for v in (np.nan, np.inf, -np.inf, -1.0, 0.0, 1.0, 2.0, 5.0)]
    assert idx == [0, 0, 0, 0, 0, 127, 255, 255]


def test_downsample_sum_order(oracle):
    rng = np.random.default_rng(0)
    src = rng.random((8, 12, 5)).astype(np.float32)
    out = oracle.downsample(src, 6, 4, 2)
    want = np.zeros((4, 6, 4), np.float32)
    for y in range(4):
        for x in range(6):
            acc = np.zeros(4, np.float32)
            for dy in range(2):
                for dx in range(2):
                    acc = acc + src[2 * y + dy, 2 * x + dx, :4]
            want[y, x] = acc * np.float32(0.25)
    assert np.array_equal(out[..., :4], want) and np.all(np.isinf(out[..., 4]))
