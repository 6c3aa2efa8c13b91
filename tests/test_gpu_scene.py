"""GPU parity for the scene ingest either side of the hot path (SURVEY.md 8f rows
f1 and f4): amrvr_scene_build (= detail::BuildSceneGeometry, SceneBuilder.cpp:
113-425) and amrvr_scene_histogram (= detail::ComputeSceneHistogram, :445-577)
through the C ABI against the oracle restatement, which tests/test_cpu_scene.py
pins bit for bit to the reference's own SceneBuilder.cpp.  Everything here is
bit-exact: doubles, floats and integer bin counts."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes
from test_cpu_scene import CASES, _make, _transform_fields

pytestmark = pytest.mark.gpu


def _assert_same_scene(boxes, summary, want_boxes, want):
    assert summary.nboxes == want.nboxes
    for i in range(want.nboxes):
        b, r = boxes[i], want_boxes[i]
        assert list(b.min_corner) == list(r.min_corner) and list(b.max_corner) == list(r.max_corner)
        assert list(b.dims) == list(r.dims) and list(b.lo) == list(r.lo) and list(b.begin) == list(r.begin)
        assert (b.jstride, b.kstride, b.nstride, b.component) == (r.jstride, r.kstride, r.nstride, r.component)
        assert b.data == r.data
    assert _transform_fields(summary.transform) == _transform_fields(want.transform)
    for name in ("scalar_range", "processed_range", "original_range", "bounds_min", "bounds_max"):
        assert list(getattr(summary, name)) == list(getattr(want, name)), name
    assert summary.has_original_range == want.has_original_range
    assert summary.global_scale == want.global_scale


@pytest.mark.parametrize("case", sorted(CASES))
def test_scene_build_bit_exact(ctx, oracle, case):
    scene, opts = _make(case)
    levels, nlevels, keep = scenes.c_levels(scene)
    want_boxes, want, status = oracle.scene_build(levels, nlevels, **opts)
    assert status == 0
    boxes, summary = ctx.scene_build(levels, nlevels, **opts)
    _assert_same_scene(boxes, summary, want_boxes, want)


def test_scene_build_ghost_cells_and_components(ctx, oracle):
    """FABs with 2 ghost cells (1e30 outside the valid box) and the field in component 1 of 3."""
    scene = scenes.scene_c3(box_cells=6)
    levels, nlevels, keep = scenes.c_levels(scene, ghost=2, ncomp=3, component=1)
    want_boxes, want, status = oracle.scene_build(levels, nlevels, component=1)
    assert status == 0
    boxes, summary = ctx.scene_build(levels, nlevels, component=1)
    _assert_same_scene(boxes, summary, want_boxes, want)
    assert summary.original_range[1] < 1e29  # no ghost cell was looked at
    # and the boxes it returns upload and quantise like the oracle's
    ds = ctx.upload(boxes, summary.nboxes, summary.transform, tuple(summary.scalar_range))
    for b in (0, summary.nboxes - 1):
        w = oracle.quantize_box(want_boxes[b], want.transform, *want.scalar_range)
        assert np.array_equal(ds.read_brick(b, w.size), w)
    got, total = ctx.scene_histogram(boxes, summary.nboxes, summary.transform, tuple(summary.scalar_range), 100)
    w, wt = oracle.scene_histogram(want_boxes, want.nboxes, want.transform, tuple(want.scalar_range), 100)
    assert total == wt == scene.total_cells() and np.array_equal(got, w)
    ds.free()


def test_upload_reuses_what_scene_build_staged(ctx, oracle):
    """amrvr_scene_upload_staged (the in-situ sequence of VolumeRendererApi.cpp:80-124 + 257-395):
    the statistics pass of amrvr_scene_build left the host FABs in the staging buffer, the
    upload quantises from there -- same bytes as a plain upload, with ghost cells and a
    component offset; any other list of views falls back to the copying upload."""
    scene = scenes.scene_c3(box_cells=6)
    levels, nlevels, keep = scenes.c_levels(scene, ghost=2, ncomp=3, component=1)
    want_boxes, want, status = oracle.scene_build(levels, nlevels, component=1)
    assert status == 0
    boxes, summary = ctx.scene_build(levels, nlevels, component=1)
    rng = tuple(summary.scalar_range)
    staged = ctx.upload(boxes, summary.nboxes, summary.transform, rng, staged=True)
    assert staged.was_staged
    again = ctx.upload(boxes, summary.nboxes, summary.transform, rng, staged=True)  # the buffer still holds them
    assert again.was_staged
    plain = ctx.upload(boxes, summary.nboxes, summary.transform, rng)
    assert not plain.was_staged
    for b in range(summary.nboxes):
        w = oracle.quantize_box(want_boxes[b], want.transform, *want.scalar_range)
        assert np.array_equal(staged.read_brick(b, w.size), w)
        assert np.array_equal(again.read_brick(b, w.size), w)
        assert np.array_equal(plain.read_brick(b, w.size), w)
    # a plain upload used the staging buffer for itself: the next staged call must copy
    after_plain = ctx.upload(boxes, summary.nboxes, summary.transform, rng, staged=True)
    assert not after_plain.was_staged
    # other views (one box fewer) than the ones staged: copies as well, and is right
    boxes2, summary2 = ctx.scene_build(levels, nlevels, component=1)
    fewer = ctx.upload(boxes2, summary2.nboxes - 1, summary2.transform, rng, staged=True)
    assert not fewer.was_staged
    w = oracle.quantize_box(want_boxes[0], want.transform, *want.scalar_range)
    assert np.array_equal(fewer.read_brick(0, w.size), w)
    for ds in (staged, again, plain, after_plain, fewer):
        ds.free()


@pytest.mark.parametrize("case,bins", [("c3_two_levels", 64), ("c2prime_rescaled", 7), ("c3_log_poisoned", 256),
                                       ("c2_raw_range", 1000), ("c1", 1), ("c4_level_window", 20000)])
def test_histogram_bit_exact(ctx, oracle, case, bins):
    """bins <= 8192 take the shared-memory path, 20000 the global-atomic one."""
    scene, opts = _make(case)
    levels, nlevels, keep = scenes.c_levels(scene)
    boxes, summary = ctx.scene_build(levels, nlevels, **opts)
    got, total = ctx.scene_histogram(boxes, summary.nboxes, summary.transform, tuple(summary.scalar_range), bins)
    want, want_total = oracle.scene_histogram(boxes, summary.nboxes, summary.transform,
                                              tuple(summary.scalar_range), bins)
    assert total == want_total == int(got.sum())
    assert np.array_equal(got, want)


def test_histogram_device_resident_full_size(ctx, oracle):
    """32^3 bricks already in HBM (in-situ): counts add up and match the oracle on a sub-sample."""
    import torch

    scene = scenes.scene_c2(box_cells=32)
    levels, nlevels, keep = scenes.c_levels(scene)
    hb, hs = ctx.scene_build(levels, nlevels)
    dev = [torch.from_numpy(np.ascontiguousarray(b.data)).cuda() for b in sorted(scene.boxes, key=lambda b: b.gen_index)]
    for i, t in enumerate(dev):
        hb[i].data = t.data_ptr()
    got, total = ctx.scene_histogram(hb, hs.nboxes, hs.transform, tuple(hs.scalar_range), 512)
    for i, b in enumerate(sorted(scene.boxes, key=lambda b: b.gen_index)):
        hb[i].data = b.data.ctypes.data
    want, want_total = oracle.scene_histogram(hb, hs.nboxes, hs.transform, tuple(hs.scalar_range), 512)
    assert total == want_total == scene.total_cells()
    assert np.array_equal(got, want)


def test_scene_build_errors(ctx):
    scene = scenes.scene_c2(box_cells=4)
    for b in scene.boxes:
        b.data[...] = -1.0
    levels, nlevels, keep = scenes.c_levels(scene)
    with pytest.raises(RuntimeError, match="no positive scalar values"):
        ctx.scene_build(levels, nlevels, log_scale=True)
    with pytest.raises(RuntimeError, match="no data"):
        ctx.scene_build(levels, nlevels, min_level=5, max_level=6)
    for b in scene.boxes:
        b.data[...] = np.nan
    levels, nlevels, keep = scenes.c_levels(scene)
    with pytest.raises(RuntimeError, match="invalid scalar range"):
        ctx.scene_build(levels, nlevels)
    with pytest.raises(RuntimeError, match="binCount"):
        ctx.scene_histogram((api.Box * 1)(), 0, api.ScalarTransform(), (0.0, 1.0), 0)
