"""Scene ingest either side of the hot path, CPU side:
the C restatement of detail::BuildSceneGeometry / ComputeSceneHistogram
(oracle/restate/orc_scene.c) against the reference's own SceneBuilder.cpp compiled
in oracle/_ref, bit for bit (doubles, floats and integer bin counts)."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes


def _poison(scene, seed=3):
    """Negative, zero, NaN and +-inf cells, so the log floor and the finite filter matter."""
    rng = np.random.default_rng(seed)
    for b in scene.boxes:
        flat = b.data.reshape(-1)
        idx = rng.choice(flat.size, size=max(flat.size // 50, 4), replace=False)
        flat[idx[0::4]] = -0.25
        flat[idx[1::4]] = 0.0
        flat[idx[2::4]] = np.nan
        flat[idx[3::4]] = np.inf
    return scene


CASES = {
    # name: (factory, kwargs, nranks, rule, build options)
    "c3_two_levels": (scenes.scene_c3, dict(box_cells=8), 2, "chunks", dict()),
    "c2prime_rescaled": (scenes.scene_c2prime, dict(box_cells=10), 3, "round_robin", dict()),
    "c4_level_window": (scenes.scene_c4, dict(box_cells=4), 4, "skewed", dict(min_level=1, max_level=2)),
    "c2_raw_range": (scenes.scene_c2, dict(box_cells=8), 1, "chunks", dict(normalize_to_data_range=False)),
    "c3_log_poisoned": (scenes.scene_c3, dict(box_cells=8), 2, "round_robin", dict(log_scale=True)),
    "c1": (scenes.scene_c1, dict(), 2, "chunks", dict()),
}


def _make(case):
    factory, kw, nranks, rule, opts = CASES[case]
    scene = factory(nranks=nranks, **kw) if factory is scenes.scene_c1 else factory(nranks=nranks, rule=rule, **kw)
    if "poisoned" in case:
        _poison(scene)
    return scene, opts


def _transform_fields(t):
    return [getattr(t, name) for name, _ in api.ScalarTransform._fields_]


def _owner_major(scene, boxes, n, opts):
    """Oracle boxes come in (level, fab) order over all ranks; the reference returns
    them rank by rank (MFIter order within a rank): stable sort by owner."""
    lo, hi = opts.get("min_level", 0), opts.get("max_level", len(scene.levels) - 1)
    gen = [b for b in sorted(scene.boxes, key=lambda b: b.gen_index) if lo <= b.level <= hi]
    assert len(gen) == n
    order = sorted(range(n), key=lambda i: gen[i].owner)
    return [boxes[i] for i in order], [gen[i].owner for i in order]


@pytest.mark.parametrize("case", sorted(CASES))
def test_scene_build_matches_reference(oracle, ref, case):
    scene, opts = _make(case)
    levels, nlevels, keep = scenes.c_levels(scene)
    boxes, summary, status = oracle.scene_build(levels, nlevels, **opts)
    assert status == 0
    built = ref.build_scene(scene, **opts)
    try:
        mine, owners = _owner_major(scene, boxes, summary.nboxes, opts)
        assert built["nboxes"] == summary.nboxes and built["owners"] == owners
        for i, b in enumerate(mine):
            r = built["boxes"][i]
            assert list(b.min_corner) == list(r.min_corner) and list(b.max_corner) == list(r.max_corner)
            assert list(b.dims) == list(r.dims) and list(b.lo) == list(r.lo)
            assert b.component == r.component
        assert _transform_fields(summary.transform) == _transform_fields(built["transform"])
        assert tuple(summary.scalar_range) == built["scalar_range"]
        assert tuple(summary.processed_range) == built["processed_range"]
        assert tuple(summary.original_range) == built["original_range"]
        assert list(summary.bounds_min) == built["bounds_min"]
        assert list(summary.bounds_max) == built["bounds_max"]
    finally:
        ref.free_scene(built)


@pytest.mark.parametrize("case,bins", [("c3_two_levels", 64), ("c2prime_rescaled", 7),
                                       ("c3_log_poisoned", 256), ("c2_raw_range", 1000), ("c1", 1)])
def test_histogram_matches_reference(oracle, ref, case, bins):
    scene, opts = _make(case)
    levels, nlevels, keep = scenes.c_levels(scene)
    boxes, summary, status = oracle.scene_build(levels, nlevels, **opts)
    assert status == 0
    built = ref.build_scene(scene, **opts)
    try:
        want, want_total = ref.scene_histogram(built, bins)
        got, got_total = oracle.scene_histogram(boxes, summary.nboxes, summary.transform,
                                                tuple(summary.scalar_range), bins)
        assert got_total == want_total == scene.total_cells()
        assert np.array_equal(got, want)
    finally:
        ref.free_scene(built)


def test_scene_build_reproduces_the_generator(oracle):
    """scenes._finish mirrors BuildSceneGeometry: corners and transform must agree."""
    scene = scenes.scene_c2prime(box_cells=10)
    levels, nlevels, keep = scenes.c_levels(scene)
    boxes, summary, status = oracle.scene_build(levels, nlevels)
    assert status == 0
    gen = sorted(scene.boxes, key=lambda b: b.gen_index)
    for i, b in enumerate(gen):
        assert np.array_equal(np.array(boxes[i].min_corner[:]), b.min_corner)
        assert np.array_equal(np.array(boxes[i].max_corner[:]), b.max_corner)
    assert _transform_fields(summary.transform) == _transform_fields(scene.transform)
    assert np.array_equal(np.array(summary.bounds_min[:]), scene.bounds_min)


def test_scene_build_errors(oracle):
    scene = scenes.scene_c2(box_cells=4)
    for b in scene.boxes:
        b.data[...] = -1.0
    levels, nlevels, keep = scenes.c_levels(scene)
    assert oracle.scene_build(levels, nlevels, log_scale=True)[2] == 3  # no positive value
    assert oracle.scene_build(levels, nlevels, min_level=5, max_level=6)[2] == 1  # no data
    for b in scene.boxes:
        b.data[...] = np.nan
    levels, nlevels, keep = scenes.c_levels(scene)
    assert oracle.scene_build(levels, nlevels)[2] == 2  # invalid scalar range


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof of every POD of include/amrvr.h as gcc lays it out == the ctypes mirror."""
    names = {"amrvr_box": api.Box, "amrvr_box_record": api.BoxRecord, "amrvr_camera": api.Camera,
             "amrvr_scalar_transform": api.ScalarTransform, "amrvr_frame_params": api.FrameParams,
             "amrvr_frame_stats": api.FrameStats, "amrvr_march_stats": api.MarchStats,
             "amrvr_colormap_pt": api.ColorMapPoint, "amrvr_fab": api.Fab, "amrvr_level": api.Level,
             "amrvr_scene_summary": api.SceneSummary}
    src = tmp_path / "sizes.c"
    body = "\n".join(f'  printf("{n} %zu\\n", sizeof({n}));' for n in names)
    src.write_text(f'#include <stdio.h>\n#include "{api.REPO_ROOT}/include/amrvr.h"\nint main(void) {{\n{body}\n  return 0;\n}}\n')
    import subprocess

    exe = tmp_path / "sizes"
    subprocess.run(["gcc", "-std=c11", "-o", str(exe), str(src)], check=True)
    out = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout
    for line in out.splitlines():
        name, size = line.split()
        assert C.sizeof(names[name]) == int(size), name
