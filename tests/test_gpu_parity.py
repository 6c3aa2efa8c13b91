"""GPU parity: libamrvr_b200.so (through its C ABI) against the CPU oracle on the
same seeded inputs.  Tolerances are BASELINE.json's: visibility order, pixel
ownership, brick bytes and sample counts bit-exact; float RGBA within 1e-4
absolute per channel; 8-bit output within 1 LSB."""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes

pytestmark = pytest.mark.gpu

RGBA_TOL = 1e-4  # BASELINE.json north_star: "float RGBA must agree within 1e-4 absolute per channel"


def _params(w, h, aa=1, transparency=0.0, graph=1, bbox=1):
    return api.FrameParams(width=w, height=h, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=graph, count_samples=1, draw_bounding_box=bbox)


def _c1_camera(oracle, scene):
    az, alt = scenes.auto_camera_draws()
    return oracle.auto_camera(scene.bounds_min, scene.bounds_max, az, alt)


def _assert_depth_equal(a, b):
    fin = np.isfinite(a)
    assert np.array_equal(fin, np.isfinite(b))
    assert np.abs(a[fin] - b[fin]).max(initial=0.0) <= RGBA_TOL


# ---------------------------------------------------------------- K5 / K2

def test_brick_quantise_bit_exact(ctx, oracle):
    scene = scenes.scene_c2(box_cells=16)
    # poison a few cells: the value path must sanitise them like the reference
    d = scene.boxes[3].data
    d[0, 0, 0], d[1, 2, 3], d[5, 5, 5] = np.nan, np.inf, -np.inf
    ds = ctx.upload_scene(scene)
    cboxes, n, _ = scene.c_boxes()
    for b in range(n):
        want = oracle.quantize_box(cboxes[b], scene.transform, *scene.scalar_range)
        got = ds.read_brick(b, want.size)
        assert np.array_equal(got, want), f"box {b}"
    ds.free()


def test_brick_quantise_ragged_and_strided(ctx, oracle):
    """nx not a multiple of 4, ghost cells around the valid region, component 1."""
    rng = np.random.default_rng(7)
    nx, ny, nz, g = 13, 7, 5, 2
    full = rng.random((2, nz + 2 * g, ny + 2 * g, nx + 2 * g))
    box = api.Box()
    for dd in range(3):
        box.min_corner[dd], box.max_corner[dd] = 0.0, 1.0
    box.dims[:] = [nx, ny, nz]
    box.lo[:] = [4, 8, 12]
    box.begin[:] = [4 - g, 8 - g, 12 - g]
    box.data = full.ctypes.data
    box.jstride = nx + 2 * g
    box.kstride = (nx + 2 * g) * (ny + 2 * g)
    box.nstride = box.kstride * (nz + 2 * g)
    box.component = 1
    arr = (api.Box * 1)(box)
    t, rng_ = scenes.data_range_transform(0.1, 0.8)
    ds = ctx.upload(arr, 1, t, (0.0, 0.9))  # range_max < 1 switches the soft clip on
    want = oracle.quantize_box(arr[0], t, 0.0, 0.9)
    got = ds.read_brick(0, want.size)
    assert np.array_equal(got, want)
    assert np.array_equal(want.reshape(nz, ny, nx)[0, 0],
                          [oracle.cell_index(v, t, 0.0, 0.9) for v in full[1, g, g, g:g + nx]])
    ds.free()


def test_brick_quantise_log_scale(ctx, oracle):
    scene = scenes.scene_c2(box_cells=8)
    vmin = min(float(b.data.min()) for b in scene.boxes)
    vmax = max(float(b.data.max()) for b in scene.boxes)
    t, _ = scenes.data_range_transform(vmin, vmax, log_scale=True, positive_floor=max(vmin, 1e-3))
    cboxes, n, _ = scene.c_boxes()
    ds = ctx.upload(cboxes, n, t, (0.0, 1.0))
    bad = total = 0
    for b in range(n):
        want = oracle.quantize_box(cboxes[b], t, 0.0, 1.0)
        got = ds.read_brick(b, want.size)
        bad += int((got != want).sum())
        total += want.size
        assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    # device log() and glibc log() may differ in the last ulp: a bin edge can flip, rarely
    assert bad <= max(1, total // 100000), (bad, total)
    ds.free()


def test_brick_stats(ctx):
    scene = scenes.scene_c2(box_cells=16)
    scene.boxes[1].data[0, 0, 0] = np.nan
    scene.boxes[2].data[1, 1, 1] = -3.5
    cboxes, n, _ = scene.c_boxes()
    lo, hi, lp, cnt = ctx.brick_stats(cboxes, n)
    allv = np.concatenate([b.data.ravel() for b in scene.boxes])
    fin = allv[np.isfinite(allv)]
    assert lo == fin.min() and hi == fin.max() and lp == fin[fin > 0].min() and cnt == fin.size


# ------------------------------------------------------------------- K1

def _check_layers(ctx, oracle, scene, camera, w, h, transparency=0.0, colormap=None):
    recs, n = scene.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    ds = ctx.upload_scene(scene)
    layers = ctx.raymarch(ds, camera, w, h, transparency, ref_d, colormap, count_samples=True)
    cboxes, nb, _ = scene.c_boxes()
    total = 0
    worst = 0.0
    for b in range(nb):
        want, s = oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, transparency, ref_d,
                               w, h, camera, colormap)
        got = layers.read(b)
        total += s
        worst = max(worst, float(np.abs(got[..., :4] - want[..., :4]).max()))
        assert np.abs(got[..., :4] - want[..., :4]).max() <= RGBA_TOL, f"box {b}"
        _assert_depth_equal(got[..., 4], want[..., 4])
        # the footprint rectangle must contain every pixel the reference touches
        x0, y0, rw, rh = layers.rect(b)
        hit = np.argwhere(want[..., 3] > 0)
        if hit.size:
            assert hit[:, 0].min() >= y0 and hit[:, 0].max() < y0 + rh
            assert hit[:, 1].min() >= x0 and hit[:, 1].max() < x0 + rw
    assert layers.stats().samples == total
    layers.free()
    ds.free()
    return worst


def test_layers_c1_auto_camera(ctx, oracle):
    scene = scenes.scene_c1()
    worst = _check_layers(ctx, oracle, scene, _c1_camera(oracle, scene), 128, 128)
    assert worst == 0.0  # power-of-two cells: identical arithmetic, identical bits


def test_layers_c2_small(ctx, oracle):
    scene = scenes.scene_c2(box_cells=16)
    _check_layers(ctx, oracle, scene, scenes.make_camera(), 160, 160, transparency=0.15)


def test_layers_non_power_of_two_cells(ctx, oracle):
    """C2': inexact step recurrence and true division in the cell index."""
    scene = scenes.scene_c2prime(box_cells=20)
    _check_layers(ctx, oracle, scene, scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45)),
                  160, 120)


def test_layers_user_colormap_and_inside_camera(ctx, oracle):
    """Eye inside the volume (full-frame footprints) and a CIELAB user map."""
    scene = scenes.scene_c2(box_cells=8)
    cmap = [(0.0, 0.1, 0.1, 0.6, 0.0), (0.3, 0.2, 0.8, 0.3, 0.05), (0.7, 0.9, 0.9, 0.1, 0.4),
            (1.0, 1.0, 0.2, 0.1, 0.9)]
    cam = scenes.make_camera(eye=(0.52, 0.47, 0.55), look_at=(0.9, 0.6, 0.1), fov_y=70.0)
    _check_layers(ctx, oracle, scene, cam, 96, 64, colormap=cmap)


def test_layers_width_not_multiple_of_four(ctx, oracle):
    """Footprints are padded to 4-column multiples; columns past the frame edge are
    not rays: the sample count must still equal the reference's (50 x 37 frame)."""
    scene = scenes.scene_c2(box_cells=8)
    _check_layers(ctx, oracle, scene, scenes.make_camera(), 50, 37)
    cam = scenes.make_camera(eye=(0.52, 0.47, 0.55), look_at=(0.9, 0.6, 0.1), fov_y=70.0)
    _check_layers(ctx, oracle, scene, cam, 50, 37)  # eye inside: full-frame footprints


# ------------------------------------------------------------- whole frame

def _check_frame(ctx, oracle, scene, camera, params, colormap=None):
    ds = ctx.upload_scene(scene)
    got = ctx.render_frame(ds, camera, params, colormap)
    want = oracle.render_frame(scene, camera, params, colormap)
    err = float(np.abs(got["final5"][..., :4] - want["final5"][..., :4]).max())
    assert err <= RGBA_TOL
    lsb = int(np.abs(got["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
    assert lsb <= 1  # "8-bit PNG output must agree within 1 LSB"
    assert got["stats"].samples == want["samples"]
    ds.free()
    return err, lsb


def test_frame_odd_size(ctx, oracle):
    """50 x 37: byte rows that are not whole words, fold tiles cut by the frame edge."""
    _check_frame(ctx, oracle, scenes.scene_c2(box_cells=8), scenes.make_camera(), _params(50, 37))
    _check_frame(ctx, oracle, scenes.scene_c4(box_cells=8), scenes.make_camera(), _params(37, 50))


def test_frame_bytes_only_matches_float_path(ctx, oracle):
    """The fold's own overlay + RGB8 epilogue (nobody asks for the float frame)
    gives the bytes of the float frame quantised afterwards."""
    scene = scenes.scene_c4(box_cells=8)
    cam = scenes.make_camera()
    for (w, h) in ((128, 96), (50, 37)):
        params = _params(w, h)
        ds = ctx.upload_scene(scene)
        both = ctx.render_frame(ds, cam, params)
        want = both["rgb8"].copy()
        only = ctx.render_frame(ds, cam, params, want_rgba=False)
        assert np.array_equal(only["rgb8"], want)
        lsb = int(np.abs(want.astype(int) - oracle.render_frame(scene, cam, params)["rgb8"].astype(int)).max())
        assert lsb <= 1
        ds.free()


def test_frame_c1(ctx, oracle):
    scene = scenes.scene_c1()
    _check_frame(ctx, oracle, scene, _c1_camera(oracle, scene), _params(128, 128))


def test_frame_c2_small(ctx, oracle):
    _check_frame(ctx, oracle, scenes.scene_c2(box_cells=16), scenes.make_camera(), _params(192, 192))


def test_frame_c2prime_aspect(ctx, oracle):
    _check_frame(ctx, oracle, scenes.scene_c2prime(box_cells=20),
                 scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45)), _params(160, 120))


def test_frame_c3_small_antialiased(ctx, oracle):
    _check_frame(ctx, oracle, scenes.scene_c3(box_cells=8), scenes.make_camera(),
                 _params(96, 96, aa=4, transparency=0.15))


def test_frame_c4_small(ctx, oracle):
    _check_frame(ctx, oracle, scenes.scene_c4(box_cells=8), scenes.make_camera(), _params(128, 128))


def test_frame_stats_carry_the_march_time_line(ctx):
    """amrvr_frame_stats: stage times from CUDA events and the march's own time line on the
    device clock (first CTA start -> last CTA end, CTA residency, table staging)."""
    scene = scenes.scene_c4(box_cells=8)
    ds = ctx.upload_scene(scene)
    st = ctx.render_frame(ds, scenes.make_camera(), _params(128, 128), want_rgba=False)["stats"]
    assert st.tiles > 0 and st.fragment_pixels > 0 and st.samples > 0
    assert st.march_ctas > 0 and st.march_span_ns > 0
    assert 0 < st.march_cta_busy_ns <= st.march_span_ns * st.march_ctas
    assert 0 <= st.march_table_ns <= st.march_cta_busy_ns
    assert st.march_ms > 0 and st.fold_ms > 0 and st.pre_ms >= 0 and st.composite_ms >= st.fold_ms
    assert st.march_span_ns * 1e-6 <= st.march_ms * 1.5 + 0.05  # the two clocks agree
    ds.free()


def test_frame_ragged_width(ctx, oracle):
    """Width not a multiple of 4 or 16: padded footprints overhang the frame."""
    _check_frame(ctx, oracle, scenes.scene_c2(box_cells=8), scenes.make_camera(), _params(101, 67))


@pytest.mark.parametrize("w,h", [(128, 96), (101, 67), (100, 50)])
def test_frame_bytes_only_path(ctx, oracle, w, h):
    """rgb8 without the float frame: on one rank the fold's own epilogue applies the
    overlay and writes the bytes (word stores when width % 4 == 0, byte stores
    otherwise; 100x50 ends in a 4-pixel tile and a 2-row band): same bytes as the
    float path and as the oracle."""
    scene = scenes.scene_c2(box_cells=8)
    camera = scenes.make_camera()
    params = _params(w, h)
    ds = ctx.upload_scene(scene)
    both = ctx.render_frame(ds, camera, params)
    want_bytes = both["rgb8"].copy()
    only = ctx.render_frame(ds, camera, params, want_rgba=False)
    assert only["final5"] is None
    assert np.array_equal(only["rgb8"], want_bytes)
    ref = oracle.render_frame(scene, camera, params)
    assert int(np.abs(only["rgb8"].astype(int) - ref["rgb8"].astype(int)).max()) <= 1
    ds.free()


def test_frame_of_a_scene_still_uploading(ctx):
    """amrvr_scene_upload_async: the first frame marches copy group by copy group while
    later groups are still on PCIe (46 MB of FABs = 3 groups here).  Same bricks, same
    samples and the same pixels, bit for bit, as after a finished upload -- with and
    without the float frame, and through the entry points that wait for the upload."""
    scene = scenes.scene_c4(box_cells=32)
    camera = scenes.make_camera()
    params = _params(320, 240)
    done = ctx.upload_scene(scene)
    want = ctx.render_frame(done, camera, params)
    want_rgb, want_f5, want_samples = want["rgb8"].copy(), want["final5"].copy(), want["stats"].samples
    ncells = int(np.prod(scene.boxes[-1].dims))
    want_brick = done.read_brick(len(scene.boxes) - 1, ncells)
    done.free()
    for want_rgba in (True, False):
        ds = ctx.upload_scene(scene, wait=False)
        got = ctx.render_frame(ds, camera, params, want_rgba=want_rgba)
        assert np.array_equal(got["rgb8"], want_rgb)
        if want_rgba:
            assert np.array_equal(got["final5"], want_f5)
        assert got["stats"].samples == want_samples
        again = ctx.render_frame(ds, camera, params, want_rgba=False)  # the upload is over by now
        assert np.array_equal(again["rgb8"], want_rgb)
        ds.free()
    ds = ctx.upload_scene(scene, wait=False)
    assert np.array_equal(ds.read_brick(len(scene.boxes) - 1, ncells), want_brick)  # waits
    ds.free()
    ds = ctx.upload_scene(scene, wait=False)
    ds.wait()
    ds.free()
    ctx.upload_scene(scene, wait=False).free()  # freed while the copies are in flight


def test_frame_bytes_only_antialiased_and_empty(ctx, oracle):
    """AA > 1 keeps the downsample -> fused overlay+bytes pass; an empty scene gives a black frame."""
    scene = scenes.scene_c2(box_cells=8)
    camera = scenes.make_camera()
    params = _params(64, 48, aa=4)
    ds = ctx.upload_scene(scene)
    only = ctx.render_frame(ds, camera, params, want_rgba=False)
    ref = oracle.render_frame(scene, camera, params)
    assert int(np.abs(only["rgb8"].astype(int) - ref["rgb8"].astype(int)).max()) <= 1
    ds.free()
    empty = ctx.upload((api.Box * 1)(), 0, scenes.identity_transform(), (0.0, 1.0))
    out = ctx.render_frame(empty, camera, _params(36, 20), want_rgba=False)
    assert out["final5"] is None and np.all(out["rgb8"] == 0)
    empty.free()


def test_frame_empty_scene(ctx):
    t = scenes.identity_transform()
    ds = ctx.upload((api.Box * 1)(), 0, t, (0.0, 1.0))
    out = ctx.render_frame(ds, scenes.make_camera(), _params(32, 16))
    assert np.all(out["final5"][..., :4] == 0) and np.all(np.isinf(out["final5"][..., 4]))
    assert np.all(out["rgb8"] == 0)
    ds.free()


def test_bad_arguments_are_reported(ctx):
    t = scenes.identity_transform()
    ds = ctx.upload((api.Box * 1)(), 0, t, (0.0, 1.0))
    with pytest.raises(api.AmrvrError) as e:
        ctx.render_frame(ds, scenes.make_camera(), _params(32, 16, aa=3))
    assert e.value.status == api.AMRVR_ERR_INVALID
    with pytest.raises(api.AmrvrError):
        ctx.raymarch(ds, scenes.make_camera(), 0, 16, 0.0, 0.1)
    ds.free()


# ------------------------------------------------------------ K7 ordering

@pytest.mark.parametrize("nranks,rule", [(2, "chunks"), (4, "round_robin"), (8, "chunks")])
def test_order_device_matches_oracle(ctx, oracle, nranks, rule):
    scene = scenes.scene_c4(box_cells=4, nranks=nranks, rule=rule)
    recs, n = scene.records()
    rng = np.random.default_rng(5)
    ctx.set_order_device_min_boxes(0 if nranks != 4 else 1024)  # device kernels; host back end for P=4
    for trial in range(12):
        eye = rng.normal(size=3)
        eye = 0.5 + eye / np.linalg.norm(eye) * rng.uniform(1.2, 3.0)
        cam = scenes.make_camera(eye=eye)
        want, ok = oracle.order(nranks, recs, n, cam, 1.25)
        got, failed = ctx.order_records(nranks, recs, n, cam, 1.25)
        assert np.array_equal(got, want), (trial, got, want)
        assert failed == (not ok)


def test_order_device_random_boxes_with_cycles(ctx, oracle):
    """Random touching boxes (cycles get split) -- device path == oracle."""
    rng = np.random.default_rng(11)
    ctx.set_order_device_min_boxes(0)
    for trial in range(20):
        n = int(rng.integers(2, 40))
        recs = (api.BoxRecord * n)()
        cuts = np.sort(rng.choice(np.arange(1, 16), size=3, replace=False)) / 16.0
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
        got, failed = ctx.order_records(4, recs, n, cam, 1.0)
        assert np.array_equal(got, want), (trial, got, want)
        assert failed == (not ok)


def test_order_device_many_boxes(ctx, oracle):
    """C5-sized table (16^3 = 4096 unit bricks, 8 owners): the default threshold
    sends this to the device radix sort + pair kernel."""
    n_axis = 16
    n = n_axis ** 3
    recs = (api.BoxRecord * n)()
    local = {}
    i = 0
    for k in range(n_axis):
        for j in range(n_axis):
            for ii in range(n_axis):
                r = recs[i]
                for d, c in enumerate((ii, j, k)):
                    r.min_corner[d], r.max_corner[d] = c / n_axis, (c + 1) / n_axis
                    r.dims[d] = 4
                r.owner = i % 8
                r.local_index = local.get(r.owner, 0)
                local[r.owner] = r.local_index + 1
                i += 1
    ctx.set_order_device_min_boxes(1024)
    launches = ctx.kernel_launches
    for eye in ((1.9, 1.3, 2.4), (-1.5, 0.2, 0.7)):
        cam = scenes.make_camera(eye=eye)
        want, ok = oracle.order(8, recs, n, cam, 1.0)
        got, failed = ctx.order_records(8, recs, n, cam, 1.0)
        assert np.array_equal(got, want) and failed == (not ok)
    assert ctx.kernel_launches >= launches + 4  # pair kernel + sort kernel per call


# --------------------------------------------- K6 fold with P-rank run structure

@pytest.mark.parametrize("nranks,rule", [(2, "chunks"), (3, "round_robin"), (8, "round_robin")])
def test_composite_virtual_ranks(ctx, oracle, nranks, rule):
    """The fold is not associative: P ranks give different pixels from 1 rank.
    Emulate P ranks on one GPU and compare with the oracle's P-rank compose."""
    base = scenes.scene_c3(box_cells=8)
    scene_p = base.with_owners(nranks, rule)  # owner-major order, as P ranks would gather
    camera = scenes.make_camera()
    w, h = 120, 90  # 10800 pixels: pieces do not end on row boundaries for P = 8
    recs, n = scene_p.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene_p.bounds_min, scene_p.bounds_max)
    # one GPU holds all boxes in the owner-major order
    single = scenes.Scene(scene_p.name, scene_p.boxes, scene_p.transform, scene_p.scalar_range,
                          scene_p.bounds_min, scene_p.bounds_max, 1, scene_p.levels)
    ds = ctx.upload_scene(single)
    layers = ctx.raymarch(ds, camera, w, h, 0.15, ref_d)
    owners = scene_p.owners()
    rank_order, _ = oracle.order(nranks, recs, n, camera, w / h)
    got, got_owner = ctx.composite_virtual(ds, layers, camera, nranks, owners, rank_order)
    dense = [layers.read(b) for b in range(n)]
    hints = [oracle.depth_hint(recs[i], camera) for i in range(n)]
    want, want_owner = oracle.compose(nranks, w, h, dense, owners, hints, rank_order)
    assert np.array_equal(got_owner, want_owner)  # pixel ownership: bit-exact
    assert np.abs(got[..., :4] - want[..., :4]).max() <= RGBA_TOL
    _assert_depth_equal(got[..., 4], want[..., 4])
    # the inputs are identical dense layers, so the fold itself must be bit-exact
    assert np.array_equal(got[..., :4], want[..., :4])
    layers.free()
    ds.free()


@pytest.mark.parametrize("n_axis,size", [(4, 512), (8, 256)])
def test_composite_c5_sweep(ctx, oracle, n_axis, size):
    """C5 (compositor sweep): n^3 unit bricks dealt round-robin to 8
    ranks; parity at reduced frame size with the same boxes (the reference's dense
    layers make 4096^2 infeasible for the oracle)."""
    nranks = 8
    scene_p = scenes.scene_c5(n_axis, nranks=nranks)
    camera = scenes.make_camera()
    w = h = size
    recs, n = scene_p.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene_p.bounds_min, scene_p.bounds_max)
    single = scenes.Scene(scene_p.name, scene_p.boxes, scene_p.transform, scene_p.scalar_range,
                          scene_p.bounds_min, scene_p.bounds_max, 1, scene_p.levels)
    ds = ctx.upload_scene(single)
    layers = ctx.raymarch(ds, camera, w, h, 0.0, ref_d)
    owners = scene_p.owners()
    rank_order, _ = oracle.order(nranks, recs, n, camera, 1.0)
    got, got_owner = ctx.composite_virtual(ds, layers, camera, nranks, owners, rank_order)
    dense = [layers.read(b) for b in range(n)]
    hints = [oracle.depth_hint(recs[i], camera) for i in range(n)]
    want, want_owner = oracle.compose(nranks, w, h, dense, owners, hints, rank_order)
    assert np.array_equal(got_owner, want_owner)
    assert np.array_equal(got[..., :4], want[..., :4])
    _assert_depth_equal(got[..., 4], want[..., 4])
    layers.free()
    ds.free()


def test_full_size_c4_layers_of_sampled_boxes(ctx, oracle):
    """BASELINE's full size (C4: 176 boxes of 64^3, 2048x2048): the layers of one
    box per AMR level against the oracle's paint at full resolution, the exact
    sample count of those boxes, and two size-independent properties of the
    whole frame: rendering twice is bit-identical, and the frame equals the
    1-virtual-rank composite of the stand-alone march."""
    scene = scenes.scene_c4()
    camera = scenes.make_camera()
    w = h = 2048
    recs, n = scene.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    ds = ctx.upload_scene(scene)
    layers = ctx.raymarch(ds, camera, w, h, 0.0, ref_d, count_samples=True)
    cboxes, nb, _ = scene.c_boxes()
    assert nb == 176
    for b in (0, 60, 150):  # level 0, 1, 2
        want, _s = oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.0, ref_d, w, h, camera,
                                None)
        got = layers.read(b)
        assert np.abs(got[..., :4] - want[..., :4]).max() <= RGBA_TOL, f"box {b}"
        _assert_depth_equal(got[..., 4], want[..., 4])
    total = layers.stats().samples
    full, owner = ctx.composite_virtual(ds, layers, camera, 1, np.zeros(n, np.int32), np.zeros(1, np.int32))
    assert np.all(owner == 0)
    params = _params(w, h)
    params.draw_bounding_box = 0
    a = ctx.render_frame(ds, camera, params)
    first = a["final5"].copy()
    assert a["stats"].samples == total
    b2 = ctx.render_frame(ds, camera, params)
    assert np.array_equal(first, b2["final5"])          # idempotent
    assert np.array_equal(first[..., :4], full[..., :4])  # frame == fold of the marched layers
    layers.free()
    ds.free()


def test_blend_known_answer(ctx, oracle):
    """ImageFullTest.cpp:253-255: over((.5,0,0,.5),(0,0,.5,.5)) = (.5,0,.25,.75)."""
    out = oracle.blend([0.5, 0, 0, 0.5, 1.0], [0, 0, 0.5, 0.5, 2.0])
    assert np.allclose(out[:4], [0.5, 0.0, 0.25, 0.75]) and out[4] == 1.0
