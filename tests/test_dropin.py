"""Drop-in check (`-m gpu`): the reference's OWN driver code (VolumeRenderer::
renderScene, SceneBuilder, the image classes -- compiled from /root/reference
into oracle/_ref/libamrvr_dropin.so) with its three hot-path translation units
replaced by amrvolumerenderer_b200/adapters/*.cpp, which keep the reference's
signatures and forward to libamrvr_b200.so.  Compared with the all-reference
build / the oracle on the same inputs."""
import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, scenes

RGBA_TOL = 1e-4


@pytest.fixture(scope="module")
def dropin():
    from oracle.oracle import DropIn, RefUnavailable

    try:
        return DropIn()
    except (RefUnavailable, OSError) as exc:
        pytest.skip(f"drop-in build unavailable: {exc}")


def _params(w, h, aa=1, transparency=0.0, bbox=1):
    return api.FrameParams(width=w, height=h, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=1, count_samples=0, draw_bounding_box=bbox)


def test_dropin_fails_loudly_without_gpu(dropin):
    """No CPU fallback behind the reference's entry points either."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    scene = scenes.scene_c1()
    cboxes, _n, _ = scene.c_boxes()
    with pytest.raises(api.AmrvrError) as e:
        dropin.paint(cboxes[0], scene.transform, *scene.scalar_range, 0.0, 1 / 64, 16, 16,
                     scenes.make_camera())
    assert "no CUDA device" in str(e.value) or "CPU fallback" in str(e.value)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c1", "c2prime"])
def test_dropin_paint(dropin, oracle, case):
    """VolumePainter::paint through the adapter == the oracle's paint."""
    if case == "c1":
        scene, cam, w, h = scenes.scene_c1(), scenes.make_camera(), 96, 72
    else:
        scene = scenes.scene_c2prime(box_cells=10)
        cam, w, h = scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45)), 80, 60
    recs, n = scene.records()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    cboxes, nb, _ = scene.c_boxes()
    for b in range(min(nb, 6)):
        want, _s = oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15, ref_d, w, h, cam)
        got, _ = dropin.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15, ref_d, w, h, cam)
        assert np.abs(got[..., :4] - want[..., :4]).max() <= RGBA_TOL
        fin = np.isfinite(want[..., 4])
        assert np.array_equal(fin, np.isfinite(got[..., 4]))
        assert np.abs(got[..., 4][fin] - want[..., 4][fin]).max(initial=0) <= RGBA_TOL


@pytest.mark.gpu
def test_dropin_order(dropin, oracle):
    """BuildVisibilityOrderedGroup through the adapter (threads as MPI ranks)."""
    scene = scenes.scene_c4(box_cells=4, nranks=4, rule="round_robin")
    recs, n = scene.records()
    for eye in ((1.9, 1.3, 2.4), (-1.2, 0.3, 2.0), (0.4, 2.8, -0.9)):
        cam = scenes.make_camera(eye=eye)
        want, _ok = oracle.order(4, recs, n, cam, 1.25)
        got, _log = dropin.order(4, recs, n, cam, 1.25)
        assert np.array_equal(got, want)


@pytest.mark.gpu
def test_dropin_compose_single_rank(dropin, oracle):
    """Compositor::compose through the adapter, one rank (one GPU)."""
    scene = scenes.scene_c3(box_cells=4)
    cam = scenes.make_camera()
    w, h = 50, 37
    recs, n = scene.records()
    cboxes, n, _ = scene.c_boxes()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    dense = [oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15, ref_d, w, h, cam)[0]
             for b in range(n)]
    hints = np.array([oracle.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    want, want_owner = oracle.compose(1, w, h, dense, scene.owners(), hints, [0])
    got, got_owner, _s = dropin.compose(1, w, h, dense, scene.owners(), hints, [0])
    assert np.array_equal(got_owner, want_owner)
    assert np.array_equal(got[..., :4], want[..., :4])  # same dense inputs -> same fold, bit for bit


@pytest.mark.gpu
def test_dropin_compose_two_ranks(dropin, oracle):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (one per thread-rank)")
    scene = scenes.scene_c3(box_cells=4, nranks=2, rule="round_robin")
    cam = scenes.make_camera()
    w, h = 50, 37
    recs, n = scene.records()
    cboxes, n, _ = scene.c_boxes()
    ref_d = oracle.reference_sample_distance(recs, n, scene.bounds_min, scene.bounds_max)
    dense = [oracle.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.15, ref_d, w, h, cam)[0]
             for b in range(n)]
    hints = np.array([oracle.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    for order in ([0, 1], [1, 0]):
        want, want_owner = oracle.compose(2, w, h, dense, scene.owners(), hints, order)
        got, got_owner, _s = dropin.compose(2, w, h, dense, scene.owners(), hints, order)
        assert np.array_equal(got_owner, want_owner)
        assert np.array_equal(got[..., :4], want[..., :4])


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["c1_auto", "c3_aa4", "c4"])
def test_dropin_render_scene(dropin, ref, case, tmp_path):
    """The reference's renderScene, end to end to PPM bytes, with the B200 hot
    path swapped in == the all-reference build (within 1 LSB)."""
    from oracle.oracle import read_ppm

    if case == "c1_auto":
        scene, p, cam = scenes.scene_c1(), _params(64, 64), None
    elif case == "c3_aa4":
        scene, p, cam = scenes.scene_c3(box_cells=4), _params(48, 48, aa=4, transparency=0.15), scenes.make_camera()
    else:
        scene, p, cam = scenes.scene_c4(box_cells=4), _params(64, 64), scenes.make_camera()
    ref.render_scene(scene, cam, p, tmp_path / "ref.ppm")
    dropin.render_scene(scene, cam, p, tmp_path / "b200.ppm")
    want, got = read_ppm(tmp_path / "ref.ppm"), read_ppm(tmp_path / "b200.ppm")
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1


# ---------------------------------------------------------------- in-situ seam
# api::Render / api::ComputeHistogram (VolumeRendererApi.hpp:22-56): in the drop-in
# library they are adapters/VolumeRendererApiB200.cpp (views of the MultiFabs ->
# amrvr_convexify_levels -> amrvr_scene_build -> amrvr_scene_upload ->
# amrvr_render_frame); in the all-reference library the reference's own file.

def _insitu_params(w=512, h=512, aa=1, transparency=0.0):
    return api.FrameParams(width=w, height=h, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=1, count_samples=0, draw_bounding_box=1)


@pytest.mark.gpu
def test_dropin_api_render_config1(dropin, ref, tmp_path):
    """BASELINE config 1 exactly as Examples/RenderFromMultiFab.cpp:20-58: 32^3 cells in
    16^3 boxes, field x^2+y^2+z^2, 512x512, automatic camera -> image within 1 LSB of
    the all-reference build; the status lines on stdout are the reference's."""
    from oracle.oracle import read_ppm

    scene = scenes.scene_c1()
    want_log = ref.api_render(scene, _insitu_params(), tmp_path / "multifab-render-ref.ppm")["log"]
    got_log = dropin.api_render(scene, _insitu_params(), tmp_path / "multifab-render.ppm")["log"]
    want, got = read_ppm(tmp_path / "multifab-render-ref.ppm"), read_ppm(tmp_path / "multifab-render.ppm")
    assert want.shape == got.shape == (512, 512, 3)
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
    assert want.max() > 0  # the frame is not empty
    for needle in ("Loaded AMReX data component 0 across 1 level(s); normalized scalar range [0, 1]",
                   "Render: composed 262144 pixels on rank 0", "Saved volume composited image to"):
        assert needle in want_log and needle in got_log, (needle, got_log)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["camera_colormap_range", "log_up_aa", "components_transparency"])
def test_dropin_api_render_options(dropin, ref, case, tmp_path):
    """RenderOptions beyond the defaults: explicit camera + colour map + scalar-range
    override; log scale + custom up vector + antialiasing; a component of a 3-component
    MultiFab + box transparency."""
    from oracle.oracle import read_ppm

    scene = scenes.scene_c2(box_cells=8)
    kw = {}
    p = _insitu_params(96, 80)
    if case == "camera_colormap_range":
        kw = dict(camera=scenes.make_camera(eye=(2.2, 1.1, 1.7)), scalar_range=(0.05, 0.6),
                  colormap=[(0.05, 0.1, 0.1, 0.6, 0.0), (0.2, 0.2, 0.8, 0.3, 0.1), (0.6, 1.0, 0.2, 0.1, 0.9)])
    elif case == "log_up_aa":
        kw = dict(log_scale=True, up_vector=(0.0, 0.0, 2.0))
        p = _insitu_params(64, 48, aa=4)
    else:
        kw = dict(component=1, ncomp=3)
        p = _insitu_params(80, 96, transparency=0.3)
    ref.api_render(scene, p, tmp_path / "ref.ppm", **kw)
    dropin.api_render(scene, p, tmp_path / "b200.ppm", **kw)
    want, got = read_ppm(tmp_path / "ref.ppm"), read_ppm(tmp_path / "b200.ppm")
    assert want.max() > 0
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1


@pytest.mark.gpu
def test_dropin_api_render_errors(dropin, ref, tmp_path):
    """Same exception class (std::invalid_argument / std::runtime_error) and text as
    VolumeRendererApi.cpp:257-276 and its validators."""
    scene = scenes.scene_c1()
    p = _insitu_params(32, 32)
    bad = [
        dict(scalar_range=(0.7, 0.2)),                                   # validateScalarRange
        dict(colormap=[(0.5, 0.1, 0.1, 0.1, 0.1)]),                      # fewer than two control points
        dict(colormap=[(0.0, 0.1, 0.1, 0.1, 0.1), (0.0, 0.2, 0.2, 0.2, 0.2)]),  # not increasing
        dict(up_vector=(0.0, 0.0, 0.0)),                                 # validateUpVector
        dict(camera=scenes.make_camera(eye=(0.5, 0.5, 0.5), look_at=(0.5, 0.5, 0.5))),  # eye == look-at
        dict(camera=scenes.make_camera(fov_y=190.0)),
        dict(component=3),                                               # component out of bounds
        dict(log_scale=True, scalar_range=(-1.0, 2.0)),                  # non-positive under log scale
    ]
    for kw in bad:
        errs = []
        for eng in (ref, dropin):
            with pytest.raises(api.AmrvrError) as e:
                eng.api_render(scene, p, tmp_path / "never.ppm", **kw)
            errs.append((e.value.status, e.value.message))
        assert errs[0] == errs[1], (kw, errs)
        assert errs[0][0] == api.AMRVR_ERR_INVALID, (kw, errs)
    # renderScene's own checks (VolumeRenderer.cpp:562-579), reached after the scene is loaded
    for w, h, aa, tr in ((0, 32, 1, 0.0), (32, 32, 3, 0.0), (32, 32, 1, 1.5)):
        errs = []
        for eng in (ref, dropin):
            with pytest.raises(api.AmrvrError) as e:
                eng.api_render(scene, _insitu_params(w, h, aa, tr), tmp_path / "never.ppm")
            errs.append((e.value.status, e.value.message))
        assert errs[0] == errs[1] and errs[0][0] == api.AMRVR_ERR_INVALID, errs
    # empty output name: std::invalid_argument before anything else
    for eng in (ref, dropin):
        with pytest.raises(api.AmrvrError) as e:
            eng.api_render(scene, p, "")
        assert e.value.status == api.AMRVR_ERR_INVALID and "output filename must not be empty" in e.value.message


@pytest.mark.gpu
@pytest.mark.parametrize("log_scale", [False, True])
def test_dropin_api_histogram(dropin, ref, log_scale):
    """api::ComputeHistogram (VolumeRendererApi.cpp:397-415): counts, sample count and
    the three ranges equal the reference's, bit for bit."""
    scene = scenes.scene_c2(box_cells=8)
    for bins in (256, 37):
        want = ref.api_histogram(scene, bins, log_scale=log_scale)
        got = dropin.api_histogram(scene, bins, log_scale=log_scale)
        assert np.array_equal(got["counts"], want["counts"])
        assert got["samples"] == want["samples"] == 32 ** 3
        for k in ("normalized_range", "processed_range", "original_range", "has_processed", "has_original"):
            assert got[k] == want[k], k
    for eng in (ref, dropin):
        with pytest.raises(api.AmrvrError) as e:
            eng.api_histogram(scene, 0)
        assert e.value.status == api.AMRVR_ERR_INVALID and "binCount must be positive" in e.value.message


@pytest.mark.gpu
def test_dropin_api_render_two_ranks(dropin, ref, tmp_path):
    """api::Render collectively on two thread-ranks (one GPU each): NCCL context from
    MPI_COMM_WORLD, bricks replicated, frame sharded, image on rank 0."""
    import torch
    from oracle.oracle import read_ppm

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (one per thread-rank)")
    scene = scenes.scene_c2(box_cells=8, nranks=2, rule="round_robin")
    p = _insitu_params(96, 80)
    ref.api_render(scene, p, tmp_path / "ref.ppm")
    dropin.api_render(scene, p, tmp_path / "b200.ppm")
    want, got = read_ppm(tmp_path / "ref.ppm"), read_ppm(tmp_path / "b200.ppm")
    assert want.max() > 0
    assert np.abs(got.astype(int) - want.astype(int)).max() <= 1
