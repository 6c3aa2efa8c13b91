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
