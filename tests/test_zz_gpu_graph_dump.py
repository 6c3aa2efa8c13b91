"""render(write_visibility_graph=True) end to end on the GPU: rank 0 leaves
visibility_graph_<n>.dot in the working directory, byte-identical to what the
reference's BuildVisibilityOrderedGroup writes for the same box table
(VisibilityOrdering.cpp:319-352).  The text path itself is pinned on the CPU
(tests/test_cpu_host_logic.py); this covers the wiring through DeviceScene."""
import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, engine
from amrvolumerenderer_b200.plotfile import synthetic_hierarchy, write_plotfile
from amrvolumerenderer_b200.render import render

pytestmark = pytest.mark.gpu


def test_render_writes_the_visibility_graph(ctx, ref, tmp_path, monkeypatch, capsys):
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=25)
    path = write_plotfile(tmp_path / "plt", levels, ("density",), plo, phi)
    work = tmp_path / "cwd"
    work.mkdir()
    monkeypatch.chdir(work)
    first = engine._graph_file_counter
    out = render(str(path), width=96, height=64, output=str(tmp_path / "g.ppm"),
                 camera_eye=(1.9, 1.3, 2.4), camera_look_at=(0.5, 0.5, 0.5), camera_far=10.0,
                 write_visibility_graph=True, context=ctx, return_frame=True)
    names = sorted(p.name for p in work.iterdir())
    assert names and names[0] == f"visibility_graph_{first}.dot"
    assert f"Wrote visibility graph to '{names[0]}'" in capsys.readouterr().out

    n = len(out["boxes"])
    recs = (api.BoxRecord * n)()
    for i, (bmin, bmax, dims, _lo) in enumerate(out["boxes"]):
        for d in range(3):
            recs[i].min_corner[d], recs[i].max_corner[d], recs[i].dims[d] = bmin[d], bmax[d], dims[d]
        recs[i].owner, recs[i].local_index = 0, i
    aspect = float(np.float32(96) / np.float32(64))
    refdir = tmp_path / "ref"
    refdir.mkdir()
    _order, want, _log = ref.order_graphs(1, recs, n, out["camera"], aspect, refdir)
    got = [(work / name).read_bytes() for name in names]
    assert got == want
