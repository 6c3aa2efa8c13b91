"""SURVEY H2's owed measurement: what contracting a*b+c into FMA in the march costs in
parity.  Run once with the shipped library and once with an -fmad=true build:

    python tools/fmad_probe.py                      # shipped build (-fmad=false)
    AMRVR_LIB=build_variants/fmad.so python tools/fmad_probe.py

Compares frames with the oracle (tests only) on a non-power-of-two grid (C2') and a
power-of-two one (C4 at 16^3 boxes): sample counts, max |dRGBA|, pixels over 1e-4, LSBs,
and times the full-size C4 march."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from amrvolumerenderer_b200 import ctypes_api as api, scenes  # noqa: E402
from amrvolumerenderer_b200.engine import Context  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402

ctx, orc = Context(device=0), Oracle()
camera = scenes.make_camera()
print("library:", os.environ.get("AMRVR_LIB", "shipped"))
for name, scene, w, h in (("C2' 27 x 30^3, dx not a power of two", scenes.scene_c2prime(box_cells=30), 320, 240),
                          ("C4 176 x 16^3, power-of-two cells", scenes.scene_c4(box_cells=16), 256, 256)):
    p = api.FrameParams(width=w, height=h, box_transparency=0.0, antialiasing=1, use_visibility_graph=1,
                        count_samples=1, draw_bounding_box=0)
    ds = ctx.upload_scene(scene)
    got = ctx.render_frame(ds, camera, p)
    want = orc.render_frame(scene, camera, p)
    d = np.abs(got["final5"][..., :4] - want["final5"][..., :4]).max(axis=-1)
    lsb = np.abs(got["rgb8"].astype(int) - want["rgb8"].astype(int)).max()
    covered = (want["final5"][..., 3] > 0).sum()
    print(f"{name}: samples {got['stats'].samples} vs oracle {want['samples']} "
          f"(diff {got['stats'].samples - want['samples']:+d}), max|dRGBA| {d.max():.3g}, "
          f"pixels over 1e-4: {(d > 1e-4).sum()} of {covered} covered, max LSB {lsb}")
    ds.free()
factory, W, H, aa, transp = bench.WORKLOADS["C4"]
scene = factory(nranks=1, rule=bench.DEAL)
ds = ctx.upload_scene(scene)
p = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa, use_visibility_graph=1,
                    count_samples=0, draw_bounding_box=1, rgb8_on_device=1)
ms = [ctx.render_frame(ds, camera, p, want_rgba=False, want_rgb8=False)["stats"].march_ms for _ in range(12)]
print(f"C4 full size march: {np.median(ms[3:]):.4f} ms")
