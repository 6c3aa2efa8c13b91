#!/usr/bin/env python
"""Generate tests/golden/*.npz from the REFERENCE ITSELF.

The reference's tests hold no golden vector for the hot path, so
the fixtures are outputs of the reference's own unmodified sources
(oracle/_ref/libamrvr_ref.so, built by oracle/Makefile from /root/reference)
on small seeded inputs.  Run in the build container, where /root/reference
exists; the .npz files travel to the GPU box, the reference does not.

    python tests/golden/make_golden.py
"""
from __future__ import annotations

import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))

from amrvolumerenderer_b200 import ctypes_api as api, scenes  # noqa: E402
from oracle.oracle import Ref, read_ppm  # noqa: E402

OUT = Path(__file__).resolve().parent

# (name, scene factory kwargs are fixed here so tests can rebuild the same input)
ORDER_CAMERAS = [(1.9, 1.3, 2.4), (-1.2, 0.4, 0.8), (0.5, 3.1, 0.45), (2.2, -0.9, -1.0),
                 (0.5, 0.5, 3.0), (-0.7, -0.6, -0.9)]


def params(w, h, aa=1, transparency=0.0, graph=1, bbox=1):
    return api.FrameParams(width=w, height=h, box_transparency=transparency, antialiasing=aa,
                           use_visibility_graph=graph, count_samples=1, draw_bounding_box=bbox)


def frame(ref, scene, camera, p):
    with tempfile.TemporaryDirectory() as tmp:
        out = ref.render_scene(scene, camera, p, Path(tmp) / "f.ppm")
        return read_ppm(Path(tmp) / "f.ppm"), out["samples"]


def main():
    ref = Ref()

    # C1 (the reference's in-situ example, auto camera from mt19937(91021))
    rgb, samples = frame(ref, scenes.scene_c1(), None, params(64, 64))
    np.savez_compressed(OUT / "c1_frame_64.npz", rgb8=rgb, samples=np.int64(samples))

    # C1 layers: VolumePainter::paint on two of the eight boxes, explicit camera
    scene = scenes.scene_c1()
    cam = scenes.make_camera()
    cboxes, n, _ = scene.c_boxes()
    ref_d = 0.5 * (1.0 / 32.0)
    layers, counts = [], []
    for b in (0, 7):
        img, s = ref.paint(cboxes[b], scene.transform, *scene.scalar_range, 0.0, ref_d, 48, 40, cam)
        layers.append(img)
        counts.append(s)
    np.savez_compressed(OUT / "c1_layers_48x40.npz", boxes=np.array([0, 7]), layers=np.stack(layers),
                        samples=np.array(counts, np.int64), ref_distance=np.float32(ref_d))

    # C2' (non-power-of-two cells, aspect != 1), three ranks round-robin, whole frame
    rgb, samples = frame(ref, scenes.scene_c2prime(box_cells=10, nranks=3, rule="round_robin"),
                         scenes.make_camera(eye=(2.1, 2.3, 1.7), look_at=(0.5, 0.68, 0.45)),
                         params(80, 60))
    np.savez_compressed(OUT / "c2prime_frame_80x60_p3.npz", rgb8=rgb, samples=np.int64(samples))

    # C3 small, AA 4, box_transparency 0.15, two ranks
    rgb, samples = frame(ref, scenes.scene_c3(box_cells=4, nranks=2), scenes.make_camera(),
                         params(48, 48, aa=4, transparency=0.15))
    np.savez_compressed(OUT / "c3_frame_48_aa4_p2.npz", rgb8=rgb, samples=np.int64(samples))

    # visibility order of C4's 176 boxes for several cameras and rank counts
    orders = {}
    for P, rule in ((2, "chunks"), (4, "round_robin"), (8, "chunks")):
        sc = scenes.scene_c4(box_cells=4, nranks=P, rule=rule)
        recs, n = sc.records()
        rows = []
        for eye in ORDER_CAMERAS:
            got, _log = ref.order(P, recs, n, scenes.make_camera(eye=eye), 1.25)
            rows.append(got)
        orders[f"p{P}_{rule}"] = np.stack(rows)
    np.savez_compressed(OUT / "c4_rank_order.npz", cameras=np.array(ORDER_CAMERAS), **orders)

    # DirectSendBase::composeLayered on P = 3 ranks of painted C3 layers
    sc = scenes.scene_c3(box_cells=4, nranks=3, rule="round_robin")
    cam = scenes.make_camera()
    recs, n = sc.records()
    cboxes, n, _ = sc.c_boxes()
    w, h = 50, 37  # 1850 pixels: pieces of 616/616/618, not row aligned
    ref_d = 0.5 * (1.0 / 16.0)
    dense = [ref.paint(cboxes[b], sc.transform, *sc.scalar_range, 0.15, ref_d, w, h, cam)[0]
             for b in range(n)]
    from oracle.oracle import Oracle

    orc = Oracle()  # depth hints only (a6 is pinned separately by the frame fixtures)
    hints = np.array([orc.depth_hint(recs[i], cam) for i in range(n)], np.float32)
    order, _ = ref.order(3, recs, n, cam, w / h)
    full, owner, _sec = ref.compose(3, w, h, dense, sc.owners(), hints, order)
    np.savez_compressed(OUT / "c3_compose_50x37_p3.npz", full5=full, piece_owner=owner,
                        rank_order=order, hints=hints, ref_distance=np.float32(ref_d))
    for f in sorted(OUT.glob("*.npz")):
        print(f.name, f.stat().st_size)


if __name__ == "__main__":
    main()
