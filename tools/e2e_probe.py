"""One GPU: end-to-end time of upload + frame with the upload waited for and not
(amrvr_scene_upload_async), with the stage times of the frame.
    python tools/e2e_probe.py [C4]"""
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np  # noqa: E402

import bench  # noqa: E402
from amrvolumerenderer_b200 import ctypes_api as api, scenes  # noqa: E402
from amrvolumerenderer_b200.engine import Context  # noqa: E402

workload = sys.argv[1] if len(sys.argv) > 1 else "C4"
factory, W, H, aa, transp = bench.WORKLOADS[workload]
scene = factory(nranks=1, rule=bench.DEAL)
camera = scenes.make_camera()
lib = api.load_library()
pinned, fab_bytes = bench.pin_scene(lib, scene, 0, 1)
ctx = Context(device=0)
params = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa, use_visibility_graph=1,
                         count_samples=0, draw_bounding_box=1)
ref = None
for wait in (True, False, True, False):
    rows = []
    for it in range(8):
        ctx.synchronize()
        t0 = time.perf_counter()
        ds = ctx.upload_scene(scene, wait=wait)
        t1 = time.perf_counter()
        out = ctx.render_frame(ds, camera, params, want_rgba=False)
        t2 = time.perf_counter()
        st = out["stats"]
        rows.append(((t2 - t0) * 1e3, (t1 - t0) * 1e3, st.pre_ms, st.march_ms, st.composite_ms, st.post_ms))
        if ref is None:
            ref = out["rgb8"].copy()
        assert np.array_equal(out["rgb8"], ref)
        ds.free()
    r = np.median(np.array(rows[2:]), axis=0)
    print(f"wait={wait}: e2e {r[0]:.3f} ms  upload call {r[1]:.3f}  frame: pre {r[2]:.3f} march {r[3]:.3f} "
          f"composite {r[4]:.3f} post {r[5]:.3f}", flush=True)
