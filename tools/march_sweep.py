"""Compare builds of libamrvr_b200.so on one GPU: march / fold / frame time of the
bench frame and a CRC of the RGB8 frame (variants must agree bit for bit).

    python tools/march_sweep.py [--workload C4] lib_a.so lib_b.so ...
Each library runs in its own process (AMRVR_LIB).  Build variants with
    make -C amrvolumerenderer_b200/csrc OUT=/abs/path/lib_x.so OBJ=/abs/path/obj_x EXTRA_NVFLAGS=-DAMRVR_...=...
"""
import os
import subprocess
import sys
import zlib
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def child(workload, frames):
    import numpy as np
    import torch  # noqa: F401  (L2 flush buffer)

    import bench
    from amrvolumerenderer_b200 import ctypes_api as api, scenes
    from amrvolumerenderer_b200.engine import Context

    factory, W, H, aa, transp = bench.WORKLOADS[workload]
    scene = factory(nranks=1, rule=bench.DEAL)
    camera = scenes.make_camera()
    ctx = Context(device=0)
    ds = ctx.upload_scene(scene)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
    base = dict(width=W, height=H, box_transparency=transp, antialiasing=aa, use_visibility_graph=1,
                draw_bounding_box=1)
    counted = ctx.render_frame(ds, camera, api.FrameParams(count_samples=1, **base), want_rgba=False)
    crc = zlib.crc32(counted["rgb8"].tobytes())
    p = api.FrameParams(count_samples=0, rgb8_on_device=1, **base)
    march, fold, frame, span, resident, pre, comp, post = [], [], [], [], [], [], [], []
    for i in range(frames + 3):
        flush.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", 0))
        e0.record(stream)
        st = ctx.render_frame(ds, camera, p, want_rgba=False, want_rgb8=False)["stats"]
        e1.record(stream)
        torch.cuda.synchronize()
        if i >= 3:
            march.append(st.march_ms)
            fold.append(st.fold_ms)
            frame.append(e0.elapsed_time(e1))
            pre.append(st.pre_ms)
            comp.append(st.composite_ms)
            post.append(st.post_ms)
            span.append(st.march_span_ns * 1e-6)
            resident.append(st.march_cta_busy_ns / max(st.march_span_ns * max(st.march_ctas, 1), 1))
    s = int(counted["stats"].samples)
    print(f"{os.environ.get('AMRVR_LIB', 'default'):60s} march {np.median(march):.4f} (min {min(march):.4f}) "
          f"pre {np.median(pre):.4f} composite {np.median(comp):.4f} post {np.median(post):.4f} "
          f"span {np.median(span):.4f} resident {np.median(resident):.3f} "
          f"fold {np.median(fold):.4f} frame {np.median(frame):.4f} ms  "
          f"{s / (np.median(march) * 1e-3) / 1e9:.1f} Gsamples/s  samples {s} crc {crc:08x}", flush=True)


if __name__ == "__main__":
    args = sys.argv[1:]
    if args and args[0] == "--child":
        child(args[1], int(args[2]))
        sys.exit(0)
    workload, frames = "C4", 15
    if args and args[0] == "--workload":
        workload, args = args[1], args[2:]
    for lib in args or [""]:
        env = dict(os.environ)
        if lib:
            env["AMRVR_LIB"] = str(Path(lib).resolve())
        subprocess.run([sys.executable, __file__, "--child", workload, str(frames)], env=env, cwd=ROOT)
