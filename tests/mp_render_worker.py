"""Worker of tests/test_multi_gpu.py::test_render_under_torchrun: the Python
entry point started the way the reference is started under mpirun.  Every rank
calls render() on the same plotfile WITHOUT a context (Context.from_env picks
rank, device and the NCCL id); rank 0 then renders the same plotfile alone and
compares with the oracle run on the P-rank scene the ranks report (float RGBA
<= 1e-4, bytes <= 1 LSB).  The same plotfile rendered by one rank alone is
printed beside it: that frame differs by a few LSB, as the reference's own
P-rank and 1-rank frames do (the run fold is not associative).

    python -m torch.distributed.run --nproc-per-node P tests/mp_render_worker.py SCRATCH_DIR
"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from amrvolumerenderer_b200.engine import Context  # noqa: E402
from amrvolumerenderer_b200.plotfile import synthetic_hierarchy, write_plotfile  # noqa: E402
from amrvolumerenderer_b200.render import render  # noqa: E402


def main():
    scratch = Path(sys.argv[1])
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")
    path = scratch / "plt"
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=31)
    if rank == 0:
        write_plotfile(path, levels, ("density",), plo, phi)
    dist.barrier()
    kw = dict(width=96, height=64, camera_eye=(1.9, 1.3, 2.4), camera_look_at=(0.5, 0.5, 0.5),
              camera_far=10.0, return_frame=True)
    out = render(str(path), output=str(scratch / "all.ppm"), **kw)
    # every rank reports its own convex boxes; rank 0 rebuilds the P-rank scene from them
    # (owner-major, local order kept = the MPI_Allgatherv order) and asks the oracle
    shares = [None] * world
    dist.all_gather_object(shares, out["boxes"])
    ok = True
    if rank == 0:
        from oracle.oracle import Oracle, read_ppm
        from amrvolumerenderer_b200 import ctypes_api as api, scenes
        from test_gpu_render_api import _scene_from_frame

        assert out["rgb8"] is not None and out["rgb8"].any()
        assert np.array_equal(read_ppm(scratch / "all.ppm"), out["rgb8"])
        boxes = []
        for owner, share in enumerate(shares):
            one = dict(out, boxes=share)
            for b in _scene_from_frame(one, levels, phi).boxes:
                b.owner = owner
                boxes.append(b)
        s = out["summary"]
        scene = scenes.Scene("plotfile", boxes, out["transform"], out["scalar_range"],
                             np.array(s.bounds_min), np.array(s.bounds_max), world, [])
        params = api.FrameParams(width=96, height=64, box_transparency=0.0, antialiasing=1,
                                 use_visibility_graph=1, count_samples=0, draw_bounding_box=1)
        want = Oracle().render_frame(scene, out["camera"], params)
        err = float(np.abs(out["final5"][..., :4] - want["final5"][..., :4]).max())
        lsb = int(np.abs(out["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
        ok = err <= 1e-4 and lsb <= 1 and sum(len(x) for x in shares) == 24 + 24 + 8
        # one rank alone renders the same plotfile: the same picture up to the fold's
        # association (the reference's P-rank and 1-rank frames differ the same way)
        alone = Context(device=local)
        try:
            single = render(str(path), output=str(scratch / "one.ppm"), context=alone, **kw)
        finally:
            alone.close()
        drift = int(np.abs(out["rgb8"].astype(int) - single["rgb8"].astype(int)).max())
        print(f"[mp_render_worker] P={world} boxes per rank={[len(x) for x in shares]} "
              f"max|dRGBA|={err:.3g} lsb={lsb} (vs one rank alone: {drift} LSB) {'OK' if ok else 'FAIL'}",
              flush=True)
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if flag[0] else 1)


if __name__ == "__main__":
    main()
