"""Worker of tests/test_multi_gpu.py::test_render_under_torchrun: the Python
entry point started the way the reference is started under mpirun.  Every rank
calls render() on the same plotfile WITHOUT a context (Context.from_env picks
rank, device and the NCCL id); rank 0 then renders the same plotfile alone and
compares bytes (<= 1 LSB: the blend fold differs in association only).

    python -m torch.distributed.run --nproc-per-node P tests/mp_render_worker.py SCRATCH_DIR
"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

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
    if rank == 0:
        levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=8, nlevels=3, seed=31)
        write_plotfile(path, levels, ("density",), plo, phi)
    dist.barrier()
    kw = dict(width=96, height=64, camera_eye=(1.9, 1.3, 2.4), camera_look_at=(0.5, 0.5, 0.5),
              camera_far=10.0, return_frame=True)
    out = render(str(path), output=str(scratch / "all.ppm"), **kw)
    ok = True
    if rank == 0:
        assert out["rgb8"] is not None and out["rgb8"].any()
        assert (scratch / "all.ppm").exists()
        alone = Context(device=local)
        try:
            want = render(str(path), output=str(scratch / "one.ppm"), context=alone, **kw)
        finally:
            alone.close()
        lsb = int(np.abs(out["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
        ok = lsb <= 1
        print(f"[mp_render_worker] P={world} lsb={lsb} {'OK' if ok else 'FAIL'}", flush=True)
    flag = [ok]
    dist.broadcast_object_list(flag, src=0)
    dist.destroy_process_group()
    sys.exit(0 if flag[0] else 1)


if __name__ == "__main__":
    main()
