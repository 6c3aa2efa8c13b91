"""Worker of tests/test_multi_gpu.py: one process per GPU (torch.distributed.run).

Every rank uploads its own boxes, all ranks render one frame through
amrvr_render_frame (NCCL alltoallv composite + gather), and rank 0 compares the
result with the oracle run with the SAME rank count and box->rank map (the
reference's blend fold is not associative).  Also checks the
stand-alone order / composite / gather entry points: rank order bit-exact,
pixel ownership (piece range of every rank) bit-exact.

    python -m torch.distributed.run --nproc-per-node P tests/mp_frame_worker.py CASE
"""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from amrvolumerenderer_b200 import ctypes_api as api, scenes  # noqa: E402
from amrvolumerenderer_b200.engine import Context, nccl_unique_id  # noqa: E402

RGBA_TOL = 1e-4

CASES = {
    # name: (scene factory kwargs, width, height, aa, transparency, rule)
    "c4": (scenes.scene_c4, dict(box_cells=8), 160, 120, 1, 0.0, "chunks"),
    "c3_aa4": (scenes.scene_c3, dict(box_cells=8), 96, 96, 4, 0.15, "round_robin"),
    "c2prime": (scenes.scene_c2prime, dict(box_cells=20), 101, 67, 1, 0.0, "round_robin"),
    "c1_few_boxes": (scenes.scene_c1, dict(), 64, 48, 1, 0.0, "chunks"),
    # C5 compositor sweep at reduced size: 512 unit bricks, many layers per pixel
    "c5": (scenes.scene_c5, dict(n=8), 192, 192, 1, 0.0, "round_robin"),
}


def main():
    case = sys.argv[1]
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    box = [nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)

    factory, kw, w, h, aa, transp, rule = CASES[case]
    if factory is scenes.scene_c1:
        scene = factory(nranks=world)  # 8 boxes: ranks beyond 8 own nothing
    else:
        scene = factory(nranks=world, rule=rule, **kw)
    camera = scenes.make_camera()
    params = api.FrameParams(width=w, height=h, box_transparency=transp, antialiasing=aa,
                             use_visibility_graph=1, count_samples=1, draw_bounding_box=1)
    ctx = Context(device=local, rank=rank, nranks=world, nccl_id=box[0])
    ds = ctx.upload_scene(scene)
    got = ctx.render_frame(ds, camera, params)
    frame_order, frame_failed = ctx.last_rank_order()  # the ordering the frame itself computed
    stolen = [None] * world
    dist.all_gather_object(stolen, (int(got["stats"].stolen_tiles), int(got["stats"].exchanged_bytes)))

    # stand-alone entry points: order, march, composite, gather
    root_aa = int(round(aa ** 0.5))
    rw, rh = w * root_aa, h * root_aa
    order, failed = ctx.order_boxes(ds, camera, w / h)
    recs, n = scene.records()
    ok = True
    msg = []
    if rank == 0:
        from oracle.oracle import Oracle

        orc = Oracle()
        want = orc.render_frame(scene, camera, params)
        want_order, graph_ok = orc.order(world, recs, n, camera, w / h)
        err = float(np.abs(got["final5"][..., :4] - want["final5"][..., :4]).max())
        lsb = int(np.abs(got["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
        samples = torch.tensor([got["stats"].samples], dtype=torch.int64, device="cuda")
    else:
        samples = torch.tensor([got["stats"].samples], dtype=torch.int64, device="cuda")
    dist.all_reduce(samples)

    ref_d = None
    # march + composite + gather by hand, to check pixel ownership per rank
    from amrvolumerenderer_b200.engine import reference_sample_distance

    ref_d = reference_sample_distance(recs, n, [min(float(r.min_corner[d]) for r in recs[:n]) for d in range(3)],
                                      [max(float(r.max_corner[d]) for r in recs[:n]) for d in range(3)])
    layers = ctx.raymarch(ds, camera, rw, rh, transp, ref_d)
    begin, end, piece = ctx.composite(ds, layers, camera, order)
    full = ctx.gather(piece, begin, end, rw * rh, root=0)
    spans = [None] * world
    dist.all_gather_object(spans, (rank, begin, end))
    if rank == 0:
        if not np.array_equal(order, want_order):
            ok = False
            msg.append(f"rank order {order} != oracle {want_order}")
        if failed != (not graph_ok) or frame_failed != (not graph_ok):
            ok = False
            msg.append("graph_failed flag differs")
        if not np.array_equal(frame_order, want_order):
            ok = False
            msg.append(f"rank order of the frame {frame_order} != oracle {want_order}")
        for piece_index, owner in enumerate(want_order):
            b, e = orc.piece_range(rw * rh, piece_index, world)
            if spans[owner] != (owner, b, e):
                ok = False
                msg.append(f"ownership: rank {owner} has {spans[owner]}, oracle [{b},{e})")
        if err > RGBA_TOL:
            ok = False
            msg.append(f"max |dRGBA| = {err}")
        if lsb > 1:
            ok = False
            msg.append(f"max LSB = {lsb}")
        if int(samples.item()) != want["samples"]:
            ok = False
            msg.append(f"samples {int(samples.item())} != {want['samples']}")
        # the hand-driven path must give the same render-size image before post
        dense = np.frombuffer(ctx.read_device(full, rw * rh * 20), np.float32).reshape(rh, rw, 5)
        if aa == 1:
            rgb_no_overlay = orc.quantize_rgb8(dense)
            bbox_free = api.FrameParams(width=w, height=h, box_transparency=transp, antialiasing=aa,
                                        use_visibility_graph=1, count_samples=0, draw_bounding_box=0)
            want2 = orc.render_frame(scene, camera, bbox_free)
            d2 = float(np.abs(dense[..., :4] - want2["final5"][..., :4]).max())
            if d2 > RGBA_TOL:
                ok = False
                msg.append(f"stand-alone composite+gather differs by {d2}")
            del rgb_no_overlay
        ctx.lib.amrvr_device_free(ctx._h, full)
        print(f"[mp_frame_worker] case={case} P={world} err={err:.3g} lsb={lsb} "
              f"samples={int(samples.item())} order={list(order)} "
              f"replicated={ds.replicated} (stolen tiles, pushed bytes) per rank={stolen} "
              f"exchange={os.environ.get('AMRVR_EXCHANGE', 'peer')} "
              f"{'OK' if ok else 'FAIL ' + '; '.join(msg)}",
              flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, src=0)
    layers.free()
    ds.free()
    ctx.close()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
