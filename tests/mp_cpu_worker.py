"""Worker of tests/test_multiprocess_cpu.py: world_size >= 2 over gloo, no GPU.

Checks the host-side logic of the N>1 path: every rank shards the scene the
same way, derives the same rank order and layer order from the global box
table, owns the piece the oracle says it owns, and the exchange plan is
symmetric (what r plans to send to q is what q plans to receive from r), so
the alltoallv needs no count exchange."""
import os
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from amrvolumerenderer_b200 import engine, scenes  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402


def main():
    dist.init_process_group("gloo")
    rank, world = dist.get_rank(), dist.get_world_size()
    orc = Oracle()
    for factory, kw, rule, w, h in ((scenes.scene_c4, dict(box_cells=4), "round_robin", 160, 120),
                                    (scenes.scene_c3, dict(box_cells=4), "chunks", 101, 67),
                                    (scenes.scene_c1, dict(), None, 64, 48)):
        scene = factory(nranks=world, **kw) if rule is None else factory(nranks=world, rule=rule, **kw)
        recs, n = scene.records()
        cam = scenes.make_camera()

        # sharding: my boxes are exactly the records I own, in local order
        _cb, mine, sel = scene.c_boxes(only_rank=rank)
        counts = [None] * world
        dist.all_gather_object(counts, mine)
        assert sum(counts) == n
        owned = [i for i in range(n) if recs[i].owner == rank]
        assert len(owned) == mine and [recs[i].local_index for i in owned] == list(range(mine))

        # every rank derives the same order; it is the oracle's
        order, failed = engine.order_boxes_host(world, recs, n, cam, w / h)
        want, ok = orc.order(world, recs, n, cam, w / h)
        assert np.array_equal(order, want) and failed == (not ok)
        orders = [None] * world
        dist.all_gather_object(orders, order.tolist())
        assert all(o == orders[0] for o in orders)

        # pixel ownership
        send, recv, (b, e) = engine.exchange_plan_host(world, rank, recs, n, cam, w, h, order)
        piece = list(order).index(rank)
        assert (b, e) == orc.piece_range(w * h, piece, world)
        assert send[rank] == 0 and recv[rank] == 0

        # symmetry of the plan: all-to-all of the planned send counts
        sent = torch.from_numpy(send.copy())
        got = torch.zeros(world, dtype=torch.int64)
        dist.all_to_all_single(got, sent)
        assert np.array_equal(got.numpy(), recv), (rank, got.numpy(), recv)
        spans = [None] * world
        dist.all_gather_object(spans, (b, e))
        covered = sorted(spans)
        assert covered[0][0] == 0 and covered[-1][1] == w * h
        assert all(a[1] == c[0] for a, c in zip(covered, covered[1:]))
    dist.barrier()
    if rank == 0:
        print("mp_cpu_worker OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
