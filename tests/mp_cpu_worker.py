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

from amrvolumerenderer_b200 import ctypes_api as api, engine, scenes  # noqa: E402
from oracle.oracle import Oracle  # noqa: E402


def check_input_side(rank, world):
    """Plotfile boxes are dealt to ranks without loss or overlap, every rank reads
    exactly its FABs, and convexify on each rank's share against the GLOBAL fine
    BoxArray yields, over all ranks, the same uncovered cells as one rank does."""
    import ctypes as C
    import tempfile

    from amrvolumerenderer_b200 import ctypes_api as api
    from amrvolumerenderer_b200.plotfile import PlotFile, synthetic_hierarchy, write_plotfile
    from oracle import amr_io_oracle as aio

    lib = api.load_library()
    levels, plo, phi = synthetic_hierarchy(base_cells=16, box_cells=4, nlevels=3, seed=5)
    box = [None]
    if rank == 0:
        box[0] = str(write_plotfile(Path(tempfile.mkdtemp()) / "plt", levels, ("rho",), plo, phi,
                                    fabs_per_file=7))
    dist.broadcast_object_list(box, src=0)
    P32 = C.POINTER(C.c_int32)
    with PlotFile(box[0]) as pf:
        nlev = pf.finest_level + 1
        clev = (api.Level * nlev)()
        keep, gboxes = [], []
        for l in range(nlev):
            allb = np.ascontiguousarray(pf.boxes(l))
            gboxes.append(allb)
            idx, off = pf.level_share(l, rank, world, rule=l % 2)
            dst = np.empty(int(off[-1]))
            pf.read_level(l, 0, dst, rank, world, rule=l % 2, threads=2)
            shares = [None] * world
            dist.all_gather_object(shares, idx.tolist())
            assert sorted(b for sh in shares for b in sh) == list(range(len(allb)))
            fabs = (api.Fab * max(len(idx), 1))()
            for k, b in enumerate(idx):
                want = levels[l].boxes[b][1][0]
                assert np.array_equal(dst[off[k]:off[k + 1]].reshape(want.shape), want)
                nx, ny, nz = (int(allb[b, 3 + d] - allb[b, d] + 1) for d in range(3))
                f = fabs[k]
                for d in range(3):
                    f.lo[d], f.hi[d], f.begin[d] = int(allb[b, d]), int(allb[b, 3 + d]), int(allb[b, d])
                f.data = dst.ctypes.data + int(off[k]) * 8
                f.jstride, f.kstride, f.nstride = nx, nx * ny, nx * ny * nz
            keep += [dst, fabs]
            clev[l].nfabs, clev[l].fabs = len(idx), fabs
        ratios = (C.c_int32 * 3 * (nlev - 1))()
        for l in range(nlev - 1):
            for d in range(3):
                ratios[l][d] = 2
        ptrs = (P32 * nlev)(*[g.ctypes.data_as(P32) for g in gboxes])
        counts_in = (C.c_int32 * nlev)(*[len(g) for g in gboxes])
        out = (api.Fab * 4096)()
        counts = (C.c_int32 * nlev)()
        assert lib.amrvr_convexify_levels(nlev, clev, ptrs, counts_in, ratios, out, 4096, counts) == 0
        mine, base = [], 0
        for l in range(nlev):
            mine.append([[out[base + k].lo[d] for d in range(3)] + [out[base + k].hi[d] for d in range(3)]
                         for k in range(counts[l])])
            base += counts[l]
        everyone = [None] * world
        dist.all_gather_object(everyone, mine)
        for l in range(nlev):
            meta = pf.level(l)
            domain = tuple(meta.domain_lo) + tuple(meta.domain_hi)
            pieces = [p for r in everyone for p in r[l]]
            got, most = aio.pieces_mask(domain, pieces)
            fine = [tuple(b) for b in gboxes[l + 1]] if l + 1 < nlev else []
            want = aio.uncovered_mask(domain, [tuple(b) for b in gboxes[l]], fine, (2, 2, 2))
            assert most <= 1 and np.array_equal(got, want), (rank, l)


def check_context_from_env(rank, world):
    """Context.from_env() under torchrun: every rank receives rank 0's NCCL id
    (seen through a wrapped constructor), picks its own rank / device from the
    environment -- and, there being no GPU here, fails loudly instead of
    falling back to anything."""
    import torch

    seen = {}
    real_init = engine.Context.__init__

    def spy(self, device=0, rank=0, nranks=1, nccl_id=None):
        seen.update(device=device, rank=rank, nranks=nranks, nccl_id=nccl_id)
        real_init(self, device=device, rank=rank, nranks=nranks, nccl_id=nccl_id)

    engine.Context.__init__ = spy
    try:
        if torch.cuda.is_available():
            return  # this check is about the CPU-only container
        try:
            engine.Context.from_env()
            raise AssertionError("a context was created without a GPU")
        except api.AmrvrError as e:
            assert "no CUDA device" in str(e) or "CPU fallback" in str(e), str(e)
    finally:
        engine.Context.__init__ = real_init
    assert seen["rank"] == rank and seen["nranks"] == world
    assert seen["device"] == int(os.environ.get("LOCAL_RANK", rank))
    ids = [None] * world
    dist.all_gather_object(ids, seen["nccl_id"])
    assert ids[0] is not None and len(ids[0]) == api.NCCL_ID_BYTES and all(i == ids[0] for i in ids)


def check_striped_plan(rank, world):
    """A frame striped over the ranks (amrvr_stripe_plan_host): every rank computes the same
    plan (the peer-mapped pools grow collectively, so the sizes must agree without a
    message), the rows the ranks fold partition the frame, and what rank r holds of every
    footprint is exactly the rows of r's strips."""
    import ctypes as C
    import zlib

    lib = api.load_library()
    height, nb = 270, 23
    rng = np.random.default_rng(99)  # the same footprints on every rank
    rects = np.zeros((nb, 4), np.int32)
    for b in range(nb):
        h = int(rng.integers(1, 100))
        rects[b] = (4 * int(rng.integers(0, 30)), int(rng.integers(0, height - h + 1)),
                    4 * int(rng.integers(1, 30)), h)
    n = world * nb
    tile_base, ry_first = np.zeros(n, np.int32), np.zeros(n, np.int32)
    layer_off, pool = np.zeros(n, np.uint64), np.zeros(world, np.uint64)
    qbegin = np.zeros(world + 1, np.int32)
    P32, P64 = C.POINTER(C.c_int32), C.POINTER(C.c_uint64)
    assert lib.amrvr_stripe_plan_host(world, nb, rects.ctypes.data_as(P32), tile_base.ctypes.data_as(P32),
                                      ry_first.ctypes.data_as(P32), layer_off.ctypes.data_as(P64),
                                      qbegin.ctypes.data_as(P32), pool.ctypes.data_as(P64)) == 0
    digest = zlib.crc32(tile_base.tobytes() + ry_first.tobytes() + layer_off.tobytes() + pool.tobytes()
                        + qbegin.tobytes())
    digests = [None] * world
    dist.all_gather_object(digests, digest)
    assert all(d == digests[0] for d in digests)
    mine = [y for y in range(height) if (y >> 4) % world == rank]
    rows = [None] * world
    dist.all_gather_object(rows, mine)
    assert sorted(y for r in rows for y in r) == list(range(height))
    held = 0
    for b in range(nb):
        y0, h = int(rects[b, 1]), int(rects[b, 3])
        for y in range(y0, y0 + h):
            row = lib.amrvr_stripe_layer_row_host(world, rank, y0, y)
            assert (row >= 0) == (y in mine)
            held += row >= 0
    total = torch.tensor([held])
    dist.all_reduce(total)
    assert int(total.item()) == int(rects[:, 3].sum())


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
    check_input_side(rank, world)
    check_striped_plan(rank, world)
    check_context_from_env(rank, world)
    dist.barrier()
    if rank == 0:
        print("mp_cpu_worker OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
