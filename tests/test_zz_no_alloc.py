"""amrvr_brick_stats_static, the entry point for callers that may not allocate on the GPU.

Some hosts of this library cannot let it call cudaMalloc: an in-situ AMReX run has handed
almost all device memory to its arena before the first frame, and many codes forbid
allocation inside the time loop because the matching cudaFree stops the whole device.
For them the brick statistics (K2) exist a second time with the reduction's partial
records in a fixed array inside the library.  This file checks, through the C ABI, that
the variant returns exactly what amrvr_brick_stats returns (the statistics are min / max /
count, so the tolerance is zero, as in test_brick_stats), on two calls in a row whose box
lists differ -- the second call finds the first call's records still in the array and must
not read them -- and that it refuses what it cannot do without allocating.

Collected last on purpose (pytest -x stops at the first failure); passes alone in a fresh
process and after the rest of the suite.  No reference counterpart.
"""
import ctypes as C

import numpy as np
import pytest

from amrvolumerenderer_b200 import ctypes_api as api, engine, scenes

pytestmark = pytest.mark.gpu


def _device_boxes(scene, ghost_box=None):
    """amrvr_box views over torch device copies of the scene's FABs; `ghost_box` gets a
    one-cell ghost frame around it (begin = lo - 1, strides of the padded block)."""
    import torch

    cboxes, n, sel = scene.c_boxes()
    out = (api.Box * n)()
    keep = []
    for i, b in enumerate(sel):
        C.memmove(C.byref(out[i]), C.byref(cboxes[i]), C.sizeof(api.Box))
        data = np.ascontiguousarray(b.data)
        if i == ghost_box:
            nz, ny, nx = data.shape
            padded = np.full((nz + 2, ny + 2, nx + 2), 1.0e30)  # ghosts must not be counted
            padded[1:-1, 1:-1, 1:-1] = data
            data = padded
            for d in range(3):
                out[i].begin[d] = out[i].lo[d] - 1
            out[i].jstride = nx + 2
            out[i].kstride = (nx + 2) * (ny + 2)
            out[i].nstride = padded.size
        t = torch.from_numpy(data).cuda()
        keep.append(t)
        out[i].data = t.data_ptr()
    torch.cuda.synchronize()
    return out, n, keep


def test_static_scratch_stats_equal_the_allocating_entry_point_twice_in_a_row(ctx):
    # call 1: 64 bricks of 16^3 with a NaN, an infinity and a negative cell
    first = scenes.scene_c2(box_cells=16)
    first.boxes[1].data[0, 0, 0] = np.nan
    first.boxes[5].data[3, 2, 1] = np.inf
    first.boxes[2].data[1, 1, 1] = -3.5
    boxes1, n1, keep1 = _device_boxes(first)
    want1 = ctx.brick_stats(boxes1, n1)
    got1 = engine.brick_stats_static(boxes1, n1, device=0)
    assert got1 == want1
    allv = np.concatenate([b.data.ravel() for b in first.boxes])
    fin = allv[np.isfinite(allv)]
    assert got1 == (fin.min(), fin.max(), fin[fin > 0].min(), fin.size)

    # call 2, straight after: fewer and smaller boxes with other values (one of them a view
    # with ghost cells), on another stream -- the array still holds call 1's records
    second = scenes.scene_c2(box_cells=8)
    second.boxes = second.boxes[:5]
    for k, b in enumerate(second.boxes):
        b.data[...] = 100.0 + k + b.data
    second.boxes[3].data[2, 2, 2] = 7.25e-3
    second.__dict__.pop("_cbox_cache", None)
    boxes2, n2, keep2 = _device_boxes(second, ghost_box=3)
    got2 = engine.brick_stats_static(boxes2, n2, device=0, stream=ctx.stream)
    want2 = ctx.brick_stats(boxes2, n2)
    assert got2 == want2
    assert got2[3] == 5 * 8 ** 3 and got2[0] == 7.25e-3 and got2[1] < 1.0e30
    assert got2 != got1

    # and once more with the first list: same answer as before
    assert engine.brick_stats_static(boxes1, n1, device=0) == want1
    # an empty list is the reduction's identity, as with amrvr_brick_stats
    assert engine.brick_stats_static(boxes1, 0, device=0) == ctx.brick_stats(boxes1, 0)


def test_static_scratch_stats_refuse_what_would_need_an_allocation(ctx):
    scene = scenes.scene_c2(box_cells=4)
    host_boxes, n, _ = scene.c_boxes()
    with pytest.raises(api.AmrvrError, match="not in device memory") as err:
        engine.brick_stats_static(host_boxes, n, device=0)  # host FABs need a staging buffer
    assert err.value.status == api.AMRVR_ERR_INVALID
    too_many = api.BRICK_STATS_STATIC_MAX_BOXES + 1
    with pytest.raises(api.AmrvrError, match="exceed the fixed scratch array") as err:
        engine.brick_stats_static((api.Box * too_many)(), too_many, device=0)
    assert err.value.status == api.AMRVR_ERR_INVALID
    # the refusals leave the entry point usable
    dev_boxes, n, keep = _device_boxes(scene)
    assert engine.brick_stats_static(dev_boxes, n, device=0) == ctx.brick_stats(dev_boxes, n)
