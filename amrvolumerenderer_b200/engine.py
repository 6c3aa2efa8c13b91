"""Host-side handles over the C ABI (include/amrvr.h): one :class:`Context` per
GPU, :class:`DeviceScene` for uploaded bricks, :class:`Layers` for a marched
frame.  Every compute call goes to libamrvr_b200.so; nothing here computes
pixels on the CPU and nothing here touches ``oracle/``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import ctypes_api as api

_P = C.POINTER


def _check(lib, ctx, status):
    if status != api.AMRVR_OK:
        raise api.AmrvrError(status, lib.amrvr_last_error(ctx).decode())


def nccl_unique_id() -> bytes:
    lib = api.load_library()
    buf = C.create_string_buffer(api.NCCL_ID_BYTES)
    _check(lib, None, lib.amrvr_nccl_unique_id(buf))
    return buf.raw


class Context:
    """amrvr_ctx: one per GPU / rank."""

    def __init__(self, device: int = 0, rank: int = 0, nranks: int = 1, nccl_id: bytes | None = None):
        self.lib = api.load_library()
        self.rank, self.nranks, self.device = rank, nranks, device
        self._h = C.c_void_p()
        self._pinned_pool = {}
        idbuf = C.create_string_buffer(nccl_id, api.NCCL_ID_BYTES) if nccl_id else None
        _check(self.lib, None, self.lib.amrvr_ctx_create(device, rank, nranks, idbuf, C.byref(self._h)))

    @classmethod
    def from_env(cls) -> "Context":
        """The context of this process under `torch.distributed.run` (one process
        per GPU), the counterpart of the reference being started under mpirun:
        RANK / WORLD_SIZE / LOCAL_RANK pick rank and device, and rank 0's NCCL
        id reaches the others through torch.distributed (the default group is
        created over gloo when the caller has not initialised one; only this one
        small object travels that way).  Without WORLD_SIZE > 1: device 0, one rank."""
        import os

        world = int(os.environ.get("WORLD_SIZE", "1"))
        if world <= 1:
            return cls(device=int(os.environ.get("LOCAL_RANK", "0")))
        rank = int(os.environ["RANK"])
        local = int(os.environ.get("LOCAL_RANK", rank))
        import torch.distributed as dist

        if not dist.is_initialized():
            dist.init_process_group("gloo")
        box = [nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        return cls(device=local, rank=rank, nranks=world, nccl_id=box[0])

    def close(self):
        if self._h:
            for ptr, _arr in self._pinned_pool.values():
                self.lib.amrvr_host_free_pinned(ptr)
            self._pinned_pool.clear()
            self.lib.amrvr_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def _pinned(self, shape, dtype) -> np.ndarray:
        key = (tuple(shape), np.dtype(dtype).str)
        hit = self._pinned_pool.get(key)
        if hit is None:
            nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
            ptr = C.c_void_p()
            _check(self.lib, None, self.lib.amrvr_host_alloc_pinned(max(nbytes, 1), C.byref(ptr)))
            raw = (C.c_uint8 * max(nbytes, 1)).from_address(ptr.value)
            arr = np.frombuffer(raw, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
            hit = self._pinned_pool[key] = (ptr, arr)
        return hit[1]

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, status):
        _check(self.lib, self._h, status)

    @property
    def stream(self) -> int:
        return int(self.lib.amrvr_ctx_stream(self._h) or 0)

    @property
    def kernel_launches(self) -> int:
        return int(self.lib.amrvr_ctx_kernel_launches(self._h))

    def last_rank_order(self):
        """(rank_order, graph_failed) of the last render_frame: the visibility order
        BuildVisibilityOrderedGroup would hand to the compositor."""
        out = np.zeros(self.nranks, np.int32)
        failed = C.c_int(0)
        self.check(self.lib.amrvr_ctx_last_rank_order(self._h, out.ctypes.data_as(_P(C.c_int)),
                                                      C.byref(failed)))
        return out, bool(failed.value)

    def peer_bandwidth(self, nbytes: int, peer_offset: int = 1, write: bool = True, iters: int = 10) -> float:
        """Measured NVLink rate of this rank (GB/s), all ranks copying at once (collective)."""
        out = C.c_float(0)
        self.check(self.lib.amrvr_peer_bandwidth(self._h, int(nbytes), int(peer_offset), int(bool(write)),
                                                 int(iters), C.byref(out)))
        return float(out.value)

    def set_order_device_min_boxes(self, n: int):
        """Box count from which the visibility ordering uses its device kernels."""
        self.check(self.lib.amrvr_ctx_set_order_device_min_boxes(self._h, int(n)))

    def synchronize(self):
        self.check(self.lib.amrvr_synchronize(self._h))

    # ------------------------------------------------------------------ scene
    def brick_stats(self, cboxes, n):
        lo, hi, lp = C.c_double(), C.c_double(), C.c_double()
        cnt = C.c_int64()
        self.check(self.lib.amrvr_brick_stats(self._h, n, cboxes, C.byref(lo), C.byref(hi),
                                              C.byref(lp), C.byref(cnt)))
        return lo.value, hi.value, lp.value, cnt.value

    def scene_build(self, levels, nlevels, min_level=0, max_level=None, component=0,
                    log_scale=False, normalize_to_data_range=True, capacity=None):
        """amrvr_scene_build (= detail::BuildSceneGeometry) -> (Box array, SceneSummary)."""
        if max_level is None:
            max_level = nlevels - 1
        if capacity is None:
            capacity = sum(int(levels[i].nfabs) for i in range(nlevels))
        boxes = (api.Box * max(capacity, 1))()
        summary = api.SceneSummary()
        self.check(self.lib.amrvr_scene_build(self._h, nlevels, levels, int(min_level), int(max_level),
                                              int(component), int(bool(log_scale)),
                                              int(bool(normalize_to_data_range)), boxes, int(capacity),
                                              C.byref(summary)))
        return boxes, summary

    def scene_histogram(self, cboxes, n, transform, scalar_range, bin_count):
        """amrvr_scene_histogram (= detail::ComputeSceneHistogram) -> (counts, total)."""
        counts = np.zeros(max(int(bin_count), 1), np.uint64)
        total = C.c_uint64(0)
        self.check(self.lib.amrvr_scene_histogram(self._h, n, cboxes, C.byref(transform),
                                                  float(scalar_range[0]), float(scalar_range[1]),
                                                  int(bin_count), counts.ctypes.data_as(_P(C.c_uint64)),
                                                  C.byref(total)))
        return counts, int(total.value)

    def upload(self, cboxes, n, transform, scalar_range, wait=True, staged=False) -> "DeviceScene":
        """wait=False: amrvr_scene_upload_async -- returns with the FAB copies in flight
        (the arrays must stay alive and unchanged until the first frame or .wait()).
        staged=True: amrvr_scene_upload_staged -- reuse what the last scene_build /
        brick_stats of this context left in the staging buffer (unchanged FABs)."""
        h = C.c_void_p()
        fn = self.lib.amrvr_scene_upload if wait else self.lib.amrvr_scene_upload_async
        if staged:
            fn = self.lib.amrvr_scene_upload_staged
        self.check(fn(self._h, n, cboxes, C.byref(transform), float(scalar_range[0]),
                      float(scalar_range[1]), C.byref(h)))
        return DeviceScene(self, h)

    def upload_scene(self, scene, wait=True) -> "DeviceScene":
        """scene: amrvolumerenderer_b200.scenes.Scene; uploads this rank's boxes."""
        cboxes, n, keep = scene.c_boxes(only_rank=self.rank if scene.nranks > 1 else None)
        ds = self.upload(cboxes, n, scene.transform, scene.scalar_range, wait=wait)
        ds._keep = keep
        return ds

    # ----------------------------------------------------------------- frames
    def render_frame(self, dscene, camera, params, colormap=None, root=0, want_rgba=True,
                     want_rgb8=True):
        pts, npts = api.colormap_array(colormap)
        is_root = self.rank == root
        # results land in pinned host buffers owned by the context (a pageable
        # destination makes the device->host copy synchronous and ~5x slower);
        # they are overwritten by the next frame of the same size
        rgba = self._pinned((params.height, params.width, 5), np.float32) if (want_rgba and is_root) else None
        rgb8 = self._pinned((params.height, params.width, 3), np.uint8) if (want_rgb8 and is_root) else None
        stats = api.FrameStats()
        self.check(self.lib.amrvr_render_frame(
            self._h, dscene._h, C.byref(camera), C.byref(params), pts, npts, root,
            rgba.ctypes.data_as(C.c_void_p) if rgba is not None else None,
            rgb8.ctypes.data_as(C.c_void_p) if rgb8 is not None else None, C.byref(stats)))
        return {"final5": rgba, "rgb8": rgb8, "stats": stats}

    def raymarch(self, dscene, camera, width, height, box_transparency, ref_distance,
                 colormap=None, count_samples=False) -> "Layers":
        pts, npts = api.colormap_array(colormap)
        h = C.c_void_p()
        self.check(self.lib.amrvr_raymarch(self._h, dscene._h, C.byref(camera), width, height,
                                           float(box_transparency), float(ref_distance), pts, npts,
                                           int(count_samples), C.byref(h)))
        return Layers(self, h, width, height)

    def order_boxes(self, dscene, camera, aspect, use_graph=True):
        out = np.zeros(self.nranks, np.int32)
        failed = C.c_int(0)
        self.check(self.lib.amrvr_order_boxes(self._h, dscene._h, C.byref(camera), float(aspect),
                                              int(use_graph), out.ctypes.data_as(_P(C.c_int)),
                                              C.byref(failed)))
        return out, bool(failed.value)

    def order_records(self, nranks, records, n, camera, aspect, use_graph=True):
        out = np.zeros(nranks, np.int32)
        failed = C.c_int(0)
        self.check(self.lib.amrvr_order_records(self._h, nranks, n, records, C.byref(camera),
                                                float(aspect), int(use_graph),
                                                out.ctypes.data_as(_P(C.c_int)), C.byref(failed)))
        return out, bool(failed.value)

    def composite_virtual(self, dscene, layers, camera, nvirtual, box_owner, rank_order):
        """Single-GPU emulation of an nvirtual-rank composite -> (full5, piece_owner)."""
        own = np.ascontiguousarray(box_owner, np.int32)
        ro = np.ascontiguousarray(rank_order, np.int32)
        npix = layers.width * layers.height
        dev = C.c_void_p()
        self.check(self.lib.amrvr_device_alloc(self._h, npix * 20, C.byref(dev)))
        piece_owner = np.full(npix, -1, np.int32)
        try:
            self.check(self.lib.amrvr_composite_virtual(
                self._h, dscene._h, layers._h, C.byref(camera), nvirtual,
                own.ctypes.data_as(_P(C.c_int)), ro.ctypes.data_as(_P(C.c_int)), dev,
                piece_owner.ctypes.data_as(_P(C.c_int))))
            full = np.zeros((layers.height, layers.width, 5), np.float32)
            self.check(self.lib.amrvr_copy_to_host(self._h, full.ctypes.data_as(C.c_void_p), dev,
                                                   npix * 20))
        finally:
            self.lib.amrvr_device_free(self._h, dev)
        return full, piece_owner

    def composite(self, dscene, layers, camera, rank_order):
        ro = np.ascontiguousarray(rank_order, np.int32)
        b, e = C.c_int(0), C.c_int(0)
        piece = C.c_void_p()
        self.check(self.lib.amrvr_composite(self._h, dscene._h, layers._h, C.byref(camera),
                                            ro.ctypes.data_as(_P(C.c_int)), C.byref(b), C.byref(e),
                                            C.byref(piece)))
        return b.value, e.value, piece

    def last_composite_ms(self):
        """Device ms of the last composite that pushed slices to peers: (wait for free pools,
        push kernel, wait for arrival, fold table copy + host gap, fold_bin + fold)."""
        out = (C.c_float * 5)()
        self.check(self.lib.amrvr_ctx_last_composite_ms(self._h, out))
        return tuple(float(x) for x in out)

    def read_device(self, ptr, nbytes) -> bytes:
        buf = np.zeros(nbytes, np.uint8)
        self.check(self.lib.amrvr_copy_to_host(self._h, buf.ctypes.data_as(C.c_void_p), ptr, nbytes))
        return buf

    def gather(self, piece, begin, end, npix, root=0):
        full = C.c_void_p()
        if self.rank == root:
            self.check(self.lib.amrvr_device_alloc(self._h, npix * 20, C.byref(full)))
        self.check(self.lib.amrvr_gather(self._h, piece, begin, end, npix, root, full))
        return full


class DeviceScene:
    def __init__(self, ctx: Context, handle):
        self.ctx, self._h, self._keep = ctx, handle, None

    def free(self):
        if self._h:
            self.ctx.lib.amrvr_scene_free(self.ctx._h, self._h)
            self._h = C.c_void_p()

    @property
    def was_staged(self) -> bool:
        """True when amrvr_scene_upload_staged found the FABs in the staging buffer."""
        return bool(self.ctx.lib.amrvr_scene_upload_was_staged(self._h))

    def wait(self):
        """Blocks until an asynchronous upload has finished."""
        self.ctx.check(self.ctx.lib.amrvr_scene_wait(self.ctx._h, self._h))

    @property
    def replicated(self) -> bool:
        return bool(self.ctx.lib.amrvr_scene_is_replicated(self._h))

    def upload_stats(self):
        """(device ms of the FAB copies + quantise, FAB bytes, cells) of this rank's upload."""
        ms, nbytes, cells = C.c_float(0), C.c_int64(0), C.c_int64(0)
        self.ctx.check(self.ctx.lib.amrvr_scene_upload_stats(self._h, C.byref(ms), C.byref(nbytes),
                                                             C.byref(cells)))
        return float(ms.value), int(nbytes.value), int(cells.value)

    def global_boxes(self):
        n = C.c_int(0)
        recs = _P(api.BoxRecord)()
        self.ctx.check(self.ctx.lib.amrvr_scene_global_boxes(self._h, C.byref(n), C.byref(recs)))
        return recs, n.value

    def read_brick(self, local_box: int, ncells: int) -> np.ndarray:
        out = np.zeros(ncells, np.uint8)
        self.ctx.check(self.ctx.lib.amrvr_scene_read_brick(self.ctx._h, self._h, local_box,
                                                           out.ctypes.data_as(C.c_void_p)))
        return out


class Layers:
    def __init__(self, ctx: Context, handle, width, height):
        self.ctx, self._h, self.width, self.height = ctx, handle, width, height

    def free(self):
        if self._h:
            self.ctx.lib.amrvr_layers_free(self.ctx._h, self._h)
            self._h = C.c_void_p()

    def __len__(self):
        return int(self.ctx.lib.amrvr_layers_count(self._h))

    def stats(self) -> api.MarchStats:
        s = api.MarchStats()
        self.ctx.check(self.ctx.lib.amrvr_layers_stats(self._h, C.byref(s)))
        return s

    def rect(self, layer: int):
        v = [C.c_int(0) for _ in range(4)]
        self.ctx.check(self.ctx.lib.amrvr_layers_rect(self._h, layer, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def read(self, layer: int) -> np.ndarray:
        out = np.zeros((self.height, self.width, 5), np.float32)
        self.ctx.check(self.ctx.lib.amrvr_layers_read(self.ctx._h, self._h, layer,
                                                      out.ctypes.data_as(C.c_void_p)))
        return out


# -------------------------------------------- calls that need no context

def brick_stats_static(cboxes, n, device: int = 0, stream: int | None = None):
    """amrvr_brick_stats_static: (min, max, min positive, finite count) of FABs that are
    already in device memory, reduced without a Context and without any device allocation.

    For hosts that cannot let a library allocate on the GPU (an in-situ simulation whose
    arena already holds the device's memory, a time loop that bans cudaMalloc): the
    reduction's partial records sit in a fixed array inside libamrvr_b200.so.  One call
    at a time per GPU (all calls share that array), at most
    ``api.BRICK_STATS_STATIC_MAX_BOXES`` boxes, device-resident data only -- anything
    else raises AmrvrError.  ``stream`` is a raw cudaStream_t (None = default stream)."""
    lib = api.load_library()
    lo, hi, lp = C.c_double(), C.c_double(), C.c_double()
    cnt = C.c_int64()
    _check(lib, None, lib.amrvr_brick_stats_static(int(device), C.c_void_p(stream or None), int(n),
                                                   cboxes, C.byref(lo), C.byref(hi), C.byref(lp),
                                                   C.byref(cnt)))
    return lo.value, hi.value, lp.value, cnt.value


# ------------------------------------------------------------ host-only calls

def color_table(alpha_scale, norm, rmin, rmax, colormap=None) -> np.ndarray:
    lib = api.load_library()
    out = np.zeros((256, 4), np.float32)
    pts, n = api.colormap_array(colormap)
    _check(lib, None, lib.amrvr_color_table(float(alpha_scale), float(norm), float(rmin),
                                            float(rmax), pts, n,
                                            out.ctypes.data_as(_P(C.c_float))))
    return out


def order_boxes_host(nranks, records, n, camera, aspect, use_graph=True):
    lib = api.load_library()
    out = np.zeros(nranks, np.int32)
    failed = C.c_int(0)
    _check(lib, None, lib.amrvr_order_boxes_host(nranks, n, records, C.byref(camera), float(aspect),
                                                 int(use_graph), out.ctypes.data_as(_P(C.c_int)),
                                                 C.byref(failed)))
    return out, bool(failed.value)


def order_graph_dots_host(nranks, records, n, camera, aspect) -> list[bytes]:
    """The DOT text of every round's visibility graph (exportGraph,
    Common/VisibilityOrdering.cpp:319-352), byte for byte."""
    lib = api.load_library()
    docs = []
    round_ = 0
    while True:
        needed, nrounds = C.c_int64(0), C.c_int(0)
        _check(lib, None, lib.amrvr_order_graph_dot_host(nranks, n, records, C.byref(camera),
                                                         float(aspect), round_, None, 0,
                                                         C.byref(needed), C.byref(nrounds)))
        if round_ >= nrounds.value:
            return docs
        buf = C.create_string_buffer(max(int(needed.value), 1))
        _check(lib, None, lib.amrvr_order_graph_dot_host(nranks, n, records, C.byref(camera),
                                                         float(aspect), round_, buf, needed.value,
                                                         None, None))
        docs.append(buf.raw[:needed.value])
        round_ += 1


_graph_file_counter = 0  # the reference's function-local static (VisibilityOrdering.cpp:324)


def write_visibility_graphs(nranks, records, n, camera, aspect, directory=".") -> list[str]:
    """What BuildVisibilityOrderedGroup does on rank 0 when writeVisibilityGraph
    is set: visibility_graph_<counter>.dot per round in the working directory,
    one "Wrote visibility graph to '...'" line each on stdout."""
    import os
    import sys

    global _graph_file_counter
    written = []
    for doc in order_graph_dots_host(nranks, records, n, camera, aspect):
        name = f"visibility_graph_{_graph_file_counter}.dot"
        _graph_file_counter += 1
        try:
            with open(os.path.join(directory, name), "wb") as fh:
                fh.write(doc)
        except OSError:
            print(f"Failed to write visibility graph to '{name}'", file=sys.stderr)
            continue
        print(f"Wrote visibility graph to '{name}'")
        written.append(name)
    return written


def layer_order_host(records, n, camera):
    lib = api.load_library()
    hints = np.zeros(max(n, 1), np.float32)
    order = np.zeros(max(n, 1), np.int32)
    _check(lib, None, lib.amrvr_layer_order_host(n, records, C.byref(camera),
                                                 hints.ctypes.data_as(_P(C.c_float)),
                                                 order.ctypes.data_as(_P(C.c_int))))
    return hints[:n], order[:n]


def piece_range(npix, piece, npieces):
    lib = api.load_library()
    b, e = C.c_int(0), C.c_int(0)
    _check(lib, None, lib.amrvr_piece_range(npix, piece, npieces, C.byref(b), C.byref(e)))
    return b.value, e.value


def reference_sample_distance(records, n, bmin, bmax) -> float:
    lib = api.load_library()
    out = C.c_float(0)
    lo = (C.c_double * 3)(*[float(x) for x in bmin])
    hi = (C.c_double * 3)(*[float(x) for x in bmax])
    _check(lib, None, lib.amrvr_reference_sample_distance(n, records, lo, hi, C.byref(out)))
    return out.value


def exchange_plan_host(nranks, rank, records, n, camera, width, height, rank_order):
    """(send_floats[P], recv_floats[P], (begin, end)) of amrvr_composite for `rank`."""
    lib = api.load_library()
    ro = np.ascontiguousarray(rank_order, np.int32)
    send = np.zeros(nranks, np.int64)
    recv = np.zeros(nranks, np.int64)
    b, e = C.c_int(0), C.c_int(0)
    _check(lib, None, lib.amrvr_exchange_plan_host(
        nranks, rank, n, records, C.byref(camera), width, height, ro.ctypes.data_as(_P(C.c_int)),
        send.ctypes.data_as(_P(C.c_int64)), recv.ctypes.data_as(_P(C.c_int64)), C.byref(b),
        C.byref(e)))
    return send, recv, (b.value, e.value)
