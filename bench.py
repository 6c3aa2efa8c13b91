#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: "Gsamples/s
ray-march + frames/s 2048^2 3-level AMR at 1/2/4/8 B200").

A step = one frame of workload C4 (synthetic 3-level AMR, 1024^3 effective, 176
boxes of 64^3, 2048x2048, visibility graph on) through
amrvr_render_frame: ray-march -> visibility order -> sort-last composite ->
byte image on the root.  The boxes are dealt to the N ranks (the box->rank map
fixes the reference's blend structure); the frame is fixed, so scaling is
"strong".  `value` = ray-march samples of the whole frame / device time per
frame, bricks resident in HBM.  `e2e` = the same frame through the public API
from pinned HOST FABs: amrvr_scene_upload (H2D + quantise) + amrvr_render_frame
+ D2H of the RGB8 image, all inside the timed region.

Every line carries `parity`: the same scene, box->rank map and rank count
rendered at a reduced frame size and compared, untimed, with the oracle (and
with the reference's own PPM); a broken tolerance makes the run exit non-zero.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
"""
from __future__ import annotations

import argparse
import ctypes as C
import functools
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

from amrvolumerenderer_b200 import ctypes_api as api, scenes  # noqa: E402

METRIC = "raymarch_gsamples_per_s"
UNIT = "Gsamples/s"
WORKLOADS = {
    # name: (scene factory, width, height, antialiasing, box_transparency)
    "C4": (scenes.scene_c4, 2048, 2048, 1, 0.0),
    "C3": (scenes.scene_c3, 1024, 1024, 4, 0.15),
    "C2": (scenes.scene_c2, 1024, 1024, 1, 0.0),
    # C5, the compositor-only sweep (BASELINE.json configs[4]): 64 / 512 / 4096 layers at
    # 4096^2; the bricks are 2^3 cells, so the frame time is the composite's.  Read the
    # line's "composite" object, not `value`.
    "C5-4": (functools.partial(scenes.scene_c5, n=4), 4096, 4096, 1, 0.0),
    "C5-8": (functools.partial(scenes.scene_c5, n=8), 4096, 4096, 1, 0.0),
    "C5-16": (functools.partial(scenes.scene_c5, n=16), 4096, 4096, 1, 0.0),
}
# Box -> rank map of the synthetic scenes.  "skewed": box i -> rank (i + i div P) mod P over the
# (level, k, j, i) order; "chunks": contiguous chunks of ceil(n/P) boxes, the rule SURVEY.md 8(d) writes
# down for C3/C4 (AMRVR_BENCH_DEAL=chunks selects it).  With replicated bricks the march is dealt by image
# strips, so the rule no longer moves the march; it changes the same-owner run structure the fold follows
# (the reference's blend fold is not associative: the oracle gets the same map) and the FAB bytes each
# rank uploads -- chunks give one rank a whole refinement level, skewed gives every rank near and far boxes.
DEAL = os.environ.get("AMRVR_BENCH_DEAL", "skewed")
DEAL_TEXT = {"skewed": "skewed round-robin (i + i div P) mod P [SURVEY 8(d) names contiguous chunks: "
                       "AMRVR_BENCH_DEAL=chunks]",
             "chunks": "in contiguous chunks of ceil(n/P) (SURVEY 8(d))",
             "round_robin": "round-robin i mod P"}[DEAL]
NVLINK_NOMINAL_GBPS = 900.0  # NVLink 5, per direction per GPU (B200_PROFILING.md)
RGBA_TOL = 1e-4  # north_star: float RGBA within 1e-4 absolute per channel
LSB_TOL = 1      # north_star: 8-bit output within 1 LSB


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _ncu_value(pattern, kernel="march"):
    """A number of the newest committed `ncu --set full` summary of `kernel` under
    profiles/ (a capture of this same workload; ncu is never run inside the bench)."""
    import re

    files = sorted((ROOT / "profiles").glob(f"r*_{kernel}_full.txt"))
    if not files:
        return None, None
    best = files[-1]
    for ln in best.read_text().splitlines():
        m = re.match(pattern, ln)
        if m:
            return m, best.name
    return None, best.name


def _ncu_traffic(kernel="march"):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel."""
    import re

    files = sorted((ROOT / "profiles").glob(f"r*_{kernel}_full.txt"))
    if not files:
        return None, None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    for ln in files[-1].read_text().splitlines():
        m = re.match(r"\s*dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", ln)
        if m and seen < 2:
            total += float(m.group(2)) * unit.get(m.group(3), 1.0)
            seen += 1
    return (total if seen == 2 else None), files[-1].name


def _ncu_issue(march_s, sm_clock_hz, n_sm=148):
    """Issue-slot use of the march at the LIVE launch duration: warp instructions per
    launch (smsp__inst_executed.sum of the committed ncu capture of this workload --
    an instruction count, not a timing) / live duration / (SMs x 4 schedulers x clock)."""
    try:
        m, name = _ncu_value(r"\s*smsp__inst_executed\.sum\s+([0-9.]+)\s+inst")
        if m is None or march_s <= 0 or sm_clock_hz <= 0:
            return None
        inst = float(m.group(1))
        return {"warp_instructions_per_launch": inst,
                "achieved_per_sm_clock": inst / march_s / (n_sm * sm_clock_hz),
                "peak_per_sm_clock": 4.0,
                "frac": inst / march_s / (n_sm * 4.0 * sm_clock_hz),
                "source": f"profiles/{name} (instruction count of the timed instantiation) / live launch duration"}
    except Exception:  # never let a reporting extra break the bench line
        return None


def nvlink_counters(index: int):
    """NVLink data/raw byte counters of one GPU, summed over its links (NVML field values
    138-141, KiB; hardware counters, not arithmetic).  None when NVML cannot give them."""
    try:
        import pynvml

        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES")
        phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
        h = pynvml.nvmlDeviceGetHandleByIndex(phys)
        ids = [(pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_TX, 0xFFFFFFFF),
               (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_DATA_RX, 0xFFFFFFFF),
               (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_TX, 0xFFFFFFFF),
               (pynvml.NVML_FI_DEV_NVLINK_THROUGHPUT_RAW_RX, 0xFFFFFFFF)]
        vals = pynvml.nvmlDeviceGetFieldValues(h, ids)
        out = {}
        for name, v in zip(("data_tx", "data_rx", "raw_tx", "raw_rx"), vals):
            if v.nvmlReturn != 0:  # 3 = NVML_ERROR_NOT_SUPPORTED on the pool's boxes (seen in round 2)
                return None
            out[name] = int(v.value.ullVal) * 1024
        return out
    except Exception:
        return None


def _delta(after, before, per):
    if after is None or before is None:
        return None
    return {k: (after[k] - before[k]) / per for k in after}


class ClockSampler(threading.Thread):
    """SM clock and throttle reasons of one GPU while the timed region runs
    (NVML in-process, every 5 ms; nvidia-smi as the fallback)."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self._halt = index, threading.Event()
        self.sm, self.max_sm, self.reasons = [], [], set()
        self.nvml = self.handle = None
        try:
            import pynvml

            pynvml.nvmlInit()
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            phys = int(visible.split(",")[index]) if visible and visible.split(",")[index].isdigit() else index
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(phys)
            self.nvml = pynvml
        except Exception:
            self.nvml = None

    def _sample_nvml(self):
        n, h = self.nvml, self.handle
        self.sm.append(float(n.nvmlDeviceGetClockInfo(h, n.NVML_CLOCK_SM)))
        self.max_sm.append(float(n.nvmlDeviceGetMaxClockInfo(h, n.NVML_CLOCK_SM)))
        mask = n.nvmlDeviceGetCurrentClocksEventReasons(h)
        for name, bit in (("hw_slowdown", n.nvmlClocksEventReasonHwSlowdown),
                          ("hw_thermal_slowdown", n.nvmlClocksEventReasonHwThermalSlowdown),
                          ("sw_thermal_slowdown", n.nvmlClocksEventReasonSwThermalSlowdown),
                          ("sw_power_cap", n.nvmlClocksEventReasonSwPowerCap)):
            if mask & bit:
                self.reasons.add(name)

    def _sample_smi(self):
        out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                              "--format=csv,noheader,nounits"], capture_output=True, text=True,
                             timeout=5).stdout.strip()
        r = [c.strip() for c in out.split(",")]
        if len(r) >= 6:
            self.sm.append(float(r[0]))
            self.max_sm.append(float(r[1]))
            for i in range(4):
                if r[2 + i].lower().startswith("active"):
                    self.reasons.add(self.NAMES[i])

    def run(self):
        while not self._halt.is_set():
            try:
                if self.nvml is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                if self.nvml is not None:
                    self.nvml = None  # fall back to nvidia-smi
            self._halt.wait(0.005 if self.nvml is not None else 0.2)

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None,
                "sm_max_mhz": max(self.max_sm) if self.max_sm else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm),
                "source": "nvml" if self.nvml is not None else "nvidia-smi"}


def pin_scene(lib, scene, rank, nranks):
    """Copy this rank's FABs into pinned host memory (the e2e H2D source)."""
    mine = [b for b in scene.boxes if nranks == 1 or b.owner == rank]
    total = sum(b.data.nbytes for b in mine)
    base = C.c_void_p()
    st = lib.amrvr_host_alloc_pinned(max(total, 8), C.byref(base))
    if st != 0:
        raise RuntimeError(lib.amrvr_last_error(None).decode())
    off = 0
    for b in mine:
        view = np.ctypeslib.as_array((C.c_double * b.data.size).from_address(base.value + off))
        view[:] = b.data.ravel()
        b.data = view.reshape(b.data.shape)
        off += b.data.nbytes
    return base, total


def _host_cores():
    return len(os.sched_getaffinity(0))


def _mem_available_gb():
    try:
        for ln in open("/proc/meminfo"):
            if ln.startswith("MemAvailable:"):
                return int(ln.split()[1]) / (1 << 20)
    except OSError:
        pass
    return 0.0


def _sample_params(workload, size):
    _, W, H, aa, transp = WORKLOADS[workload]
    root = int(round(aa ** 0.5))
    return api.FrameParams(width=size // root, height=size // root, box_transparency=transp,
                           antialiasing=aa, use_visibility_graph=1, count_samples=1, draw_bounding_box=1)


def reference_arm(args, workload):
    """The reference's own CPU implementation (oracle/_ref: its unmodified sources
    against the AMReX/MPI shim; ranks are threads).  Each step is a bounded sample
    of the workload -- the same boxes and camera at 512x512, 1/16 of the rays --
    because a full 2048^2 frame takes its dense layers 14.8 GB and ~20 s; ONE such
    full-size frame is timed as well (`full_size_check`) so that the rate the
    sample gives can be compared with the rate of the real frame."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import Oracle, Ref, RefUnavailable

    factory, W, H, aa, transp = WORKLOADS[workload]
    cores = _host_cores()
    P = max(1, min(8, cores))
    scene = factory(nranks=P, rule=DEAL)
    camera = scenes.make_camera()
    sw = sh = 512
    params = _sample_params(workload, sw)
    kind = "reference"
    try:
        eng = Ref()
    except (RefUnavailable, OSError):
        eng, kind = Oracle(), "port"
    times, samples, stages = [], 0, {}
    for it in range(args.warmup + args.steps):
        t0 = time.perf_counter()
        if kind == "reference":
            out = eng.render_scene(scene, camera, params, "/tmp/amrvr_bench_ref.ppm")
            stages = eng.stage_seconds(out["log"])
        else:
            out = eng.render_frame(scene, camera, params)
        dt = time.perf_counter() - t0
        samples = out["samples"]
        if it >= args.warmup:
            times.append(dt)
    ms = 1e3 * float(np.mean(times))
    value = samples / (ms * 1e-3) / 1e9
    full = None
    if kind == "reference" and workload == "C4" and not args.no_full_size:
        need_gb = len(scene.boxes) * W * H * 20 / (1 << 30) * 1.25 + 2.0
        if _mem_available_gb() >= need_gb:
            fp = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa,
                                 use_visibility_graph=1, count_samples=1, draw_bounding_box=1)
            t0 = time.perf_counter()
            o = eng.render_scene(scene, camera, fp, "/tmp/amrvr_bench_ref_full.ppm")
            dt = time.perf_counter() - t0
            full = {"size": f"{W}x{H}", "ms": dt * 1e3, "samples": o["samples"],
                    "gsamples_per_s": o["samples"] / dt / 1e9,
                    "rate_vs_sample": (o["samples"] / dt / 1e9) / value if value > 0 else None,
                    "stages": eng.stage_seconds(o["log"]),
                    "note": "one full-size frame of the same scene, timed once: the reference's rate on "
                            "the real frame next to the rate its 512x512 sample gives"}
        else:
            full = {"skipped": f"needs ~{need_gb:.0f} GB of host memory for the reference's dense layers, "
                               f"{_mem_available_gb():.0f} GB available"}
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        # the same workload string as the GPU arm prints at this N; what is timed per step is
        # the bounded sample named beside it
        "config": {"workload": f"{workload}: {len(scene.boxes)} boxes, {scene.total_cells()} cells, "
                               f"{W}x{H} AA{aa}, visibility graph on, boxes dealt {DEAL_TEXT} to "
                               f"{args.gpus} rank(s)",
                   "sample": f"{sw}x{sh} frame of the {W}x{H} workload, {P} ranks-as-threads "
                             "(the reference's dense layers make the full frame 14.8 GB and ~20 s: "
                             "one is timed in full_size_check)",
                   "boxes": len(scene.boxes), "cells": scene.total_cells()},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": P, "host_cores": cores, "kind": kind,
                         "sample": f"{workload} scene, {sw}x{sh} render size, {P} ranks-as-threads on "
                                   f"{cores} host cores, full frame incl. composite; stages {stages}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "full_size_check": full,
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(workload):
    """Bounded CPU sample for the product line (rank 0, N=1)."""
    from oracle.oracle import Oracle, Ref, RefUnavailable

    cores = _host_cores()
    P = max(1, min(8, cores))
    factory = WORKLOADS[workload][0]
    scene = factory(nranks=P, rule=DEAL)
    sw = sh = 512
    params = _sample_params(workload, sw)
    try:
        eng, kind = Ref(), "reference"
    except (RefUnavailable, OSError):
        eng, kind = Oracle(), "port"
    t0 = time.perf_counter()
    if kind == "reference":
        out = eng.render_scene(scene, scenes.make_camera(), params, "/tmp/amrvr_bench_cpu.ppm")
        stages = eng.stage_seconds(out["log"])
    else:
        out = eng.render_frame(scene, scenes.make_camera(), params)
        stages, P = {}, 1
    dt = time.perf_counter() - t0
    return {"value": out["samples"] / dt / 1e9, "unit": UNIT, "cores": P, "host_cores": cores, "kind": kind,
            "sample": f"{workload} scene at {sw}x{sh} render size ({out['samples']} samples, "
                      f"{dt:.2f} s wall, {P} ranks-as-threads on {cores} host cores); reference stage "
                      f"times {stages}",
            "march_only_value": (out["samples"] / stages["march_s"] / 1e9) if stages.get("march_s") else None}


# --------------------------------------------------------------------- parity

def parity_check(ctx, dist, scene, camera, workload, world, rank, device):
    """Untimed: the bench's own scene, box->rank map and rank count at a reduced frame
    size, against the oracle (same rank count: the reference's blend fold is not
    associative) and against the PPM the reference itself writes.  Visibility order
    and pixel ownership bit-exact, sample count exact, float RGBA <= 1e-4, bytes <= 1 LSB."""
    import torch

    size = 512 if len(scene.boxes) <= 512 else 128  # the oracle's layers are dense frames
    params = _sample_params(workload, size)
    root = int(round(params.antialiasing ** 0.5))
    rw, rh = params.width * root, params.height * root
    ds = ctx.upload_scene(scene)
    got = ctx.render_frame(ds, camera, params)
    samples = torch.tensor([got["stats"].samples], dtype=torch.int64, device=device)
    if world > 1:
        dist.all_reduce(samples)
    frame_order, frame_failed = ctx.last_rank_order()
    # ownership through the Compositor::compose seam: every rank's region
    order, failed = ctx.order_boxes(ds, camera, params.width / params.height)
    recs, n = scene.records()
    from amrvolumerenderer_b200.engine import reference_sample_distance

    ref_d = reference_sample_distance(
        recs, n, [min(float(r.min_corner[d]) for r in recs[:n]) for d in range(3)],
        [max(float(r.max_corner[d]) for r in recs[:n]) for d in range(3)])
    layers = ctx.raymarch(ds, camera, rw, rh, params.box_transparency, ref_d)
    begin, end, piece = ctx.composite(ds, layers, camera, order)
    # ... and its pixels: the pieces the NCCL exchange + fold left on every rank, gathered on rank 0
    # (a12 content + a14), against the oracle's frame without the overlay, before any downsample
    full = ctx.gather(piece, begin, end, rw * rh, root=0)
    spans = [None] * world
    if world > 1:
        dist.all_gather_object(spans, (rank, begin, end))
    else:
        spans = [(0, begin, end)]
    # at N > 1 one more frame with replication switched off for this upload: every rank marches its
    # own boxes and the fragments cross with NCCL (the path a scene too large to replicate takes)
    got_os, ds_os = None, None
    if world > 1:
        keep = os.environ.get("AMRVR_REPLICATE_MAX_MB")
        os.environ["AMRVR_REPLICATE_MAX_MB"] = "0"
        try:
            ds_os = ctx.upload_scene(scene)
        finally:
            if keep is None:
                del os.environ["AMRVR_REPLICATE_MAX_MB"]
            else:
                os.environ["AMRVR_REPLICATE_MAX_MB"] = keep
        got_os = ctx.render_frame(ds_os, camera, params)
        samples_os = torch.tensor([got_os["stats"].samples], dtype=torch.int64, device=device)
        dist.all_reduce(samples_os)
    out = None
    if rank == 0:
        from oracle.oracle import Oracle, Ref, RefUnavailable, read_ppm

        orc = Oracle()
        want = orc.render_frame(scene, camera, params)
        want_order, graph_ok = orc.order(world, recs, n, camera, params.width / params.height)
        own_ok = True
        for piece_index, owner in enumerate(want_order):
            b, e = orc.piece_range(rw * rh, piece_index, world)
            own_ok = own_ok and tuple(spans[owner]) == (int(owner), b, e)
        err = float(np.abs(got["final5"][..., :4] - want["final5"][..., :4]).max())
        lsb = int(np.abs(got["rgb8"].astype(int) - want["rgb8"].astype(int)).max())
        out = {"ranks": world, "size": f"{params.width}x{params.height} AA{params.antialiasing}",
               "rank_order_equal": bool(np.array_equal(order, want_order) and
                                        np.array_equal(frame_order, want_order) and
                                        failed == (not graph_ok) and frame_failed == (not graph_ok)),
               "ownership_equal": bool(own_ok),
               "samples_equal": bool(int(samples.item()) == int(want["samples"])),
               "max_abs_rgba": err, "max_lsb": lsb, "against": "oracle, same rank count and box->rank map"}
        dense = np.frombuffer(ctx.read_device(full, rw * rh * 20), np.float32).reshape(rh, rw, 5)
        bare = api.FrameParams(width=rw, height=rh, box_transparency=params.box_transparency, antialiasing=1,
                               use_visibility_graph=1, count_samples=0, draw_bounding_box=0)
        want_bare = orc.render_frame(scene, camera, bare)
        out["seam_piece_max_abs_rgba"] = float(np.abs(dense[..., :4] - want_bare["final5"][..., :4]).max())
        out["seam"] = ("amrvr_raymarch -> amrvr_composite (slices pushed into the peers' receive pools, or NCCL; fold) -> amrvr_gather at "
                       f"{rw}x{rh}, vs the oracle's frame without overlay")
        ctx.lib.amrvr_device_free(ctx._h, full)
        if got_os is not None:
            out["object_space_frame"] = {
                "replicated": bool(ds_os.replicated),
                "max_abs_rgba": float(np.abs(got_os["final5"][..., :4] - want["final5"][..., :4]).max()),
                "max_lsb": int(np.abs(got_os["rgb8"].astype(int) - want["rgb8"].astype(int)).max()),
                "samples_equal": bool(int(samples_os.item()) == int(want["samples"])),
                "path": "AMRVR_REPLICATE_MAX_MB=0 for one upload: boxes marched by their owner, "
                        "fragments exchanged through the compose seam, same frame"}
        try:  # the reference's own frame: real VolumeRenderer::renderScene, ranks as threads
            ref = Ref()
            ref.render_scene(scene, camera, params, "/tmp/amrvr_bench_parity.ppm")
            ppm = read_ppm("/tmp/amrvr_bench_parity.ppm")
            out["reference_ppm_max_lsb"] = int(np.abs(got["rgb8"].astype(int) - ppm.astype(int)).max())
        except (RefUnavailable, OSError) as exc:
            out["reference_ppm_max_lsb"] = None
            out["reference_ppm_note"] = f"oracle/_ref unavailable: {exc}"
        osf = out.get("object_space_frame")
        out["ok"] = bool(out["rank_order_equal"] and out["ownership_equal"] and out["samples_equal"] and
                         err <= RGBA_TOL and lsb <= LSB_TOL and
                         out["seam_piece_max_abs_rgba"] <= RGBA_TOL and
                         (osf is None or (not osf["replicated"] and osf["samples_equal"] and
                                          osf["max_abs_rgba"] <= RGBA_TOL and osf["max_lsb"] <= LSB_TOL)) and
                         (out["reference_ppm_max_lsb"] is None or out["reference_ppm_max_lsb"] <= LSB_TOL))
        eq = lambda flag: "equal" if flag else "DIFFERENT"
        out["summary"] = (
            f"{world} rank(s) {out['size']} vs oracle: order {eq(out['rank_order_equal'])}, ownership "
            f"{eq(out['ownership_equal'])}, samples {eq(out['samples_equal'])}, max|dRGBA| {err:.3g}, "
            f"LSB {lsb}, reference PPM LSB {out['reference_ppm_max_lsb']}, compose-seam piece (exchange "
            f"+ fold + gather) max|d| {out['seam_piece_max_abs_rgba']:.3g}"
            + (f", object-space frame (unreplicated, own boxes) max|d| {osf['max_abs_rgba']:.3g} LSB "
               f"{osf['max_lsb']} samples {eq(osf['samples_equal'])}" if osf else "")
            + (" -- OK" if out["ok"] else " -- BROKEN"))
    layers.free()
    ds.free()
    if ds_os is not None:
        ds_os.free()
    return out


def composite_seam(ctx, dist, dscene, scene, camera, workload, world, rank, stream, flush, steps, gather_all):
    """N > 1, untimed part of the line: the Compositor::compose seam on RANK-OWNED layers at the
    workload's full render size (BASELINE.json configs[4] words the compositor sweep this way):
    every rank marches its own boxes (amrvr_raymarch), then amrvr_composite = host plan + the row
    slices other pieces need stored straight into their ranks' receive pools over NVLink
    (push_slices_kernel between two flag barriers; pack + ncclSend/ncclRecv without peer access) + fold.  Device time
    from CUDA events on the ctx stream around the call, host time around call + synchronise;
    bytes from amrvr_exchange_plan_host (every rank derives the same plan)."""
    import torch
    from amrvolumerenderer_b200.engine import exchange_plan_host, reference_sample_distance

    _, W, H, aa, transp = WORKLOADS[workload]
    root = int(round(aa ** 0.5))
    rw, rh = W * root, H * root
    recs, n = scene.records()
    order, _ = ctx.order_boxes(dscene, camera, W / H)
    ref_d = reference_sample_distance(
        recs, n, [min(float(r.min_corner[d]) for r in recs[:n]) for d in range(3)],
        [max(float(r.max_corner[d]) for r in recs[:n]) for d in range(3)])
    layers = ctx.raymarch(dscene, camera, rw, rh, transp, ref_d)
    send, recv, _span = exchange_plan_host(world, rank, recs, n, camera, rw, rh, order)
    sent = int(4 * (send.sum() - send[rank]))
    received = int(4 * (recv.sum() - recv[rank]))
    dev_ms, host_ms, stages = [], [], []
    pushing = os.environ.get("AMRVR_EXCHANGE") != "nccl"
    for it in range(2 + steps):
        flush.fill_(it & 0xFF)
        dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(stream)
        ctx.composite(dscene, layers, camera, order)
        e1.record(stream)
        ctx.synchronize()
        t1 = time.perf_counter()
        if it >= 2:
            dev_ms.append(e0.elapsed_time(e1))
            host_ms.append((t1 - t0) * 1e3)
            if pushing:
                try:
                    stages.append(ctx.last_composite_ms())
                except api.AmrvrError:  # no peer mapping on this box: the call took the NCCL path
                    pushing = False
    layers.free()
    stage = [float(x) for x in np.mean(np.array(stages), axis=0)] if stages else None
    rows = gather_all({"rank": rank, "device_ms": float(np.mean(dev_ms)), "host_ms": float(np.mean(host_ms)),
                       "sent_bytes": sent, "received_bytes": received,
                       "stage_ms": dict(zip(("wait_pools_free", "push_kernel", "wait_arrival", "table_copy_gap",
                                             "fold_bin_and_fold"), stage)) if stage else None,
                       "push_GBps": sent / (stage[1] * 1e-3) / 1e9 if stage and stage[1] > 0 else None})
    if rank != 0:
        return None
    worst = max(r["device_ms"] for r in rows)
    return {"what": f"amrvr_raymarch (own boxes) -> amrvr_composite at {rw}x{rh}, {n} layers owned by "
                    f"{world} ranks, mean of {steps} calls after 2 warm-up, L2 flushed between calls",
            "device_ms_max_rank": worst, "host_ms_max_rank": max(r["host_ms"] for r in rows),
            "exchange": "push over NVLink (peer-mapped receive pools)" if rows[0]["stage_ms"]
                        else "pack + ncclSend/ncclRecv",
            "push_GBps_per_rank": [r["push_GBps"] for r in rows],
            # what a rank RECEIVES over the time its peers' stores were in flight (its own push + the wait
            # for arrival): with contiguous pieces the centre of the image receives most, and that GPU's
            # ingress is the link that bounds the exchange
            "ingress_GBps_per_rank": [
                r["received_bytes"] / ((r["stage_ms"]["push_kernel"] + r["stage_ms"]["wait_arrival"]) * 1e-3) / 1e9
                if r["stage_ms"] and (r["stage_ms"]["push_kernel"] + r["stage_ms"]["wait_arrival"]) > 0 else None
                for r in rows],
            "sent_bytes_per_rank": [r["sent_bytes"] for r in rows],
            "sent_GBps_per_rank_over_device_ms": [r["sent_bytes"] / (r["device_ms"] * 1e-3) / 1e9 for r in rows],
            "aggregate_GBps": sum(r["sent_bytes"] for r in rows) / (worst * 1e-3) / 1e9,
            "per_rank": rows,
            "note": "device_ms spans the host plan, barrier, push (or pack + NCCL), barrier, fold_bin and fold of one call; the rate divides the "
                    "bytes a rank SENDS by that whole time, so it is a lower bound on the link rate during "
                    "the exchange itself"}


def full_size_check(scene, camera, workload, rgb8):
    """N=1, once, untimed: the timed frame itself (full size) against the oracle's."""
    from oracle.oracle import Oracle

    _, W, H, aa, transp = WORKLOADS[workload]
    root = int(round(aa ** 0.5))
    need_gb = len(scene.boxes) * W * root * H * root * 20 / (1 << 30) * 1.1 + 2.0
    if _mem_available_gb() < need_gb:
        return {"skipped": f"the oracle's dense layers need ~{need_gb:.0f} GB of host memory, "
                           f"{_mem_available_gb():.0f} GB available"}
    params = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa,
                             use_visibility_graph=1, count_samples=1, draw_bounding_box=1)
    t0 = time.perf_counter()
    want = Oracle().render_frame(scene, camera, params)
    lsb = np.abs(rgb8.astype(int) - want["rgb8"].astype(int))
    return {"size": f"{W}x{H}", "max_lsb": int(lsb.max()), "pixels_off_by_one": int((lsb.max(axis=-1) > 0).sum()),
            "samples_oracle": int(want["samples"]), "oracle_s": time.perf_counter() - t0,
            "against": "oracle (1 rank), whole RGB8 frame", "ok": bool(lsb.max() <= LSB_TOL)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-seam", action="store_true",
                    help="skip the untimed compose-seam measurement on rank-owned layers (N > 1)")
    ap.add_argument("--no-full-size", action="store_true",
                    help="skip the one full-size comparison with the oracle (N=1) / full-size reference frame")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    if args.impl == "reference":
        reference_arm(args, args.workload)
        return

    import torch
    import torch.distributed as dist

    from amrvolumerenderer_b200.engine import Context, nccl_unique_id

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torch.distributed.run")
    torch.cuda.set_device(local)
    device = f"cuda:{local}"
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        box = [nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=0)
        nccl_id = box[0]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def gather_all(obj):
        if world == 1:
            return [obj]
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out

    factory, W, H, aa, transp = WORKLOADS[args.workload]
    # box -> rank: skewed round-robin over the (level, k, j, i) order, so every rank
    # holds the same number of boxes of every level and of every grid column
    scene = factory(nranks=world, rule=DEAL)
    camera = scenes.make_camera()
    lib = api.load_library()
    pinned, fab_bytes = pin_scene(lib, scene, rank, world)
    ctx = Context(device=local, rank=rank, nranks=world, nccl_id=nccl_id)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)  # > 126 MB L2

    def fp(**kw):
        base = dict(width=W, height=H, box_transparency=transp, antialiasing=aa, use_visibility_graph=1,
                    count_samples=0, draw_bounding_box=1)
        base.update(kw)
        return api.FrameParams(**base)

    params = fp()
    resident = fp(rgb8_on_device=1)  # device-resident arm: the RGB8 frame stays on the root GPU
    counting = fp(count_samples=1)

    # ---- parity first (untimed); a broken tolerance fails the run after the line is printed
    parity = None
    if not args.no_parity:
        parity = parity_check(ctx, dist, scene, camera, args.workload, world, rank, device)

    dscene = ctx.upload_scene(scene)
    replicated = dscene.replicated
    # untimed: one frame of the sample-counting instantiation.  Its per-rank figures
    # (samples marched, bytes pushed, march ms) come from ONE launch on ONE rank each.
    out = ctx.render_frame(dscene, camera, counting, want_rgba=False)
    cs = out["stats"]
    counted = gather_all({"rank": rank, "samples": int(cs.samples), "march_ms": float(cs.march_ms),
                          "pushed_bytes": int(cs.exchanged_bytes), "stolen_tiles": int(cs.stolen_tiles),
                          "fragment_pixels": int(cs.fragment_pixels), "tiles": int(cs.tiles)})
    total_samples = sum(c["samples"] for c in counted)
    full_check = None
    if rank == 0 and world == 1 and not args.no_full_size and not args.no_parity and args.workload == "C4":
        full_check = full_size_check(scene, camera, args.workload, out["rgb8"])

    def one_frame():
        return ctx.render_frame(dscene, camera, resident, want_rgba=False, want_rgb8=False)

    # ---- device-resident arm: K frames, L2 flushed between frames, events on the ctx stream
    for _ in range(args.warmup):
        one_frame()
    sampler = ClockSampler(local)
    launches0 = ctx.kernel_launches
    frame_ms = []
    keys = ("pre_ms", "march_ms", "order_ms", "composite_ms", "fold_ms", "gather_ms", "post_ms")
    stage = {k: [] for k in keys}
    pushed, stolen = [], []
    span_ns, busy_ns, table_ns = [], [], []
    barrier()
    link0 = nvlink_counters(local) if world > 1 else None
    sampler.start()
    for _ in range(args.steps):
        with torch.cuda.stream(stream):
            flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(stream)
        res = one_frame()
        e1.record(stream)
        barrier()
        frame_ms.append(e0.elapsed_time(e1))
        for k in keys:
            stage[k].append(getattr(res["stats"], k))
        pushed.append(res["stats"].exchanged_bytes)
        stolen.append(res["stats"].stolen_tiles)
        span_ns.append(res["stats"].march_span_ns)
        busy_ns.append(res["stats"].march_cta_busy_ns)
        table_ns.append(res["stats"].march_table_ns)
    clocks = sampler.finish()
    link1 = nvlink_counters(local) if world > 1 else None
    launches = ctx.kernel_launches - launches0
    mine = {"rank": rank, "frame_ms": float(np.mean(frame_ms)),
            **{k: float(np.mean(v)) for k, v in stage.items()},
            "pushed_bytes": float(np.mean(pushed)), "stolen_tiles": float(np.mean(stolen)),
            "fragment_pixels": int(res["stats"].fragment_pixels), "tiles": int(res["stats"].tiles)}
    ctas = max(int(res["stats"].march_ctas), 1)
    # the march launch on the device clock: first CTA start -> last CTA end, how much of
    # (span x CTAs) the CTAs were resident, and the time a CTA spent staging transfer tables
    # NVLink bytes per frame as the hardware counted them over the timed frames (the frames'
    # own barriers ride on NCCL and are in there too)
    mine["nvlink_counted"] = _delta(link1, link0, args.steps)
    mine["march_span_ms"] = float(np.mean(span_ns)) * 1e-6
    mine["march_cta_resident_frac"] = float(np.mean(busy_ns)) / max(float(np.mean(span_ns)) * ctas, 1.0)
    mine["march_table_ms_per_cta"] = float(np.mean(table_ns)) * 1e-6 / ctas
    per_rank = gather_all(mine)
    ms_per_step = max(r["frame_ms"] for r in per_rank)

    # ---- NVLink: measured per-GPU rate of plain stores / loads into the ring neighbour, all ranks at once
    link = None
    if world > 1 and replicated:
        c0 = nvlink_counters(local)
        wr = ctx.peer_bandwidth(256 << 20, 1, True, 10)
        c1 = nvlink_counters(local)
        rd = ctx.peer_bandwidth(256 << 20, 1, False, 10)
        c2 = nvlink_counters(local)
        # the probe moves 11 x 256 MiB (warm-up + 10) each way: what the counters say about it
        link = gather_all({"rank": rank, "write_GBps": wr, "read_GBps": rd,
                           "counted_write_probe": _delta(c1, c0, 1), "counted_read_probe": _delta(c2, c1, 1)})

    # ---- the compose seam on rank-owned layers (N > 1): the path that moves fragments between ranks
    seam = None
    if world > 1 and not args.no_seam:
        seam = composite_seam(ctx, dist, dscene, scene, camera, args.workload, world, rank, stream, flush,
                              min(args.steps, 10), gather_all)

    # ---- end-to-end arm: pinned host FABs -> upload -> frame -> host RGB8
    e2e_ms, up_ms = [], []
    for it in range(2 + min(args.steps, 10)):
        barrier()
        t0 = time.perf_counter()
        # one rank: the upload returns with the copies in flight and the frame marches each
        # copy group as it lands (amrvr_scene_upload_async); several ranks: the bricks are
        # allgathered after the upload, so it is waited for
        ds2 = ctx.upload_scene(scene, wait=(world > 1))
        t1 = time.perf_counter()
        o2 = ctx.render_frame(ds2, camera, params, want_rgba=False)
        assert o2["rgb8"] is not None or rank != 0
        barrier()
        dt = (time.perf_counter() - t0) * 1e3
        ds2.free()
        if it >= 2:
            e2e_ms.append(dt)
            up_ms.append((t1 - t0) * 1e3)
    e2e_all = gather_all({"rank": rank, "e2e_ms": float(np.mean(e2e_ms)), "upload_ms": float(np.mean(up_ms)),
                          "fab_bytes": int(fab_bytes)})
    e2e_ms_mean = max(r["e2e_ms"] for r in e2e_all)
    h2d_total = sum(r["fab_bytes"] for r in e2e_all)

    # ---- in-situ case: the FABs already live in HBM (SceneBuilder.cpp:160-187 hands out
    # views of device memory when the simulation runs on the GPU)
    dev_fabs = None
    if world == 1:
        keep = []
        host_data = [b.data for b in scene.boxes]
        for b in scene.boxes:
            t = torch.from_numpy(np.ascontiguousarray(b.data)).to(device)
            keep.append(t)
            b.data = _DeviceArray(t)
        scene.__dict__.pop("_cbox_cache", None)
        dms, kms = [], []
        for it in range(2 + min(args.steps, 10)):
            torch.cuda.synchronize()
            with torch.cuda.stream(stream):
                flush.zero_()
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            ds3 = ctx.upload(*_device_boxes(scene), scene.transform, scene.scalar_range)
            o3 = ctx.render_frame(ds3, camera, params, want_rgba=False)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) * 1e3
            k5_ms, k5_bytes, k5_cells = ds3.upload_stats()
            ds3.free()
            if it >= 2:
                dms.append(dt)
                kms.append(k5_ms)
        for b, d in zip(scene.boxes, host_data):
            b.data = d
        scene.__dict__.pop("_cbox_cache", None)
        peak, _src = _peaks()
        k5 = float(np.mean(kms))
        dev_fabs = {"ms_per_step": float(np.mean(dms)), "value": total_samples / (np.mean(dms) * 1e-3) / 1e9,
                    "unit": UNIT, "scope": "FABs resident in HBM -> quantise -> frame -> RGB8 on the host",
                    "quantise": {"kernel": "brick_quantise_batch_kernel", "ms": k5, "cells": int(k5_cells),
                                 "algorithmic_bytes": 9 * int(k5_cells),
                                 "achieved_GBps": 9 * k5_cells / (k5 * 1e-3) / 1e9 if k5 > 0 else None,
                                 "frac_of_hbm_peak": (9 * k5_cells / (k5 * 1e-3) / 1e9 / peak) if k5 > 0 else None,
                                 "note": "K5: 8 B read + 1 B written per cell, one launch over all bricks, L2 "
                                         "flushed first, CUDA events on the ctx stream"}}
        del keep

    fail = False
    if rank == 0:
        peak, peak_src = _peaks()
        cells = scene.total_cells()
        frag_pixels = sum(r["fragment_pixels"] for r in per_rank)  # pieces partition the frame
        # algorithmic bytes of the march (SURVEY.md 8d PRIMARY): 8 B per brick cell (the
        # reference's FAB doubles, each needed at least once) + 20 B per fragment written
        alg_bytes = 8 * cells + 20 * frag_pixels
        march_max = max(r["march_ms"] for r in per_rank) * 1e-3
        sm_clock = (clocks["sm_mhz"] or 1965.0) * 1e6
        stage_ms = {k[:-3]: max(r[k] for r in per_rank) for k in keys}
        stage_ms["frame_minus_march"] = ms_per_step - stage_ms["march"]
        line = {
            "metric": METRIC, "value": total_samples / (ms_per_step * 1e-3) / 1e9, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {len(scene.boxes)} boxes, {cells} cells, "
                                   f"{W}x{H} AA{aa}, visibility graph on, boxes dealt "
                                   f"{DEAL_TEXT} to {world} rank(s)",
                       "value_scope": "bricks resident in HBM, RGB8 frame left on the root GPU",
                       "e2e_scope": "pinned host FABs -> upload+quantise -> frame -> RGB8 on the host"
                                    + ("; the frame marches each 16 MB copy group as it lands "
                                       "(amrvr_scene_upload_async)" if world == 1 else ""),
                       "l2": "256 MiB buffer written between timed frames (L2 flush)",
                       "placement": ("bricks replicated on every GPU at upload; march tiles dealt by image "
                                     "piece, stolen across GPUs, fragments stored into the folding GPU's pool")
                       if replicated else "one rank" if world == 1 else "boxes marched by their owner, NCCL exchange",
                       "samples_per_frame": total_samples,
                       "frames_per_s": 1e3 / ms_per_step,
                       "stage_ms": stage_ms,
                       "parity": parity["summary"] if parity else "not run (--no-parity)"},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": total_samples / (e2e_ms_mean * 1e-3) / 1e9, "unit": UNIT,
                    "ms_per_step": e2e_ms_mean, "frames_per_s": 1e3 / e2e_ms_mean,
                    "h2d_bytes_per_step": int(h2d_total), "d2h_bytes_per_step": W * H * 3,
                    # one rank: the upload call returns with the copies in flight (they end inside
                    # the frame), so the H2D rate is taken over the whole e2e time there
                    "upload_ms_max_rank": max(r["upload_ms"] for r in e2e_all),
                    "upload_call": "asynchronous (amrvr_scene_upload_async)" if world == 1 else "waited for",
                    "h2d_GBps_per_rank": [r["fab_bytes"] / ((r["e2e_ms"] if world == 1 else r["upload_ms"]) * 1e-3)
                                          / 1e9 for r in e2e_all]},
            "parity": parity,
            "roofline": {
                "kernel": "raymarch_bricks_kernel", "bound": "hbm",
                "achieved": alg_bytes / world / march_max / 1e9, "peak": peak, "unit": "GB/s",
                "frac": alg_bytes / world / march_max / 1e9 / peak,
                "traffic": _ncu_traffic()[0] if args.workload == "C4" and world == 1 else None,
                "traffic_source": f"profiles/{_ncu_traffic()[1]} (ncu --set full, C4, 1 GPU, per launch)",
                "peak_source": peak_src,
                "algorithmic_bytes": alg_bytes // world,
                "note": "PRIMARY accounting (8 B/cell + 20 B/fragment of the whole frame / GPUs, over the "
                        "slowest GPU's march); the march is issue-bound, see compute and sample_fetch",
                "march_gsamples_per_s": total_samples / march_max / 1e9,
                "sample_fetch": {"achieved_GBps": 8 * total_samples / world / march_max / 1e9,
                                 "frac_of_hbm_peak": 8 * total_samples / world / march_max / 1e9 / peak,
                                 "note": "TERTIARY, per GPU: 8 B x samples, the reference's own access "
                                         "pattern; NOT DRAM traffic"},
                "compute": {"samples_per_s_per_sm_clock": total_samples / world / march_max / (148 * sm_clock),
                            "note": "samples per SM per clock; 128 FP32 lanes/SM",
                            "issue": _ncu_issue(march_max, sm_clock)
                            if args.workload == "C4" and world == 1 else None},
                # one launch, one rank each: samples and bytes of the counting frame over ITS march time
                "per_rank_counting_frame": [
                    {"rank": c["rank"], "samples": c["samples"], "march_ms": c["march_ms"],
                     "gsamples_per_s": c["samples"] / (c["march_ms"] * 1e-3) / 1e9 if c["march_ms"] > 0 else None,
                     "sample_fetch_frac": 8 * c["samples"] / (c["march_ms"] * 1e-3) / 1e9 / peak
                     if c["march_ms"] > 0 else None,
                     "pushed_bytes": c["pushed_bytes"], "stolen_tiles": c["stolen_tiles"]} for c in counted],
            },
        }
        # the composite: fold of the own piece from local HBM; the exchange is the march's
        # stores into other GPUs' pools (stolen tiles), overlapped with the march
        fold_rows = []
        for r in per_rank:
            px = (W * H * aa) // world
            rd = 20 * r["fragment_pixels"]
            wr = (3 * px if aa == 1 else 0) + (20 * px if (world > 1 or aa > 1) else 0)
            fold_rows.append({"rank": r["rank"], "fold_ms": r["fold_ms"], "composite_ms": r["composite_ms"],
                              "fragment_bytes_read": rd, "bytes_written": wr,
                              "hbm_GBps": (rd + wr) / (r["fold_ms"] * 1e-3) / 1e9 if r["fold_ms"] > 0 else None,
                              "pushed_bytes": r["pushed_bytes"], "stolen_tiles": r["stolen_tiles"],
                              "pre_ms": r["pre_ms"], "march_ms": r["march_ms"], "march_span_ms": r["march_span_ms"],
                              "nvlink_counted_bytes_per_frame": r.get("nvlink_counted"),
                              "march_cta_resident_frac": r["march_cta_resident_frac"],
                              "march_table_ms_per_cta": r["march_table_ms_per_cta"],
                              "push_GBps": r["pushed_bytes"] / (r["march_ms"] * 1e-3) / 1e9
                              if r["march_ms"] > 0 else None})
        link_peak = None
        if link:
            link_peak = {"write_GBps_min": min(l["write_GBps"] for l in link),
                         "write_GBps_max": max(l["write_GBps"] for l in link),
                         "read_GBps_min": min(l["read_GBps"] for l in link),
                         "read_GBps_max": max(l["read_GBps"] for l in link),
                         "how": "256 MiB between ring neighbours, all ranks at once, plain 16 B loads/stores "
                                "from 148x8 CTAs, CUDA events, 10 iterations",
                         "counted_by_nvml": {"write_probe_rank0": link[0].get("counted_write_probe"),
                                             "read_probe_rank0": link[0].get("counted_read_probe"),
                                             "expected_bytes_each_way": 11 * (256 << 20),
                                             "note": "NVML NVLink throughput counters (field values 138-141, all "
                                                     "links of the GPU) around the two probes; null = NVML answers "
                                                     "NOT_SUPPORTED on this box, the rates above are CUDA-event times "
                                                     "of a copy kernel and pushed_bytes is counted by the march itself"}}
        line["composite"] = {
            "fold_ms_max_rank": max(r["fold_ms"] for r in per_rank),
            "composite_ms_max_rank": max(r["composite_ms"] for r in per_rank),
            "layers": len(scene.boxes), "per_rank": fold_rows,
            "pushed_bytes_per_rank": [int(r["pushed_bytes"]) for r in per_rank],
            "stolen_tiles_per_rank": [int(r["stolen_tiles"]) for r in per_rank],
            "tiles": per_rank[0]["tiles"],
            "nvlink_measured": link_peak, "nvlink_nominal_GBps": NVLINK_NOMINAL_GBPS,
            "nvlink_frac_of_measured_write_peak": [
                (fr["push_GBps"] / link_peak["write_GBps_min"]) if (link_peak and fr["push_GBps"] is not None)
                else None for fr in fold_rows] if link_peak else None,
            "note": "fold_ms includes the wait for the slowest GPU's march (the hand-shake is inside the fold "
                    "kernel); hbm_GBps = (20 B per stored fragment of the piece + bytes written) / fold_ms of "
                    "the same rank; push_GBps = fragment bytes this rank's march stored into other GPUs' "
                    "pools / its own march time -- the exchange rides on the march, it has no time of its own",
        }
        if seam is not None:
            peak_w = link_peak["write_GBps_min"] if link_peak else None
            seam["nvlink_measured_write_GBps"] = peak_w
            seam["frac_of_measured_write_peak_per_rank"] = (
                [g / peak_w for g in seam["sent_GBps_per_rank_over_device_ms"]] if peak_w else None)
            seam["hottest_ingress_frac_of_measured_write_peak"] = (
                max((g or 0.0) for g in seam["ingress_GBps_per_rank"]) / peak_w if peak_w else None)
            seam["push_kernel_frac_of_measured_write_peak_per_rank"] = (
                [g / peak_w if g else None for g in seam["push_GBps_per_rank"]] if peak_w else None)
            line["composite_seam"] = seam
        if dev_fabs is not None:
            line["e2e_device_fabs"] = dev_fabs
        if full_check is not None:
            line["parity_full_size"] = full_check
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(args.workload)
            except Exception as exc:  # the oracle always exists; report rather than hide
                line["cpu_baseline"] = {"error": repr(exc)}
        line["parity"] = line.pop("parity")  # last in the line: a tail of stdout still shows it
        print(json.dumps(line), flush=True)
        fail = (parity is not None and not parity["ok"]) or \
               (full_check is not None and full_check.get("ok") is False)

    dscene.free()
    ctx.close()
    lib.amrvr_host_free_pinned(pinned)
    if world > 1:
        flag = torch.tensor([1 if fail else 0], device=device)
        dist.broadcast(flag, src=0)
        fail = bool(int(flag.item()))
        dist.destroy_process_group()
    if fail:
        print("bench.py: PARITY BROKEN (see the line's parity objects)", file=sys.stderr)
        sys.exit(3)


class _DeviceArray:
    """A FAB that lives in device memory: just enough of the ndarray face for Scene.c_boxes()."""

    def __init__(self, tensor):
        self.tensor = tensor
        self.nbytes = tensor.numel() * 8
        self.shape = tuple(tensor.shape)
        self.size = tensor.numel()


def _device_boxes(scene):
    """(amrvr_box array, n) whose data pointers are device addresses."""
    arr = (api.Box * max(len(scene.boxes), 1))()
    for i, b in enumerate(scene.boxes):
        nx, ny, nz = b.dims
        cb = arr[i]
        for d in range(3):
            cb.min_corner[d] = float(b.min_corner[d])
            cb.max_corner[d] = float(b.max_corner[d])
            cb.dims[d] = int(b.dims[d])
            cb.lo[d] = int(b.lo[d])
            cb.begin[d] = int(b.lo[d])
        cb.data = b.data.tensor.data_ptr()
        cb.jstride = nx
        cb.kstride = nx * ny
        cb.nstride = nx * ny * nz
        cb.component = 0
    return arr, len(scene.boxes)


if __name__ == "__main__":
    main()
