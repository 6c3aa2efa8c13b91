#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: "Gsamples/s
ray-march + frames/s 2048^2 3-level AMR at 1/2/4/8 B200").

A step = one frame of workload C4 (synthetic 3-level AMR, 1024^3 effective, 176
boxes of 64^3, 2048x2048, visibility graph on) through
amrvr_render_frame: ray-march of this rank's bricks -> visibility order ->
sort-last composite -> gather -> byte image.  The boxes are sharded over the N
ranks (contiguous chunks of the (level,k,j,i) order); the frame is fixed, so
scaling is "strong".  `value` = ray-march samples of the whole frame / device
time per frame, bricks resident in HBM.  `e2e` = the same frame through the
public API from pinned HOST FABs: amrvr_scene_upload (H2D + quantise) +
amrvr_render_frame + D2H of the RGB8 image, all inside the timed region.

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
DEAL = "skewed"  # box i -> rank (i + i div P) mod P over the (level, k, j, i) order
NVLINK_PEAK_GBPS = 900.0  # NVLink 5, per direction per GPU (B200_PROFILING.md); no measured link figure


def _peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def _ncu_traffic(kernel_tag="march"):
    """dram__bytes_read.sum + dram__bytes_write.sum of the dominant kernel, per launch,
    from the newest committed `ncu --set full` summary under profiles/ (a capture of
    the same workload; ncu is never run inside the bench)."""
    import re

    best = None
    for f in sorted((ROOT / "profiles").glob(f"r*_{kernel_tag}_v*_full.txt")):
        best = f
    for f in sorted((ROOT / "profiles").glob(f"r*_{kernel_tag}_fold_current_full.txt")):
        best = f  # the capture of the code as committed; its first block is the march
    if best is None:
        return None, None
    unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total = 0.0
    seen = 0
    for ln in best.read_text().splitlines():
        m = re.match(r"\s*dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", ln)
        if m and seen < 2:
            total += float(m.group(2)) * unit.get(m.group(3), 1.0)
            seen += 1
    return (total if seen == 2 else None), best.name


def _ncu_issue(march_s, sm_clock_hz, n_sm=148):
    """Issue-slot use of the march at the LIVE launch duration: warp instructions
    per launch (smsp__inst_executed.sum of the committed ncu capture of this same
    workload -- an instruction count, not a timing) / live duration / (SMs x 4
    schedulers x SM clock).  None when there is no capture to read."""
    import re

    try:
        _, name = _ncu_traffic()
        if name is None or march_s <= 0 or sm_clock_hz <= 0:
            return None
        for ln in (ROOT / "profiles" / name).read_text().splitlines():
            m = re.match(r"\s*smsp__inst_executed\.sum\s+([0-9.]+)\s+inst", ln)
            if m:  # first block = the march
                inst = float(m.group(1))
                return {"warp_instructions_per_launch": inst,
                        "achieved_per_sm_clock": inst / march_s / (n_sm * sm_clock_hz),
                        "peak_per_sm_clock": 4.0,
                        "frac": inst / march_s / (n_sm * 4.0 * sm_clock_hz),
                        "source": f"profiles/{name} (instruction count, captured on the "
                                  "sample-counting instantiation: an upper bound for the timed "
                                  "one) / live launch duration"}
    except Exception:  # never let a reporting extra break the bench line
        return None
    return None


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


def reference_arm(args, workload):
    """The reference's own CPU implementation (oracle/_ref: its unmodified
    sources against the AMReX/MPI shim; ranks are threads) on a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import Oracle, Ref, RefUnavailable

    factory, W, H, aa, transp = WORKLOADS[workload]
    cores = len(os.sched_getaffinity(0))
    P = max(1, min(8, cores))
    scene = factory(nranks=P, rule=DEAL)
    camera = scenes.make_camera()
    sw, sh = 512, 512  # bounded sample: same boxes and camera, 1/16 of the rays
    params = api.FrameParams(width=sw // int(round(aa ** 0.5)), height=sh // int(round(aa ** 0.5)),
                             box_transparency=transp, antialiasing=aa, use_visibility_graph=1,
                             count_samples=1, draw_bounding_box=1)
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
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload, "sample": f"{sw}x{sh} frame of the {W}x{H} workload",
                   "boxes": len(scene.boxes), "cells": scene.total_cells()},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": P, "kind": kind,
                         "sample": f"{workload} scene, {sw}x{sh} render size, {P} ranks-as-threads, "
                                   f"full frame incl. composite; stages {stages}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def cpu_baseline(workload):
    """Bounded CPU sample for the product line (rank 0, N=1)."""
    from oracle.oracle import Oracle, Ref, RefUnavailable

    factory, W, H, aa, transp = WORKLOADS[workload]
    cores = len(os.sched_getaffinity(0))
    P = max(1, min(8, cores))
    scene = factory(nranks=P, rule=DEAL)
    sw = sh = 512
    root = int(round(aa ** 0.5))
    params = api.FrameParams(width=sw // root, height=sh // root, box_transparency=transp,
                             antialiasing=aa, use_visibility_graph=1, count_samples=1,
                             draw_bounding_box=1)
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
    return {"value": out["samples"] / dt / 1e9, "unit": UNIT, "cores": P, "kind": kind,
            "sample": f"{workload} scene at {sw}x{sh} render size ({out['samples']} samples, "
                      f"{dt:.2f} s wall, {P} ranks-as-threads); reference stage times {stages}",
            "march_only_value": (out["samples"] / stages["march_s"] / 1e9) if stages.get("march_s") else None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="C4", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
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

    factory, W, H, aa, transp = WORKLOADS[args.workload]
    # box -> rank: skewed round-robin over the (level, k, j, i) order, so every rank
    # holds the same number of boxes of every level and of every grid column
    # (contiguous chunks would give rank 0 all of level 0, i.e. 3/4 of the samples)
    scene = factory(nranks=world, rule=DEAL)
    camera = scenes.make_camera()
    lib = api.load_library()
    pinned, fab_bytes = pin_scene(lib, scene, rank, world)
    ctx = Context(device=local, rank=rank, nranks=world, nccl_id=nccl_id)
    stream = torch.cuda.ExternalStream(ctx.stream, device=torch.device("cuda", local))
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=f"cuda:{local}")  # > 126 MB L2

    params = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa,
                             use_visibility_graph=1, count_samples=0, draw_bounding_box=1)
    # device-resident arm: the RGB8 frame is built on the root GPU and stays there
    resident = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa,
                               use_visibility_graph=1, count_samples=0, draw_bounding_box=1,
                               rgb8_on_device=1)
    counting = api.FrameParams(width=W, height=H, box_transparency=transp, antialiasing=aa,
                               use_visibility_graph=1, count_samples=1, draw_bounding_box=1)

    dscene = ctx.upload_scene(scene)
    out = ctx.render_frame(dscene, camera, counting, want_rgba=False)  # untimed: count the samples
    samples = torch.tensor([out["stats"].samples], dtype=torch.int64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(samples)
    total_samples = int(samples.item())

    def one_frame():
        return ctx.render_frame(dscene, camera, resident, want_rgba=False, want_rgb8=False)

    # ---- device-resident arm: K frames, L2 flushed between frames, events on the ctx stream
    for _ in range(args.warmup):
        one_frame()
    sampler = ClockSampler(local)
    launches0 = ctx.kernel_launches
    frame_ms, march_ms = [], []
    stage = {"order": [], "composite": [], "fold": [], "gather": [], "post": []}
    barrier()
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
        march_ms.append(res["stats"].march_ms)
        stage["order"].append(res["stats"].order_ms)
        stage["composite"].append(res["stats"].composite_ms)
        stage["fold"].append(res["stats"].fold_ms)
        stage["gather"].append(res["stats"].gather_ms)
        stage["post"].append(res["stats"].post_ms)
    clocks = sampler.finish()
    launches = ctx.kernel_launches - launches0
    exchanged = res["stats"].exchanged_bytes

    t = torch.tensor([float(np.mean(frame_ms)), float(np.mean(march_ms)), float(np.mean(stage["fold"])),
                      float(np.mean(stage["composite"])), float(exchanged)], dtype=torch.float64,
                     device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step, march_mean = float(t[0].item()), float(t[1].item())
    fold_max_ms, composite_max_ms, pulled_max = float(t[2].item()), float(t[3].item()), float(t[4].item())

    # ---- end-to-end arm: pinned host FABs -> upload -> frame -> host RGB8
    e2e_ms = []
    for it in range(2 + min(args.steps, 10)):
        barrier()
        t0 = time.perf_counter()
        ds2 = ctx.upload_scene(scene)
        o2 = ctx.render_frame(ds2, camera, params, want_rgba=False)
        assert o2["rgb8"] is not None or rank != 0
        barrier()
        dt = (time.perf_counter() - t0) * 1e3
        ds2.free()
        if it >= 2:
            e2e_ms.append(dt)
    t = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64, device=f"cuda:{local}")
    fab = torch.tensor([fab_bytes], dtype=torch.int64, device=f"cuda:{local}")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(fab)
    e2e_ms_mean = float(t[0].item())

    if rank == 0:
        peak, peak_src = _peaks()
        cells = scene.total_cells()
        frag_pixels = out["stats"].samples  # placeholder replaced below
        # algorithmic bytes of the march (SURVEY.md 8d PRIMARY): 8 B per brick cell
        # (the reference's FAB doubles, each needed at least once) + 20 B per
        # fragment written.  Fragments = footprint pixels of all layers.
        lay = ctx.raymarch(dscene, camera, W * int(round(aa ** 0.5)), H * int(round(aa ** 0.5)),
                           transp, 0.001)
        frag_pixels = lay.stats().fragment_pixels
        lay.free()
        local_cells = sum(int(np.prod(b.dims)) for b in scene.boxes if world == 1 or b.owner == 0)
        alg_bytes = 8 * local_cells + 20 * frag_pixels
        march_s = march_mean * 1e-3
        sm_clock = (clocks["sm_mhz"] or 1965.0) * 1e6
        line = {
            "metric": METRIC, "value": total_samples / (ms_per_step * 1e-3) / 1e9, "unit": UNIT,
            "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_per_step,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": f"{args.workload}: {len(scene.boxes)} boxes, {cells} cells, "
                                   f"{W}x{H} AA{aa}, visibility graph on, boxes dealt "
                                   f"{DEAL} round-robin to {world} rank(s)",
                       "value_scope": "bricks resident in HBM, RGB8 frame left on the root GPU",
                       "e2e_scope": "pinned host FABs -> upload+quantise -> frame -> RGB8 on the host",
                       "l2": "256 MiB buffer written between timed frames (L2 flush)",
                       "samples_per_frame": total_samples,
                       "frames_per_s": 1e3 / ms_per_step,
                       "stage_ms": {"march": march_mean, **{k: float(np.mean(v)) for k, v in stage.items()}},
                       "exchanged_bytes_rank0": int(exchanged)},
            "clocks": clocks, "gpu_launches": int(launches),
            "e2e": {"value": total_samples / (e2e_ms_mean * 1e-3) / 1e9, "unit": UNIT,
                    "ms_per_step": e2e_ms_mean, "frames_per_s": 1e3 / e2e_ms_mean,
                    "h2d_bytes_per_step": int(fab.item()), "d2h_bytes_per_step": W * H * 3},
            "roofline": {
                "kernel": "raymarch_bricks_kernel", "bound": "hbm",
                "achieved": alg_bytes / march_s / 1e9, "peak": peak, "unit": "GB/s",
                "frac": alg_bytes / march_s / 1e9 / peak,
                "traffic": _ncu_traffic()[0] if args.workload == "C4" and world == 1 else None,
                "traffic_source": f"profiles/{_ncu_traffic()[1]} (ncu --set full, C4, 1 GPU, per launch)",
                "peak_source": peak_src,
                "algorithmic_bytes": alg_bytes,
                "note": "PRIMARY accounting (8 B/cell + 20 B/fragment); the march is issue-bound, "
                        "see compute and sample_fetch",
                "march_gsamples_per_s": (out["stats"].samples / march_s / 1e9),
                "sample_fetch": {"achieved_GBps": 8 * out["stats"].samples / march_s / 1e9,
                                 "frac_of_hbm_peak": 8 * out["stats"].samples / march_s / 1e9 / peak,
                                 "note": "TERTIARY: 8 B x samples, the reference's own access "
                                         "pattern; NOT DRAM traffic"},
                "compute": {"samples_per_s_per_sm_clock": out["stats"].samples / march_s / (148 * sm_clock),
                            "note": "samples per SM per clock; 128 FP32 lanes/SM",
                            "issue": _ncu_issue(march_s, sm_clock)
                            if args.workload == "C4" and world == 1 else None},
            },
        }
        # the composite: in peer-memory mode the fold kernel reads the other GPUs'
        # fragments over NVLink itself, so link bytes / fold time is the link rate
        frag_bytes = 20 * frag_pixels
        line["composite"] = {
            "fold_ms_max_rank": fold_max_ms, "composite_ms_max_rank": composite_max_ms,
            "layers": len(scene.boxes), "fragment_bytes_rank0": int(frag_bytes),
            "link_bytes_max_rank": int(pulled_max),
            "nvlink_GBps": (pulled_max / (fold_max_ms * 1e-3) / 1e9) if world > 1 and fold_max_ms > 0 else None,
            "nvlink_peak_GBps": NVLINK_PEAK_GBPS,
            "nvlink_frac": (pulled_max / (fold_max_ms * 1e-3) / 1e9 / NVLINK_PEAK_GBPS)
            if world > 1 and fold_max_ms > 0 else None,
            "fold_hbm_GBps": ((frag_bytes + 20 * W * H * aa) / (fold_max_ms * 1e-3) / 1e9)
            if world == 1 and fold_max_ms > 0 else None,
            "note": "fold_hbm_GBps (N=1): 20 B per stored fragment read + 20 B per pixel written; "
                    "nvlink_GBps (N>1): bytes the busiest rank pulled from peers / its fold time; "
                    "peak = NVLink 5 nominal per direction (no measured link figure on this pool)",
        }
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = cpu_baseline(args.workload)
            except Exception as exc:  # the oracle always exists; report rather than hide
                line["cpu_baseline"] = {"error": repr(exc)}
        print(json.dumps(line), flush=True)

    dscene.free()
    ctx.close()
    lib.amrvr_host_free_pinned(pinned)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
