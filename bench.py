#!/usr/bin/env python
"""Benchmark of the scale / pixel-format hot path (BASELINE.json metric: frames/s 720x480 -> 1920x1080 lanczos,
% of HBM roofline).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--config C1..C5] [--impl ours|reference]

One "step" = one pass of the hot path over one batch of synthetic frames (uniform noise,
numpy default_rng(1234 + frame index)).  Our arm times K steps with the batch resident in HBM (`value`),
then the same batch through the host-buffer C-ABI call with pinned host memory, copies inside the timed
region (`e2e`).  N > 1: one process per GPU (torchrun), segments sharded across ranks, no data-path
collective; NCCL is used only for the barrier and the max-over-ranks of the timings.
`--impl reference` times the reference's own CPU implementation of the path — the genuine libswscale that
ffmpeg's `scale` filter calls (worker/tasks.py:1565-1580) — on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from thinvids_b200 import frames as F  # noqa: E402

METRIC = "frames/s 720x480->1920x1080 lanczos (scale + pixel-format stage)"   # BASELINE.json's metric, quoted on C1


def metric_name(cfg) -> str:
    if cfg.name == "C1":
        return METRIC
    return f"frames/s {cfg.src_w}x{cfg.src_h}->{cfg.dst_w}x{cfg.dst_h} {cfg.alg} (scale + pixel-format stage)"
DISTINCT_FRAMES = 8  # distinct synthetic frames generated on the host; the batch tiles them


def workload_name(cfg: F.ScaleConfig) -> str:
    return (f"{cfg.name}: {cfg.n_frames} x {cfg.src_fmt} {cfg.src_w}x{cfg.src_h} -> {cfg.dst_fmt} "
            f"{cfg.dst_w}x{cfg.dst_h} {cfg.alg}")


def synthetic_batch(cfg: F.ScaleConfig, n: int) -> np.ndarray:
    """uint8 [n, src_frame_bytes]: frame i is the SURVEY §8d frame with seed 1234 + (i mod DISTINCT_FRAMES)."""
    from thinvids_b200.swscale import PackedLayout
    lay = PackedLayout(F.fmt_code(cfg.src_fmt), cfg.src_w, cfg.src_h)
    base = lay.pack([F.random_frame(cfg.src_fmt, cfg.src_w, cfg.src_h, 1234 + i) for i in range(min(n, DISTINCT_FRAMES))])
    reps = -(-n // base.shape[0])
    return np.tile(base, (reps, 1))[:n]


# ---------------------------------------------------------------------------------------------------------
# CPU side: the reference's own code for this path (genuine libswscale), or the oracle port if it is absent.
# Used only as the reported baseline / the `--impl reference` arm, never on the product path.
# ---------------------------------------------------------------------------------------------------------
def cpu_scaler_factory(cfg: F.ScaleConfig):
    from oracle import port, ref_swscale
    args = (cfg.src_w, cfg.src_h, F.fmt_code(cfg.src_fmt), cfg.dst_w, cfg.dst_h, F.fmt_code(cfg.dst_fmt), F.ALG_BY_NAME[cfg.alg])
    if ref_swscale.available():
        return "reference", ref_swscale.version(), lambda: ref_swscale.RefScaler(*args)
    return "port", "oracle/swscale_port.c", lambda: port.PortScaler(*args)


def cpu_run(cfg: F.ScaleConfig, seconds: float, threads: int):
    """All host threads scale frames (one context per thread, frames striped) for ~`seconds`; returns frames/s."""
    kind, desc, make = cpu_scaler_factory(cfg)
    src = [F.random_frame(cfg.src_fmt, cfg.src_w, cfg.src_h, 1234 + i) for i in range(DISTINCT_FRAMES)]
    counts = [0] * threads
    start_evt = threading.Event()
    deadline = [0.0]

    def worker(t):
        sc = make()
        dst = F.alloc_frame(cfg.dst_fmt, cfg.dst_w, cfg.dst_h)
        sc.scale_into(src[t % len(src)], dst)  # warm-up, tables built
        start_evt.wait()
        i = t
        while time.perf_counter() < deadline[0]:
            sc.scale_into(src[i % len(src)], dst)  # ctypes releases the GIL
            counts[t] += 1
            i += threads
        sc.close()

    ths = [threading.Thread(target=worker, args=(t,)) for t in range(threads)]
    for t in ths:
        t.start()
    time.sleep(0.3)
    t0 = time.perf_counter()
    deadline[0] = t0 + seconds
    start_evt.set()
    for t in ths:
        t.join()
    elapsed = time.perf_counter() - t0
    total = sum(counts)
    return {"value": total / elapsed, "unit": "frames/s", "cores": threads, "kind": kind,
            "sample": f"{total} frames of {cfg.name} in {elapsed:.1f} s, {threads} threads, one SwsContext each ({desc})"}


# ---------------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
              "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.FIELDS}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def run_reference(args, cfg, rank, world):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    per_step = max(2.0, min(8.0, 40.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        cpu_run(cfg, min(1.0, per_step), threads)
    results = [cpu_run(cfg, per_step, threads) for _ in range(args.steps)]
    fps = statistics.mean(r["value"] for r in results)
    base = dict(results[-1], value=fps)
    print(json.dumps({
        "impl": "reference", "metric": metric_name(cfg), "value": fps, "unit": "frames/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u8" if F.bit_depth(cfg.src_fmt) == 8 else "u16",
        "data": "synthetic", "config": {"workload": workload_name(cfg), "threads": threads,
                                        "step": f"{per_step:.1f} s bounded sample per step"},
        "cpu_baseline": base, "gpu_launches": 0,
        "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="C1", choices=sorted(F.CONFIGS))
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--frames", type=int, default=0, help="frames per segment (default: the config's batch)")
    ap.add_argument("--segments", type=int, default=0, help="segments per step (default: 1 per GPU; C5: 64 in total)")
    ap.add_argument("--path", default="auto", choices=("auto", "fused", "two-pass", "stream", "slab"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    cfg = F.CONFIGS[args.config]
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        return run_reference(args, cfg, rank, world)

    import torch
    import torch.distributed as dist

    from thinvids_b200.swscale import Scaler

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: thinvids_b200 has no CPU path")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    n = args.frames or cfg.n_frames          # frames per segment (one segment = one call = one kernel launch)
    if cfg.name == "C5":
        # the farm job of BASELINE.json: 64 segments of 240 frames striped over the GPUs (strong scaling)
        from thinvids_b200.farm import shard_segments
        my_segments = len(shard_segments(args.segments or 64, rank, world))
        total_segments = args.segments or 64
        scaling = "strong"
    else:
        # every GPU owns its own segment(s) of the same size (weak scaling)
        my_segments = args.segments or 1
        total_segments = my_segments * world
        scaling = "weak"
    sc = Scaler(cfg.src_w, cfg.src_h, cfg.src_fmt, cfg.dst_w, cfg.dst_h, cfg.dst_fmt, cfg.alg, device=local)
    if args.path != "auto":
        sc.set_path(args.path)
    host_src = torch.from_numpy(synthetic_batch(cfg, n)).pin_memory()
    host_dst = torch.empty((n, sc.dst_layout.frame_bytes), dtype=torch.uint8).pin_memory()
    d_src = host_src.cuda()
    d_dst = torch.empty((n, sc.dst_layout.frame_bytes), dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()

    # ---- device-resident throughput (`value`) ------------------------------------------------------------
    def step():
        for _ in range(my_segments):     # segments are independent; the synthetic segment buffer is reused
            sc.scale_packed_device(d_src, d_dst, stream)

    for _ in range(args.warmup):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = sc.kernel_launches
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    evs[0].record(stream)
    for k in range(args.steps):
        step()
        evs[k + 1].record(stream)
    barrier()
    launches = sc.kernel_launches - launches0
    step_ms = [evs[k].elapsed_time(evs[k + 1]) for k in range(args.steps)]
    total_ms = max_over_ranks(evs[0].elapsed_time(evs[-1]))
    clocks = sampler.stop()
    fps = total_segments * n * args.steps / (total_ms * 1e-3)

    # ---- end to end through the host-buffer C-ABI call (`e2e`) ------------------------------------------
    e2e = None
    if not args.no_e2e:
        e2e_segments = min(my_segments, 4)               # bounded: every segment crosses PCIe both ways
        for _ in range(2):
            sc.scale_packed_pinned(host_src, host_dst)
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            for _ in range(e2e_segments):
                sc.scale_packed_pinned(host_src, host_dst)   # returns after the last byte is back in host memory
                _ = int(host_dst[-1, -1])                    # touch the result on the host
        dt = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": world * e2e_segments * n * args.steps / dt, "unit": "frames/s",
               "h2d_bytes_per_step": int(host_src.numel()) * e2e_segments, "d2h_bytes_per_step": int(host_dst.numel()) * e2e_segments,
               "ms_per_step": dt / args.steps * 1e3, "api": "tvs_scale_host (pinned host buffers)"}
        # the two entry points must agree byte for byte
        sc.scale_packed_device(d_src, d_dst, stream)
        torch.cuda.synchronize()
        if not torch.equal(d_dst[:4].cpu(), host_dst[:4]):
            raise SystemExit("host and device entry points disagree")

    # ---- roofline of the dominant kernel ----------------------------------------------------------------
    peak, peak_src = measured_peak_gbs()
    launches_per_step = launches // max(1, args.steps)
    alg_bytes = cfg.algorithmic_bytes * n                      # per launch: one kernel does one whole segment
    # one launch per segment on the fused path; the two-pass path needs several launches per segment, and the
    # figure is then the time of the whole group of launches that processes one segment
    kernel_ms = statistics.mean(step_ms) / max(1, my_segments)
    achieved = alg_bytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    prof = os.path.join(ROOT, "profiles", "traffic.json")    # written from an `ncu --set full` capture, per launch
    if os.path.exists(prof):
        try:
            traffic = json.load(open(prof)).get(f"{cfg.name}:{n}")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "peak_source": peak_src, "kernel": sc.path, "launches_per_step": launches_per_step,
                "algorithmic_bytes_per_launch": alg_bytes,
                "kernel_ms": kernel_ms}

    # ---- CPU baseline on this box's host cores (rank 0, N = 1 only) -------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_run(cfg, 10.0, os.cpu_count() or 1)

    sc.close()
    if rank == 0:
        print(json.dumps({
            "metric": metric_name(cfg), "value": fps, "unit": "frames/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": scaling, "vs_baseline": None, "dtype": "u8" if F.bit_depth(cfg.src_fmt) == 8 else "u16",
            "data": "synthetic",
            "config": {"workload": workload_name(cfg), "frames_per_segment": n, "segments_per_step": total_segments,
                       "l2": f"batch of {(cfg.algorithmic_bytes * n) >> 20} MiB per step exceeds the 126 MB L2",
                       "sharding": "independent segments per GPU, no collective on the data path"},
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
