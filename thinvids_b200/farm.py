"""Segment farm: independent per-GPU queues, no cross-GPU traffic.

The reference shards a job by time segment — `split` enqueues one `encode(job, idx, ...)` task per ~10 MB chunk
(worker/tasks.py:1097,1276) and every node pulls from the same `tasks:encode` queue with one consumer thread
(`huey_consumer tasks.encode_huey -k thread -w 1`, ansible_workers.yml:351).  Segments never exchange data.  On an
8-GPU box that becomes one consumer per GPU: each owns its scaler context, stream and buffers, and completion is a
host-side join.  There is no NCCL and no peer access on the data path.

Two forms are provided:
* `shard_segments(n, rank, world)` — the static stripe used when every GPU is its own process (torchrun, or one
  Huey consumer per GPU pinned with CUDA_VISIBLE_DEVICES);
* `SegmentFarm` — one host thread per GPU inside a single process, pulling segment indices from a shared queue
  (ctypes releases the GIL inside the C-ABI calls).
"""
from __future__ import annotations

import queue
import threading
import time
from dataclasses import dataclass
from typing import Callable, Sequence


def bind_to_gpu_cpus(device: int) -> str:
    """Restrict the calling process to the CPUs next to GPU `device` (NVML's ideal affinity), so that the pinned staging
    buffers it allocates afterwards land on that GPU's NUMA node and its copies do not cross the socket interconnect.
    One consumer per GPU should call this before it allocates anything; with a single consumer it is not needed.
    Returns a short description of what was done ("" when NVML or the topology is not available: nothing is changed)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        visible = os.environ.get("CUDA_VISIBLE_DEVICES", "")
        index = device
        if visible:
            entry = visible.split(",")[device].strip()
            if not entry.isdigit():
                return ""        # UUID lists: leave the affinity alone
            index = int(entry)
        handle = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return ""
        os.sched_setaffinity(0, cpus)
        return f"{len(cpus)} CPUs local to GPU {index}"
    except Exception:
        return ""


def shard_segments(n_segments: int, rank: int, world: int) -> list[int]:
    """Segment i goes to GPU i mod world — the same striping the reference gets from its shared queue."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    return list(range(rank, n_segments, world))


@dataclass
class SegmentResult:
    index: int
    device: int
    frames: int
    seconds: float
    output: object = None


class SegmentFarm:
    """Runs `work(scaler, segment_index)` for every segment, one worker thread and one scaler per GPU."""

    def __init__(self, make_scaler: Callable[[int], object], devices: Sequence[int]):
        self.make_scaler = make_scaler
        self.devices = list(devices)

    def run(self, n_segments: int, work: Callable[[object, int], tuple[int, object]]) -> list[SegmentResult]:
        todo: "queue.Queue[int]" = queue.Queue()
        for i in range(n_segments):
            todo.put(i)
        results: list[SegmentResult] = []
        errors: list[BaseException] = []
        lock = threading.Lock()

        def consumer(dev: int):
            try:
                scaler = self.make_scaler(dev)
                try:
                    while True:
                        try:
                            idx = todo.get_nowait()
                        except queue.Empty:
                            return
                        t0 = time.perf_counter()
                        frames, out = work(scaler, idx)
                        with lock:
                            results.append(SegmentResult(idx, dev, frames, time.perf_counter() - t0, out))
                finally:
                    scaler.close()
            except BaseException as e:  # surfaced to the caller after the join
                with lock:
                    errors.append(e)

        threads = [threading.Thread(target=consumer, args=(d,), name=f"tvs-gpu{d}") for d in self.devices]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]
        return sorted(results, key=lambda r: r.index)
