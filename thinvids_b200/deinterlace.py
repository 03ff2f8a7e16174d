"""bwdif on the GPU: the de-interlacer the reference puts in front of ``scale`` on its DVD path
(``bwdif=mode=send_frame:parity=auto:deint=all``: worker/tasks.py:62-63,446-447; tools/upscale_benchmark.py:128-132), bound to
`tvs_bwdif_host` / `tvs_bwdif_batch` (include/tvscale.h).

PARITY UNPINNED: the image has no libavfilter to compare with.  The kernel follows the published algorithm and is tested
against a numpy restatement of it and against its invariants (tests/test_bwdif.py); by default the `encode` task keeps
ffmpeg's own bwdif in its decode child, and this stage is what `EncodeEnv.gpu_deinterlace` switches on.
Nothing here computes pixels on the CPU.
"""
from __future__ import annotations

import numpy as np

from . import frames as F
from . import lib as _lib

STREAM_START, STREAM_END = 1, 2


def bwdif_packed_host(frames: np.ndarray, width: int, height: int, fmt="yuv420p", tff: bool = True, stream_start: bool = True,
                      stream_end: bool = True, device: int = 0) -> np.ndarray:
    """`frames`: uint8 [n, frame_bytes] tightly packed planar frames in stream order.  Returns the frames that are written
    (all n, minus the first / last when they are context only), de-interlaced."""
    lib = _lib.load()
    code = F.fmt_code(fmt)
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n = frames.shape[0]
    assert frames.shape[1] == F.frame_bytes(code, width, height)
    out = np.empty_like(frames)
    flags = (STREAM_START if stream_start else 0) | (STREAM_END if stream_end else 0)
    written = _lib.check(lib.tvs_bwdif_host(device, code, width, height, frames.ctypes.data, out.ctypes.data, n, int(bool(tff)), flags))
    first = 0 if stream_start else 1
    return out[first:first + written]


class BwdifStream:
    """A stream de-interlaced batch by batch: keeps the one frame of context each side needs.  `push(batch)` returns the
    frames that can be finished so far (one frame is held back until its successor is known); `flush()` the last one."""

    def __init__(self, width: int, height: int, fmt="yuv420p", tff: bool = True, device: int = 0):
        self.args = dict(width=width, height=height, fmt=fmt, tff=tff, device=device)
        self._tail = None          # the last one or two frames seen: [predecessor of the pending frame,] pending frame
        self._started = False

    def push(self, batch: np.ndarray) -> np.ndarray:
        batch = np.ascontiguousarray(batch, dtype=np.uint8)
        if batch.shape[0] == 0:
            return batch
        work = batch if self._tail is None else np.concatenate([self._tail, batch])
        start = not self._started                      # frame 0 of `work` is the stream's first frame
        if work.shape[0] < 2:
            self._tail = work
            return work[:0]
        # the last frame has no successor yet: it is context only; so is the first one unless the stream starts here
        out = bwdif_packed_host(work, stream_start=start, stream_end=False, **self.args)
        self._started = True
        self._tail = work[-2:].copy()
        return out

    def flush(self) -> np.ndarray:
        if self._tail is None:
            return np.empty((0, 0), dtype=np.uint8)
        work, self._tail = self._tail, None
        start = not self._started
        self._started = True
        return bwdif_packed_host(work, stream_start=start, stream_end=True, **self.args)
