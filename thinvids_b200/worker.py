"""The pixel stage of the reference's `encode` task, with the scaler on the GPU.

The reference's task (worker/tasks.py:1354-1742) downloads a segment, spawns ONE ffmpeg child that decodes, filters
(``-vf setpts=...,[bwdif,]scale=-2:H[,format=nv12,hwupload]``) and encodes, watches its progress lines for halts and
heartbeats, and uploads the result.  Everything except the filter is out of scope here: Redis
bookkeeping, retries, download and upload stay where they are.  What this module mirrors is the part in between --
"build ffmpeg command" (:1532) to "Encoded part" (:1652) -- in the shape INTEGRATION.md describes: decoded frames
arrive as a raw byte stream (what ``ffmpeg ... -f rawvideo -pix_fmt yuv420p pipe:1`` writes, or any object with
``read``), are scaled in batches by `thinvids_b200.swscale.Scaler`, and leave as a raw byte stream for the encoder
(``ffmpeg -f rawvideo -pix_fmt <dst> -s WxH -i pipe:0 ...``).  Same observable behaviour as the task around it:

* the filter text comes from the same job fields through the same builders (`filtergraph.encode_filtergraph`);
* a halted job stops the part: ``{'status': 'ABORTED'}`` (worker/tasks.py:1590, 1634-1648), checked before the first
  batch and after every batch (the reference checks per ffmpeg progress line);
* the job heartbeat is fed at start and per batch with the reference's stage names ("encode_start", "encode";
  worker/tasks.py:1588, 1630) -- the 15 s throttle lives in the caller's `_job_heartbeat`, as in the reference;
* any failure is reported, not raised: ``{'status': 'FAILED', 'reason', 'stage'}``, the dict `_fail` builds
  (:1385-1464) minus the retry accounting, which is the caller's;
* success is ``{'status': 'COMPLETED', 'idx': idx, ...}`` (:1733) plus what the encoder needs to know about the frames.

There is no CPU path: without a CUDA device `Scaler` raises and the part FAILS with that reason.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import BinaryIO, Callable, Iterator, Mapping, Optional

import numpy as np

from . import frames as F
from .filtergraph import FilterGraphError, ScaleStage, encode_filtergraph
from .swscale import PackedLayout


@dataclass
class EncodeHooks:
    """The two things the reference's loop does while ffmpeg runs (worker/tasks.py:1628-1648)."""

    is_halted: Callable[[], bool] = lambda: False
    heartbeat: Callable[[str], None] = lambda stage: None


def read_frames(stream: BinaryIO, frame_bytes: int, batch_frames: int) -> Iterator[np.ndarray]:
    """Batches of whole frames, ``[n, frame_bytes]`` uint8, from a raw video byte stream; a trailing partial frame is
    an error (the decoder died mid-frame), not something to pad."""
    want = frame_bytes * batch_frames
    while True:
        buf = bytearray()
        while len(buf) < want:
            chunk = stream.read(want - len(buf))
            if not chunk:
                break
            buf += chunk
        if not buf:
            return
        if len(buf) % frame_bytes:
            raise EOFError(f"raw video stream ended inside a frame ({len(buf) % frame_bytes} of {frame_bytes} bytes)")
        yield np.frombuffer(bytes(buf), dtype=np.uint8).reshape(-1, frame_bytes)
        if len(buf) < want:
            return


def encode_part(job: Mapping, idx: int, frames_in: BinaryIO, frames_out: BinaryIO, *, src_w: int, src_h: int,
                src_fmt: str = "yuv420p", device: int = 0, hooks: Optional[EncodeHooks] = None, batch_frames: int = 64,
                deinterlaced_upstream: bool = False, scaler_factory: Optional[Callable[[ScaleStage, int], object]] = None) -> dict:
    """Scale one segment's decoded frames as the job's filter text says.  `job` is the job hash (the fields
    `_effective_target_height_for_job` and `encode` read: target_height, software_encode, filename, source_codec,
    source_resolution).  `scaler_factory(stage, device)` exists for tests of the control flow; the default builds the
    CUDA scaler."""
    hooks = hooks or EncodeHooks()
    stage_name = "encode"
    try:
        if hooks.is_halted():
            return {"status": "ABORTED"}
        vf = encode_filtergraph(job)
        try:
            # VAAPI text ends in format=nv12,hwupload; the libx264 text names no format and x264 takes yuv420p
            stage = ScaleStage.from_filtergraph(vf, src_w, src_h, src_fmt, deinterlaced_upstream=deinterlaced_upstream,
                                                default_dst_fmt="yuv420p" if int(job.get("software_encode") or 0) else None)
        except FilterGraphError as e:
            return {"status": "FAILED", "reason": f"part {idx} ({stage_name}): {e}", "stage": stage_name}
        hooks.heartbeat("encode_start")
        if hooks.is_halted():
            return {"status": "ABORTED"}

        src_layout = PackedLayout(F.FMT_BY_NAME[stage.src_fmt], stage.src_w, stage.src_h)
        n_frames = 0
        if stage.is_identity:
            # same size, same format: ffmpeg's scale filter passes frames through untouched
            for batch in read_frames(frames_in, src_layout.frame_bytes, batch_frames):
                frames_out.write(batch.tobytes())
                n_frames += len(batch)
                hooks.heartbeat("encode")
                if hooks.is_halted():
                    return {"status": "ABORTED"}
        else:
            make = scaler_factory or (lambda st, dev: st.make_scaler(device=dev))
            scaler = make(stage, device)
            try:
                for batch in read_frames(frames_in, src_layout.frame_bytes, batch_frames):
                    frames_out.write(np.ascontiguousarray(scaler.scale_packed_host(batch)).tobytes())
                    n_frames += len(batch)
                    hooks.heartbeat("encode")
                    if hooks.is_halted():
                        return {"status": "ABORTED"}
            finally:
                close = getattr(scaler, "close", None)
                if close:
                    close()
        return {"status": "COMPLETED", "idx": idx, "frames": n_frames, "filtergraph": vf,
                "dst_w": stage.dst_w, "dst_h": stage.dst_h, "dst_fmt": stage.dst_fmt}
    except Exception as e:   # the reference's catch-all (worker/tasks.py:1740)
        return {"status": "FAILED", "reason": f"part {idx} ({stage_name}): {e}", "stage": stage_name}


# ---- decode / encode ends for a box without the ffmpeg binary: OpenCV's bundled FFmpeg ----
def cv2_decode_yuv420p(path: str) -> Iterator[np.ndarray]:
    """Frames of a video file as packed yuv420p (``[h * 3 / 2 * w]`` uint8).  OpenCV hands out BGR; the conversion back
    to I420 is OpenCV's, which is fine for a smoke test of decode -> scale -> encode and nothing more."""
    import cv2
    cap = cv2.VideoCapture(path)
    if not cap.isOpened():
        raise IOError(f"cannot open {path}")
    try:
        while True:
            ok, bgr = cap.read()
            if not ok:
                return
            yield cv2.cvtColor(bgr, cv2.COLOR_BGR2YUV_I420).reshape(-1)
    finally:
        cap.release()


class Cv2Encoder:
    """File-like sink for packed yuv420p / nv12 frames: converts to BGR and feeds cv2.VideoWriter."""

    def __init__(self, path: str, w: int, h: int, fmt: str, fps: float = 24.0, fourcc: str = "mp4v"):
        import cv2
        self._cv2 = cv2
        self._w, self._h = w, h
        self._code = {"yuv420p": cv2.COLOR_YUV2BGR_I420, "nv12": cv2.COLOR_YUV2BGR_NV12}[fmt]
        self._frame_bytes = w * h * 3 // 2
        self._writer = cv2.VideoWriter(path, cv2.VideoWriter_fourcc(*fourcc), fps, (w, h))
        if not self._writer.isOpened():
            raise IOError(f"cannot open {path} for writing")
        self.frames = 0

    def write(self, data: bytes) -> int:
        arr = np.frombuffer(data, dtype=np.uint8).reshape(-1, self._h * 3 // 2, self._w)
        for yuv in arr:
            self._writer.write(self._cv2.cvtColor(yuv, self._code))
            self.frames += 1
        return len(data)

    def close(self) -> None:
        self._writer.release()
