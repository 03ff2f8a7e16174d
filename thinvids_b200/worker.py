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

import ctypes
import os
import queue
import subprocess
import threading
from collections import deque
from dataclasses import dataclass, field
from typing import BinaryIO, Callable, Iterator, Mapping, Optional

import numpy as np

from . import frames as F
from .filtergraph import (FilterGraphError, ScaleStage, effective_target_height_for_job, encode_filtergraph,
                          parse_filtergraph)
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


class PinnedRing:
    """A few page-locked host buffers (tvs_host_alloc) of one batch each, source and destination side.  Raw video is
    read straight into a source buffer and written straight from a destination buffer, so a frame crosses host memory
    once in each direction and the host entry point can overlap its PCIe copies with its kernels (pageable memory makes
    the driver stage every copy)."""

    def __init__(self, slots: int, src_batch_bytes: int, dst_batch_bytes: int):
        from . import lib as _lib
        self._lib = _lib.load()
        self._ptrs = []
        self.src, self.dst = [], []
        try:
            for _ in range(slots):
                for nbytes, out in ((src_batch_bytes, self.src), (dst_batch_bytes, self.dst)):
                    p = self._lib.tvs_host_alloc(nbytes)
                    if not p:
                        raise MemoryError(_lib.last_error())
                    self._ptrs.append(p)
                    out.append(np.ctypeslib.as_array((ctypes.c_uint8 * nbytes).from_address(p)))
        except Exception:
            self.close()
            raise

    def close(self) -> None:
        self.src, self.dst = [], []
        for p in self._ptrs:
            self._lib.tvs_host_free(p)
        self._ptrs = []


def _readinto_exact(stream: BinaryIO, view: memoryview) -> int:
    """Fill `view` from the stream; returns the bytes read (less than len(view) only at end of stream)."""
    got = 0
    readinto = getattr(stream, "readinto", None)
    while got < len(view):
        if readinto is not None:
            n = readinto(view[got:])
        else:
            chunk = stream.read(len(view) - got)
            n = len(chunk)
            view[got:got + n] = chunk
        if not n:
            break
        got += n
    return got


def _scale_stream_pinned(scaler, frames_in: BinaryIO, frames_out: BinaryIO, src_frame: int, dst_frame: int, batch_frames: int,
                         hooks: "EncodeHooks", slots: int = 3):
    """read -> scale -> write as three overlapped stages over a ring of pinned batches: a reader thread fills source
    buffers, this thread scales them on the GPU (tvs_scale_host releases the GIL), a writer thread drains destination
    buffers.  Returns (frames, aborted)."""
    ring = PinnedRing(slots, src_frame * batch_frames, dst_frame * batch_frames)
    free_src, full_src = queue.Queue(), queue.Queue()
    free_dst, full_dst = queue.Queue(), queue.Queue()
    for i in range(slots):
        free_src.put(i)
        free_dst.put(i)
    errors = []
    stop = threading.Event()

    def reader():
        try:
            while not stop.is_set():
                i = free_src.get()
                if i is None:
                    return
                got = _readinto_exact(frames_in, memoryview(ring.src[i]))
                if got % src_frame:
                    raise EOFError(f"raw video stream ended inside a frame ({got % src_frame} of {src_frame} bytes)")
                full_src.put((i, got // src_frame))
                if got < len(ring.src[i]):
                    return
        except Exception as e:   # surfaced by the scaling thread
            errors.append(e)
        finally:
            full_src.put(None)

    def writer():
        try:
            while True:
                item = full_dst.get()
                if item is None:
                    return
                i, n = item
                frames_out.write(memoryview(ring.dst[i])[:n * dst_frame])
                free_dst.put(i)
        except Exception as e:
            errors.append(e)
            stop.set()
            while True:   # keep the scaling thread from blocking on a dead sink: hand its buffers straight back
                item = full_dst.get()
                if item is None:
                    return
                free_dst.put(item[0])

    tr, tw = threading.Thread(target=reader, daemon=True), threading.Thread(target=writer, daemon=True)
    tr.start()
    tw.start()
    n_frames, aborted = 0, False
    try:
        while True:
            item = full_src.get()
            if item is None or errors:
                break
            i, n = item
            if n:
                j = free_dst.get()
                scaler._scale_host_ptr(ring.src[i].ctypes.data, ring.dst[j].ctypes.data, n)
                full_dst.put((j, n))
                n_frames += n
            free_src.put(i)
            hooks.heartbeat("encode")
            if hooks.is_halted():
                aborted = True
                break
    finally:
        stop.set()
        free_src.put(None)
        full_dst.put(None)
        tw.join()
        tr.join(timeout=5)
        ring.close()
    if errors:
        raise errors[0]
    return n_frames, aborted


class _GpuDeinterlacedReader:
    """File-like view of a raw video stream with bwdif applied on the GPU (thinvids_b200.deinterlace; parity unpinned):
    what the scaler reads when the job's graph de-interlaces and nothing upstream has done it."""

    def __init__(self, stream: BinaryIO, width: int, height: int, fmt: str, tff: bool, device: int, batch_frames: int):
        from .deinterlace import BwdifStream
        self._in, self._frame = stream, F.frame_bytes(F.FMT_BY_NAME[fmt], width, height)
        self._batch = max(2, batch_frames)
        self._bw = BwdifStream(width, height, fmt, tff, device)
        self._buf, self._eof = b"", False

    def read(self, n: int = -1) -> bytes:
        while not self._eof and (n < 0 or len(self._buf) < n):
            raw = bytearray(self._frame * self._batch)
            got = _readinto_exact(self._in, memoryview(raw))
            if got % self._frame:
                raise EOFError(f"raw video stream ended inside a frame ({got % self._frame} of {self._frame} bytes)")
            if got:
                out = self._bw.push(np.frombuffer(raw, dtype=np.uint8, count=got).reshape(-1, self._frame))
                self._buf += out.tobytes()
            if got < len(raw):
                self._eof = True
                self._buf += self._bw.flush().tobytes()
        if n < 0:
            n = len(self._buf)
        out, self._buf = self._buf[:n], self._buf[n:]
        return out


def encode_part(job: Mapping, idx: int, frames_in: BinaryIO, frames_out: BinaryIO, *, src_w: int, src_h: int,
                src_fmt: str = "yuv420p", device: int = 0, hooks: Optional[EncodeHooks] = None, batch_frames: int = 64,
                deinterlaced_upstream: bool = False, scaler_factory: Optional[Callable[[ScaleStage, int], object]] = None,
                gpu_deinterlace: bool = False, tff: bool = True) -> dict:
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
            stage = ScaleStage.from_filtergraph(vf, src_w, src_h, src_fmt, deinterlaced_upstream=deinterlaced_upstream or gpu_deinterlace,
                                                default_dst_fmt="yuv420p" if int(job.get("software_encode") or 0) else None)
        except FilterGraphError as e:
            return {"status": "FAILED", "reason": f"part {idx} ({stage_name}): {e}", "stage": stage_name}
        hooks.heartbeat("encode_start")
        if hooks.is_halted():
            return {"status": "ABORTED"}
        if stage.deinterlace and gpu_deinterlace and not deinterlaced_upstream:
            # the graph asks for bwdif and nothing upstream has run it: run it here, on the GPU, ahead of the scaler
            frames_in = _GpuDeinterlacedReader(frames_in, src_w, src_h, stage.src_fmt, tff, device, batch_frames)

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
                if hasattr(scaler, "_scale_host_ptr"):
                    # the CUDA scaler: pinned ring, read / scale / write overlapped
                    dst_frame = F.frame_bytes(F.FMT_BY_NAME[stage.dst_fmt], stage.dst_w, stage.dst_h)
                    n_frames, aborted = _scale_stream_pinned(scaler, frames_in, frames_out, src_layout.frame_bytes, dst_frame,
                                                             batch_frames, hooks)
                    if aborted:
                        return {"status": "ABORTED"}
                else:
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


# ---- the task itself, with the reference's signature -----------------------------------------------------------------
# Environment knobs of worker/tasks.py:56,66-68, same names and defaults.
VAAPI_DEVICE = os.environ.get("VAAPI_DEVICE", "/dev/dri/renderD128")
VAAPI_RC_MODE = os.environ.get("VEM_RC_MODE", "CQP")
VAAPI_QP = os.environ.get("VEM_QP", "27")
AUDIO_ARGS = os.environ.get("VEM_AUDIO_ARGS", "-c:a aac -ac 2 -b:a 192k")
_FFMPEG_HEAD = ["ffmpeg", "-hide_banner", "-nostats", "-loglevel", "error"]


@dataclass
class EncodeEnv:
    """Everything `encode` touches outside the pixel path.  In the reference these are module globals bound to Redis,
    HTTP and the local tmpfs (worker/tasks.py:29-39, 106, 270, 392, 396); they stay out of scope here
    and arrive as callables, so the deployment binds them to its own Redis/HTTP helpers and the tests bind fakes.

    job_fields(job_id) -> the job hash as a mapping      (redis.hgetall(_job_key(job_id)))
    set_job_fields(job_id, mapping)                       (redis.hset(job_key, mapping=...), :1537)
    is_halted(job_id) -> bool                             (_is_job_halted, :392)
    heartbeat(job_id, stage, force=False, note="")        (_job_heartbeat, :106)
    token_is_current(job_id, run_token) -> bool           (_task_token_is_current, :396)
    base_dir(job_id) -> str                               (_job_base_dir, :270; created if missing)
    fetch_part(job_id, idx, master_host, in_path)         (the GET of :1498-1523; raises on failure)
    upload_result(job_id, idx, stitch_host, out_path)     (the PUT of :1657-1674 and the counters after it)
    probe(path) -> (width, height, pix_fmt, fps)          (ffprobe: _ffprobe_video_stream, :190-268)
    popen(argv, **kw)                                     (subprocess.Popen)
    """

    job_fields: Callable[[str], Mapping]
    probe: Callable[[str], tuple]
    set_job_fields: Callable[[str, Mapping], None] = lambda job_id, mapping: None
    is_halted: Callable[[str], bool] = lambda job_id: False
    heartbeat: Callable[..., None] = lambda job_id, stage, force=False, note="": None
    token_is_current: Callable[[str, Optional[str]], bool] = lambda job_id, run_token: True
    base_dir: Callable[[str], str] = lambda job_id: os.path.join("/tmp", "thinvids", str(job_id))
    fetch_part: Optional[Callable[[str, int, str, str], None]] = None
    upload_result: Optional[Callable[[str, int, str, str], None]] = None
    popen: Callable[..., object] = subprocess.Popen
    device: int = 0
    batch_frames: int = 32
    scaler_factory: Optional[Callable[[ScaleStage, int], object]] = None
    gpu_deinterlace: bool = False    # run the graph's bwdif on the GPU (parity unpinned) instead of in the decode child
    top_field_first: bool = True     # field order handed to the GPU de-interlacer (raw video carries no flags)
    last_commands: dict = field(default_factory=dict)     # the two argv lists of the last part, for logs and tests


_env: Optional[EncodeEnv] = None


def configure(env: EncodeEnv) -> None:
    """Bind `encode` to a deployment (or a test double)."""
    global _env
    _env = env


def build_commands(job: Mapping, in_path: str, out_path: str, stage: ScaleStage, fps: float, *, v_sel: int = 0, a_sel: int = 0,
                   source_path: Optional[str] = None, source_start_s: Optional[float] = None,
                   source_duration_s: Optional[float] = None, deinterlace_in_decoder: bool = True) -> tuple:
    """The reference's single ffmpeg command (worker/tasks.py:1558-1586) cut in two around the scaler.

    decode:  same `*input_args`, the selected video stream only, `-vf setpts=PTS-STARTPTS[,bwdif=...]` (the part of the
             graph that is not this stage's: the PTS reset and, on the DVD path, ffmpeg's own de-interlacer), raw frames
             in the source pixel format on stdout;
    encode:  raw frames of the scaled size / format on stdin as input 0, the segment again as input 1 for its audio,
             the reference's encoder arguments unchanged (`libx264 -preset veryfast -crf 23`, or `h264_vaapi` in CQP mode
             behind `hwupload`: the frames already are nv12), `AUDIO_ARGS`, `-progress pipe:2`, same output path.
    """
    direct_source = bool(source_path) and source_start_s is not None and source_duration_s is not None
    input_args = ["-fflags", "+genpts", "-i", in_path]
    audio_map = ["-map", "1:a?"]
    video_sel = "0:v:0"
    extra_maps = []
    if direct_source:
        input_args = ["-ss", f"{float(source_start_s or 0):.6f}", "-i", source_path,
                      "-t", f"{max(0.05, float(source_duration_s or 0.0)):.6f}"]
        video_sel = f"0:v:{int(v_sel)}"
        audio_map = ["-map", f"1:a:{int(a_sel)}?"]
        extra_maps = ["-sn", "-dn", "-map_metadata", "-1", "-map_chapters", "-1"]
    upstream = [f for f in parse_filtergraph(encode_filtergraph(job))
                if f.name == "setpts" or (f.name == "bwdif" and deinterlace_in_decoder)]
    upstream_text = ",".join(_filter_text(f) for f in upstream)
    decode = [*_FFMPEG_HEAD, *input_args, "-map", video_sel, "-an", "-sn", "-dn", "-vf", upstream_text,
              "-f", "rawvideo", *stage.src_raw_args, "pipe:1"]
    raw_in = ["-f", "rawvideo", *stage.dst_raw_args, "-s", f"{stage.dst_w}x{stage.dst_h}", "-r", f"{fps:.6f}", "-i", "pipe:0"]
    if int(job.get("software_encode") or 0):
        encode = [*_FFMPEG_HEAD, "-y", "-progress", "pipe:2", *raw_in, *input_args, "-map", "0:v:0", *audio_map, *extra_maps,
                  "-c:v", "libx264", "-preset", "veryfast", "-crf", "23", *AUDIO_ARGS.split(), out_path]
    else:
        encode = [*_FFMPEG_HEAD, "-y", "-progress", "pipe:2", "-vaapi_device", VAAPI_DEVICE, *raw_in, *input_args,
                  "-map", "0:v:0", *audio_map, *extra_maps, "-vf", "hwupload", "-c:v", "h264_vaapi", "-rc_mode", VAAPI_RC_MODE,
                  "-qp", VAAPI_QP, *AUDIO_ARGS.split(), out_path]
    return decode, encode


def _filter_text(f) -> str:
    parts = list(f.args) + [f"{k}={v}" for k, v in f.kwargs.items()]
    return f.name + ("=" + ":".join(parts) if parts else "")


def encode(job_id: str, idx: int, master_host: str, v_sel: int = 0, a_sel: int = 0, stitch_host: Optional[str] = None,
           source_path: Optional[str] = None, source_start_s: Optional[float] = None, source_duration_s: Optional[float] = None,
           run_token: Optional[str] = None):
    """The reference's Huey task, same arguments and same result dicts (worker/tasks.py:1354-1366, 1380, 1469, 1733;
    decorate it with `@encode_huey.task()` in the deployment).  One ffmpeg child decodes the part to raw frames, the
    scaler runs on the GPU, a second ffmpeg child encodes; halts, heartbeats and return codes are handled as at
    :1588-1651.  Needs `configure(EncodeEnv(...))` first."""
    env = _env
    if env is None:
        raise RuntimeError("thinvids_b200.worker.encode: call configure(EncodeEnv(...)) first")
    if not env.token_is_current(job_id, run_token):
        return {"status": "STALE", "job_id": job_id, "idx": idx}
    stage_name = "prepare"

    def _fail(reason: str, stage: str):
        # the reference re-queues the part here up to PART_FAILURE_MAX_RETRIES (:1385-1464); that accounting is the
        # deployment's (it lives in Redis) - this returns the terminal dict
        return {"status": "FAILED", "reason": f"part {idx} ({stage}): {reason}", "stage": stage}

    pr_dec = pr_enc = None
    try:
        if env.is_halted(job_id):
            return {"status": "ABORTED"}
        env.heartbeat(job_id, "encode_prepare", force=True, note=f"part={idx}")
        stitch = stitch_host or master_host
        wtmp = env.base_dir(job_id)
        os.makedirs(wtmp, exist_ok=True)
        in_path = os.path.join(wtmp, f"in_{idx:03d}.ts")
        out_path = os.path.join(wtmp, f"out_{idx:03d}.mp4")
        direct_source = bool(source_path) and source_start_s is not None and source_duration_s is not None
        if not direct_source:
            stage_name = "download"
            if env.is_halted(job_id):
                return {"status": "ABORTED"}
            if env.fetch_part is not None:
                try:
                    env.fetch_part(job_id, idx, master_host, in_path)
                except Exception as e:
                    return _fail(f"download failed: {e}", stage_name)

        stage_name = "encode"
        job = dict(env.job_fields(job_id))
        target_height, force_deinterlace = effective_target_height_for_job(job)
        try:
            env.set_job_fields(job_id, {"target_height_effective": target_height,
                                        "deinterlace_effective": "1" if force_deinterlace else "0"})
        except Exception:
            pass
        src_w, src_h, src_fmt, fps = env.probe(source_path if direct_source else in_path)
        software = int(job.get("software_encode") or 0)
        try:
            # bwdif, when the graph has it, runs in the decode child: frames reach this stage progressive
            stage = ScaleStage.from_filtergraph(encode_filtergraph(job), src_w, src_h, src_fmt, deinterlaced_upstream=True,
                                                default_dst_fmt="yuv420p" if software else None)
        except FilterGraphError as e:
            return _fail(str(e), stage_name)
        decode_cmd, encode_cmd = build_commands(job, in_path, out_path, stage, fps, v_sel=v_sel, a_sel=a_sel, source_path=source_path,
                                                source_start_s=source_start_s, source_duration_s=source_duration_s,
                                                deinterlace_in_decoder=not env.gpu_deinterlace)
        env.last_commands = {"decode": decode_cmd, "encode": encode_cmd}
        env.heartbeat(job_id, "encode_start", force=True, note=f"part={idx}")
        if env.is_halted(job_id):
            return {"status": "ABORTED"}

        stderr_tail = deque(maxlen=120)
        try:
            pr_dec = env.popen(decode_cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL)
            pr_enc = env.popen(encode_cmd, stdin=subprocess.PIPE, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)
        except Exception as e:
            return _fail(f"spawn ffmpeg failed: {e}", stage_name)

        def drain():   # progress lines keep ffmpeg responsive and feed the heartbeat, as at :1624-1631
            try:
                for raw in pr_enc.stderr:
                    line = raw.decode("utf-8", "replace").rstrip("\n") if isinstance(raw, bytes) else raw.rstrip("\n")
                    env.heartbeat(job_id, "encode", note=f"part={idx}")
                    if line:
                        stderr_tail.append(line)
            except Exception:
                pass

        t_drain = threading.Thread(target=drain, daemon=True)
        if getattr(pr_enc, "stderr", None) is not None:
            t_drain.start()
        hooks = EncodeHooks(is_halted=lambda: env.is_halted(job_id),
                            heartbeat=lambda st: env.heartbeat(job_id, st, note=f"part={idx}"))
        res = encode_part(job, idx, pr_dec.stdout, pr_enc.stdin, src_w=src_w, src_h=src_h, src_fmt=src_fmt, device=env.device,
                          hooks=hooks, batch_frames=env.batch_frames, deinterlaced_upstream=not env.gpu_deinterlace,
                          scaler_factory=env.scaler_factory, gpu_deinterlace=env.gpu_deinterlace, tff=env.top_field_first)
        try:
            pr_enc.stdin.close()
        except Exception:
            pass
        if res.get("status") != "COMPLETED":
            for pr in (pr_dec, pr_enc):   # halt or failure: terminate both children, remove the partial output (:1634-1648)
                try:
                    pr.terminate()
                    try:
                        pr.wait(timeout=5)
                    except subprocess.TimeoutExpired:
                        pr.kill()
                except Exception:
                    pass
            try:
                if os.path.exists(out_path):
                    os.remove(out_path)
            except Exception:
                pass
            return res if res.get("status") == "ABORTED" else _fail(res.get("reason", "scale stage failed"), stage_name)
        rc_dec, rc_enc = pr_dec.wait(), pr_enc.wait()
        if t_drain.is_alive():
            t_drain.join(timeout=5)
        if rc_dec != 0 or rc_enc != 0 or not os.path.exists(out_path):
            tail = "\n".join(list(stderr_tail)[-40:])
            return _fail(f"ffmpeg rc={rc_enc} (decoder rc={rc_dec}); tail:\n{tail}", stage_name)

        stage_name = "upload"
        env.heartbeat(job_id, "upload_start", force=True, note=f"part={idx}")
        if env.upload_result is not None:
            try:
                env.upload_result(job_id, idx, stitch, out_path)
            except Exception as e:
                return _fail(f"upload failed: {e}", stage_name)
        return {"status": "COMPLETED", "idx": idx}
    except Exception as e:   # the reference's catch-all (:1736-1740)
        return _fail(str(e), stage_name)


# ---- decoder surfaces in place (north_star: "NVDEC decoded surfaces are consumed in place") --------------------------
@dataclass(frozen=True)
class DecoderSurface:
    """One mapped NVDEC output surface (what cuvidMapVideoFrame64 returns: a device pointer and a pitch): luma rows first,
    the interleaved chroma plane `pitch x aligned_height` bytes further down, NV12 for 8-bit streams and P016 - whose
    upper 10 bits are P010 - for 10-bit ones.  `frame_stride` is the distance between consecutive surfaces when a batch
    of them sits in one allocation (0 for a single surface)."""

    dev_ptr: int
    pitch: int
    width: int
    height: int
    aligned_height: int = 0          # coded height of the surface (defaults to height rounded up to 2)
    bit_depth: int = 8
    frame_stride: int = 0

    @property
    def fmt(self) -> str:
        return "nv12" if self.bit_depth == 8 else "p010le"

    @property
    def chroma_offset(self) -> int:
        return self.pitch * (self.aligned_height or (self.height + 1) // 2 * 2)

    def plane_args(self):
        """(pointers, pitches, frame strides) in the form Scaler.scale_device_planes takes."""
        return [self.dev_ptr, self.dev_ptr + self.chroma_offset], [self.pitch, self.pitch], [self.frame_stride, self.frame_stride]


def scale_surfaces(scaler, surface: DecoderSurface, dst_ptrs, dst_pitch, dst_fstride, n_frames: int = 1, stream: int = 0) -> None:
    """tvs_scale_batch straight from decoder surfaces: no copy, no repack; the kernels read the pitched NV12 / P016 rows
    where the decoder left them.  `scaler` must have been created for `surface.fmt` at `surface.width x surface.height`."""
    if (scaler.src_w, scaler.src_h) != (surface.width, surface.height) or F.NAME_BY_FMT[scaler.src_fmt] != surface.fmt:
        raise ValueError("scaler geometry / format does not match the decoder surface")
    src_ptrs, src_pitch, src_fstride = surface.plane_args()
    scaler.scale_device_planes(src_ptrs, src_pitch, src_fstride, dst_ptrs, dst_pitch, dst_fstride, n_frames, stream)


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
