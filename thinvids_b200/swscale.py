"""Host-side mirror of the libswscale trio the reference's ffmpeg child calls for ``scale=...``
(sws_getContext / sws_scale / sws_freeContext; reached from worker/tasks.py:1565-1580 and
tools/upscale_benchmark.py:163,182), bound to the sm_100a kernels through the C ABI of
include/tvscale.h.  Same argument meaning and error behaviour: ``Scaler(...)`` raises where
sws_getContext would return NULL, ``scale`` returns the destination planes and raises on a negative
return code.  Nothing here computes pixels on the CPU.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import numpy as np

from . import frames as F
from . import lib as _lib

PtrArr3 = ctypes.c_void_p * 3
DiffArr3 = ctypes.c_ssize_t * 3


@dataclass(frozen=True)
class PackedLayout:
    """Tightly packed frame: planes back to back, frames back to back (the layout of SURVEY.md §8d)."""

    fmt: int
    width: int
    height: int

    @property
    def planes(self):
        bps = 1 if F.bit_depth(self.fmt) == 8 else 2
        out, off = [], 0
        for rows, cols in F.plane_shapes(self.fmt, self.width, self.height):
            out.append((off, rows, cols * bps))  # (byte offset, rows, row bytes)
            off += rows * cols * bps
        return out

    @property
    def frame_bytes(self) -> int:
        return F.frame_bytes(self.fmt, self.width, self.height)

    def pack(self, frames) -> np.ndarray:
        """List of frames (each a list of numpy planes) -> uint8 array [n, frame_bytes]."""
        out = np.empty((len(frames), self.frame_bytes), dtype=np.uint8)
        for i, fr in enumerate(frames):
            for (off, rows, rb), p in zip(self.planes, fr):
                out[i, off:off + rows * rb] = np.ascontiguousarray(p).view(np.uint8).reshape(-1)
        return out

    def unpack(self, packed: np.ndarray):
        dt = F.sample_dtype(self.fmt)
        frames = []
        for i in range(packed.shape[0]):
            fr = []
            for (off, rows, rb), (r, c) in zip(self.planes, F.plane_shapes(self.fmt, self.width, self.height)):
                fr.append(packed[i, off:off + rows * rb].view(dt).reshape(r, c).copy())
            frames.append(fr)
        return frames

    def pointer_args(self, base_ptr: int):
        pl = self.planes
        ptrs = PtrArr3(*([base_ptr + off for off, _, _ in pl] + [None] * (3 - len(pl))))
        pitch = DiffArr3(*([rb for _, _, rb in pl] + [0] * (3 - len(pl))))
        fstride = DiffArr3(*([self.frame_bytes] * 3))
        return ptrs, pitch, fstride


class Scaler:
    """One scaling context on one GPU (the SwsContext of this package)."""

    def __init__(self, src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, flags=F.SWS_BICUBIC, chr_pos=None, device=0,
                 src_range=None, dst_range=None, params=None):
        """`src_range` / `dst_range`: "full" (JPEG) or "limited" (MPEG, the default), ffmpeg's in_range / out_range.  The
        deprecated format name "yuvj420p" is yuv420p with full range on that side (include/tvscale.h TVS_FMT_YUVJ420P).
        `params`: (param0, param1) of ffmpeg's scale filter - bicubic B and C, lanczos lobes - None entries = defaults."""
        self._lib = _lib.load()
        if src_fmt == "yuvj420p":
            src_fmt, src_range = "yuv420p", "full"
        if dst_fmt == "yuvj420p":
            dst_fmt, dst_range = "yuv420p", "full"
        self.src_fmt, self.dst_fmt = F.fmt_code(src_fmt), F.fmt_code(dst_fmt)
        self.src_w, self.src_h, self.dst_w, self.dst_h = src_w, src_h, dst_w, dst_h
        self.flags = F.ALG_BY_NAME[flags] if isinstance(flags, str) else int(flags)
        for side, bit in ((src_range, _lib.SRC_FULL_RANGE), (dst_range, _lib.DST_FULL_RANGE)):
            if side not in (None, "limited", "mpeg", "tv", "full", "jpeg", "pc", 0, 1):
                raise ValueError(f"unknown range {side!r}")
            if side in ("full", "jpeg", "pc", 1):
                self.flags |= bit
        self.device = device
        cp = (ctypes.c_int * 4)(*chr_pos) if chr_pos is not None else None
        pr = None
        if params is not None and any(v is not None for v in params):
            pr = (ctypes.c_double * 2)(*[_lib.PARAM_DEFAULT if v is None else float(v) for v in (tuple(params) + (None, None))[:2]])
        self._ctx = self._lib.tvs_create_ex(device, src_w, src_h, self.src_fmt, dst_w, dst_h, self.dst_fmt, self.flags, cp, pr)
        if not self._ctx:
            raise _lib.TvscaleError(-1, _lib.last_error())
        self.src_layout = PackedLayout(self.src_fmt, src_w, src_h)
        self.dst_layout = PackedLayout(self.dst_fmt, dst_w, dst_h)

    # -- introspection ---------------------------------------------------------------------------
    def table(self, which: int):
        taps = _lib.check(self._lib.tvs_context_table(self._ctx, which, None, 0, None, 0))
        n = (self.dst_w, (self.dst_w + 1) // 2, self.dst_h, (self.dst_h + 1) // 2)[which]
        coef = np.empty((n, taps), dtype=np.int16)
        pos = np.empty(n, dtype=np.int32)
        _lib.check(self._lib.tvs_context_table(self._ctx, which, coef.ctypes.data, coef.size, pos.ctypes.data, pos.size))
        return coef, pos

    def info(self, what: int) -> int:
        return _lib.check(self._lib.tvs_info(self._ctx, what))

    @property
    def kernel_launches(self) -> int:
        return self.info(_lib.INFO_KERNEL_LAUNCHES)

    @property
    def path(self) -> str:
        return ("fused", "two-pass", "repack", "stream", "box2", "slab", "cascade")[self.info(_lib.INFO_PATH)]

    def set_path(self, path) -> None:
        code = {"auto": -1, "fused": 0, "two-pass": 1, "stream": 2, "slab": 3}.get(path, path)
        _lib.check(self._lib.tvs_set_path(self._ctx, code))

    # -- sws_scale on host planes ------------------------------------------------------------------
    def scale(self, src_planes, dst_planes=None):
        """sws_scale(): one frame, host planes in, host planes out (any row strides)."""
        if dst_planes is None:
            dst_planes = F.alloc_frame(self.dst_fmt, self.dst_w, self.dst_h)
        n_s, n_d = len(src_planes), len(dst_planes)
        sp = PtrArr3(*([p.ctypes.data for p in src_planes] + [None] * (3 - n_s)))
        ss = DiffArr3(*([p.strides[0] for p in src_planes] + [0] * (3 - n_s)))
        dp = PtrArr3(*([p.ctypes.data for p in dst_planes] + [None] * (3 - n_d)))
        ds = DiffArr3(*([p.strides[0] for p in dst_planes] + [0] * (3 - n_d)))
        zero = DiffArr3(0, 0, 0)
        rows = _lib.check(self._lib.tvs_scale_host(self._ctx, sp, ss, zero, dp, ds, zero, 1))
        assert rows == self.dst_h
        return dst_planes

    def scale_packed_host(self, src: np.ndarray, dst: np.ndarray | None = None) -> np.ndarray:
        """Batch of tightly packed frames in host memory: uint8 [n, src_frame_bytes] -> [n, dst_frame_bytes]."""
        n = src.shape[0]
        assert src.dtype == np.uint8 and src.flags.c_contiguous and src.shape[1] == self.src_layout.frame_bytes
        if dst is None:
            dst = np.empty((n, self.dst_layout.frame_bytes), dtype=np.uint8)
        return self._scale_host_ptr(src.ctypes.data, dst.ctypes.data, n, dst)

    def _scale_host_ptr(self, src_ptr: int, dst_ptr: int, n: int, ret=None):
        sp, ss, sf = self.src_layout.pointer_args(src_ptr)
        dp, ds, df = self.dst_layout.pointer_args(dst_ptr)
        _lib.check(self._lib.tvs_scale_host(self._ctx, sp, ss, sf, dp, ds, df, n))
        return ret

    def scale_packed_pinned(self, src, dst):
        """Same, for torch CPU (ideally pinned) uint8 tensors [n, frame_bytes]; returns dst."""
        n = src.shape[0]
        assert src.is_contiguous() and dst.is_contiguous()
        assert src.shape[1] == self.src_layout.frame_bytes and dst.shape[1] == self.dst_layout.frame_bytes
        return self._scale_host_ptr(src.data_ptr(), dst.data_ptr(), n, dst)

    # -- device-resident batches -------------------------------------------------------------------
    def scale_packed_device(self, src, dst, stream=None):
        """torch CUDA uint8 tensors [n, frame_bytes] already in HBM; asynchronous on `stream`
        (a torch.cuda.Stream, default: the current stream)."""
        import torch

        assert src.is_cuda and dst.is_cuda and src.dtype == torch.uint8 and dst.dtype == torch.uint8
        assert src.is_contiguous() and dst.is_contiguous() and src.device.index == self.device
        n = src.shape[0]
        assert src.shape[1] == self.src_layout.frame_bytes and dst.shape == (n, self.dst_layout.frame_bytes)
        st = (stream or torch.cuda.current_stream(src.device)).cuda_stream
        sp, ss, sf = self.src_layout.pointer_args(src.data_ptr())
        dp, ds, df = self.dst_layout.pointer_args(dst.data_ptr())
        _lib.check(self._lib.tvs_scale_batch(self._ctx, sp, ss, sf, dp, ds, df, n, ctypes.c_void_p(st)))
        return dst

    def scale_packed_device_static(self, src, dst, stream=None):
        """`scale_packed_device` through tvs_scale_batch_static: the variant for callers that cannot allocate device
        memory while scaling (a GPU whose memory is already fully handed out, or a frame loop that must not take the
        device-wide stall of cudaMalloc).  Runs the two-pass kernels with their intermediate in the library's fixed
        array: at most `lib.STATIC_SCRATCH_BYTES` of intermediate per call (TvscaleError -E2BIG beyond that), one array
        per GPU shared by all calls, so keep these calls on one stream.  Output bytes equal `scale_packed_device`."""
        import torch

        assert src.is_cuda and dst.is_cuda and src.dtype == torch.uint8 and dst.dtype == torch.uint8
        assert src.is_contiguous() and dst.is_contiguous() and src.device.index == self.device
        n = src.shape[0]
        assert src.shape[1] == self.src_layout.frame_bytes and dst.shape == (n, self.dst_layout.frame_bytes)
        st = (stream or torch.cuda.current_stream(src.device)).cuda_stream
        sp, ss, sf = self.src_layout.pointer_args(src.data_ptr())
        dp, ds, df = self.dst_layout.pointer_args(dst.data_ptr())
        _lib.check(self._lib.tvs_scale_batch_static(self._ctx, sp, ss, sf, dp, ds, df, n, ctypes.c_void_p(st)))
        return dst

    def scale_device_planes(self, src_ptrs, src_pitch, src_fstride, dst_ptrs, dst_pitch, dst_fstride, n_frames, stream=0):
        """Raw form of tvs_scale_batch for pitched surfaces (e.g. NVDEC output): lists of ints."""
        pad = lambda v, fill: list(v) + [fill] * (3 - len(v))
        _lib.check(self._lib.tvs_scale_batch(
            self._ctx, PtrArr3(*pad(src_ptrs, None)), DiffArr3(*pad(src_pitch, 0)), DiffArr3(*pad(src_fstride, 0)),
            PtrArr3(*pad(dst_ptrs, None)), DiffArr3(*pad(dst_pitch, 0)), DiffArr3(*pad(dst_fstride, 0)),
            n_frames, ctypes.c_void_p(stream)))

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.tvs_destroy(self._ctx)
            self._ctx = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def filter_table(src_len, dst_len, alg, one, src_pos=128, dst_pos=128):
    """Host-only coefficient table of one axis (tvs_filter_table)."""
    lib = _lib.load()
    alg = F.ALG_BY_NAME[alg] if isinstance(alg, str) else alg
    taps = _lib.check(lib.tvs_filter_table(src_len, dst_len, alg, one, src_pos, dst_pos, None, 0, None, 0))
    coef = np.empty((dst_len, taps), dtype=np.int16)
    pos = np.empty(dst_len, dtype=np.int32)
    _lib.check(lib.tvs_filter_table(src_len, dst_len, alg, one, src_pos, dst_pos, coef.ctypes.data, coef.size,
                                    pos.ctypes.data, pos.size))
    return coef, pos
