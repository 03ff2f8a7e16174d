"""TEST INFRASTRUCTURE: the genuine libswscale, loaded through ctypes.

This is the code path the reference's ffmpeg child runs for ``scale=...`` (worker/tasks.py:1565-1580,
tools/upscale_benchmark.py:163,182): FFmpeg 8.0.1 libswscale 9.1.100, LGPL build with x86 SIMD, as
bundled in this image's opencv-python-headless wheel.  It is the parity
anchor (flags = alg | SWS_ACCURATE_RND | SWS_BITEXACT) and the ``kind: "reference"`` CPU baseline.
Never imported by the product.
"""
from __future__ import annotations

import ctypes
import glob
import os

import numpy as np

from . import build as _build

_PIXFMT_NAMES = {0: b"yuv420p", 1: b"nv12", 2: b"yuv420p10le", 3: b"p010le", 4: b"rgb24", 6: b"bgr24", 7: b"rgb48le", 8: b"rgb48be", 9: b"bgr48le", 10: b"bgr48be"}
_libs = None


def _load():
    global _libs
    if _libs is not None:
        return _libs
    dirs = [d for d in (_build.REF_DIR, _build.wheel_libs_dir()) if d and glob.glob(os.path.join(d, "libswscale-*"))]
    err = None
    for d in dirs:
        try:
            handles = []
            for prefix in ("libdrm-", "libcrypto-", "libavutil-", "libswscale-"):
                path = glob.glob(os.path.join(d, prefix + "*.so*"))[0]
                handles.append(ctypes.CDLL(path, mode=ctypes.RTLD_GLOBAL))
            avu, sws = handles[2], handles[3]
            break
        except (OSError, IndexError) as e:  # try the next location
            err = e
    else:
        raise ImportError(f"genuine libswscale not found (looked in {dirs}): {err}")
    sws.swscale_version.restype = ctypes.c_uint
    avu.av_version_info.restype = ctypes.c_char_p
    avu.av_get_pix_fmt.restype = ctypes.c_int
    avu.av_get_pix_fmt.argtypes = [ctypes.c_char_p]
    avu.av_force_cpu_flags.argtypes = [ctypes.c_int]
    avu.av_opt_set_int.restype = ctypes.c_int
    avu.av_opt_set_int.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_int64, ctypes.c_int]
    avu.av_opt_set_double.restype = ctypes.c_int
    avu.av_opt_set_double.argtypes = [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_double, ctypes.c_int]
    sws.sws_getContext.restype = ctypes.c_void_p
    sws.sws_getContext.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p] * 3
    sws.sws_alloc_context.restype = ctypes.c_void_p
    sws.sws_init_context.restype = ctypes.c_int
    sws.sws_init_context.argtypes = [ctypes.c_void_p] * 3
    sws.sws_scale.restype = ctypes.c_int
    sws.sws_scale.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                              ctypes.c_void_p, ctypes.c_void_p]
    sws.sws_freeContext.argtypes = [ctypes.c_void_p]
    _libs = (avu, sws)
    return _libs


def available() -> bool:
    try:
        _load()
        return True
    except ImportError:
        return False


def version() -> str:
    avu, sws = _load()
    v = sws.swscale_version()
    return "libswscale %d.%d.%d (FFmpeg %s)" % (v >> 16, (v >> 8) & 255, v & 255, avu.av_version_info().decode())


def force_plain_c(on: bool = True) -> None:
    """av_force_cpu_flags(0): disable SIMD.  Call before creating contexts."""
    _load()[0].av_force_cpu_flags(0 if on else -1)


class RefScaler:
    """One SwsContext.  Not safe for concurrent ``scale`` calls (same rule as libswscale)."""

    def __init__(self, src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, alg, chr_pos=None, bitexact=True, ranges=None, params=None):
        avu, sws = _load()
        self._sws = sws
        self.src = (src_w, src_h, src_fmt)
        self.dst = (dst_w, dst_h, dst_fmt)
        sf = avu.av_get_pix_fmt(_PIXFMT_NAMES[src_fmt])
        df = avu.av_get_pix_fmt(_PIXFMT_NAMES[dst_fmt])
        flags = alg | (0x40000 | 0x80000 if bitexact else 0)
        if (ranges is not None or params is not None) and chr_pos is None:
            chr_pos = (-513, -513, -513, -513)
        if chr_pos is None:
            self._ctx = sws.sws_getContext(src_w, src_h, sf, dst_w, dst_h, df, flags, None, None, None)
        else:
            self._ctx = sws.sws_alloc_context()
            opts = {b"srcw": src_w, b"srch": src_h, b"dstw": dst_w, b"dsth": dst_h, b"src_format": sf,
                    b"dst_format": df, b"sws_flags": flags, b"src_h_chr_pos": chr_pos[0],
                    b"src_v_chr_pos": chr_pos[1], b"dst_h_chr_pos": chr_pos[2], b"dst_v_chr_pos": chr_pos[3]}
            if ranges is not None:      # (src_range, dst_range), 1 = full (JPEG) range: what vf_scale's in_range / out_range set
                opts[b"src_range"], opts[b"dst_range"] = int(ranges[0]), int(ranges[1])
            for k, v in opts.items():
                if avu.av_opt_set_int(self._ctx, k, v, 0) < 0:
                    raise RuntimeError(f"av_opt_set_int({k!r}) failed")
            for k, v in zip((b"param0", b"param1"), params or ()):      # vf_scale's param0 / param1
                if v is not None and avu.av_opt_set_double(self._ctx, k, float(v), 0) < 0:
                    raise RuntimeError(f"av_opt_set_double({k!r}) failed")
            if sws.sws_init_context(self._ctx, None, None) < 0:
                raise RuntimeError("sws_init_context failed")
        if not self._ctx:
            raise RuntimeError("sws_getContext failed")

    def scale_into(self, src_planes, dst_planes) -> int:
        sp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in src_planes])
        ss = (ctypes.c_int * 4)(*[p.strides[0] for p in src_planes])
        dp = (ctypes.c_void_p * 4)(*[p.ctypes.data for p in dst_planes])
        ds = (ctypes.c_int * 4)(*[p.strides[0] for p in dst_planes])
        return self._sws.sws_scale(self._ctx, sp, ss, 0, self.src[1], dp, ds)

    def scale(self, src_planes):
        from thinvids_b200.frames import alloc_frame, plane_shapes, sample_dtype
        if self.src[2] in (4, 6, 7, 8, 9, 10):
            # packed RGB sources: the library's SIMD input stage reads and its output stage writes a few bytes past tight rows, so
            # give it padded buffers and crop
            h, row = src_planes[0].shape
            pad = np.zeros((h + 2, row + 64), dtype=src_planes[0].dtype)
            pad[:h, :row] = src_planes[0]
            shapes = plane_shapes(self.dst[2], self.dst[0], self.dst[1])
            big = [np.zeros((r + 2, c + 64), dtype=sample_dtype(self.dst[2])) for r, c in shapes]
            rows = self.scale_into([pad[:h]], [b[:r] for b, (r, c) in zip(big, shapes)])
            if rows != self.dst[1]:
                raise RuntimeError(f"sws_scale returned {rows}")
            return [np.ascontiguousarray(b[:r, :c]) for b, (r, c) in zip(big, shapes)]
        dst = alloc_frame(self.dst[2], self.dst[0], self.dst[1])
        rows = self.scale_into(src_planes, dst)
        if rows != self.dst[1]:
            raise RuntimeError(f"sws_scale returned {rows}")
        return dst

    def close(self):
        if getattr(self, "_ctx", None):
            self._sws.sws_freeContext(self._ctx)
            self._ctx = None

    __del__ = close
