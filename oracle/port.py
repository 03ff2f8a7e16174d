"""TEST INFRASTRUCTURE: ctypes wrapper around oracle/swscale_port.c (the plain-C restatement).

Same call shape as ``ref_swscale.RefScaler`` so tests can swap one for the other.  Pinned against the
genuine library and the golden vectors by tests/test_oracle.py.  Never imported by the product.
"""
from __future__ import annotations

import ctypes

import numpy as np

from . import build as _build

_lib = None


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(_build.build_port())
        lib.tvo_create.restype = ctypes.c_void_p
        lib.tvo_create.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p]
        lib.tvo_create_ex.restype = ctypes.c_void_p
        lib.tvo_create_ex.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p, ctypes.c_void_p]
        lib.tvo_destroy.argtypes = [ctypes.c_void_p]
        lib.tvo_scale.restype = ctypes.c_int
        lib.tvo_scale.argtypes = [ctypes.c_void_p] * 5
        lib.tvo_set_ranges.restype = ctypes.c_int
        lib.tvo_set_ranges.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int]
        lib.tvo_table.restype = ctypes.c_int
        lib.tvo_table.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
        _lib = lib
    return _lib


class PortScaler:
    def __init__(self, src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, alg, chr_pos=None, ranges=None, params=None):
        self._lib = _load()
        self.src = (src_w, src_h, src_fmt)
        self.dst = (dst_w, dst_h, dst_fmt)
        cp = (ctypes.c_int * 4)(*chr_pos) if chr_pos is not None else None
        pr = None
        if params is not None:      # (param0, param1) of ffmpeg's scale filter; None = the algorithm's default
            pr = (ctypes.c_double * 2)(*[123456.0 if v is None else float(v) for v in params])
        self._ctx = self._lib.tvo_create_ex(src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, alg, cp, pr)
        if not self._ctx:
            raise ValueError("libswscale refuses this conversion (a cascade whose source is not 4x larger than its destination)")
        if ranges is not None and self._lib.tvo_set_ranges(self._ctx, int(ranges[0]), int(ranges[1])) < 0:
            raise NotImplementedError("range conversion is not restated for RGB sources")

    def table(self, which: int):
        """(int16 coefficients [n, taps], int32 positions [n]) of table 0 hLum, 1 hChr, 2 vLum, 3 vChr."""
        n = ctypes.c_int()
        f = ctypes.POINTER(ctypes.c_int16)()
        pos = ctypes.POINTER(ctypes.c_int32)()
        taps = self._lib.tvo_table(self._ctx, which, ctypes.byref(n), ctypes.byref(f), ctypes.byref(pos))
        coef = np.ctypeslib.as_array(f, shape=(n.value, taps)).copy()
        return coef, np.ctypeslib.as_array(pos, shape=(n.value,)).copy()

    def scale_into(self, src_planes, dst_planes) -> int:
        sp = (ctypes.c_void_p * 3)(*([p.ctypes.data for p in src_planes] + [None] * (3 - len(src_planes))))
        ss = (ctypes.c_int * 3)(*([p.strides[0] for p in src_planes] + [0] * (3 - len(src_planes))))
        dp = (ctypes.c_void_p * 3)(*([p.ctypes.data for p in dst_planes] + [None] * (3 - len(dst_planes))))
        ds = (ctypes.c_int * 3)(*([p.strides[0] for p in dst_planes] + [0] * (3 - len(dst_planes))))
        return self._lib.tvo_scale(self._ctx, sp, ss, dp, ds)

    def scale(self, src_planes):
        from thinvids_b200.frames import alloc_frame
        dst = alloc_frame(self.dst[2], self.dst[0], self.dst[1])
        rows = self.scale_into(src_planes, dst)
        if rows != self.dst[1]:
            raise RuntimeError(f"tvo_scale returned {rows}")
        return dst

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.tvo_destroy(self._ctx)
            self._ctx = None

    __del__ = close
