"""ctypes binding of libtvscale.so (include/tvscale.h).  Fails loudly when the library is missing:
there is no CPU fallback anywhere in this package."""
from __future__ import annotations

import ctypes
import os

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libtvscale.so")

# Every symbol include/tvscale.h declares; tests check the built library exports each one.
EXPORTS = ("tvs_create", "tvs_create_ex", "tvs_destroy", "tvs_scale_batch", "tvs_scale_host", "tvs_scale_batch_static", "tvs_filter_table",
           "tvs_context_table", "tvs_info", "tvs_set_path", "tvs_last_error", "tvs_version", "tvs_host_alloc", "tvs_host_free", "tvs_bwdif_batch", "tvs_bwdif_host")

STATIC_SCRATCH_BYTES = 32 << 20   # TVS_STATIC_SCRATCH_BYTES
INFO_KERNEL_LAUNCHES, INFO_PATH, INFO_SCRATCH_BYTES, INFO_SRC_FRAME_BYTES, INFO_DST_FRAME_BYTES = range(5)
PARAM_DEFAULT = 123456.0                                 # TVS_PARAM_DEFAULT
FMT_YUVJ420P = 5                                         # TVS_FMT_YUVJ420P (frames.BGR24 = 6 is TVS_FMT_BGR24)
SRC_FULL_RANGE, DST_FULL_RANGE = 0x10000000, 0x20000000  # TVS_SRC_FULL_RANGE, TVS_DST_FULL_RANGE

_lib = None


class TvscaleError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"tvscale error {code}: {message}")
        self.code = code


def load() -> ctypes.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -m thinvids_b200.build` "
                          "(nvcc, sm_100a).  thinvids_b200 has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    P, I, Z = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t
    lib.tvs_create.restype = P
    lib.tvs_create.argtypes = [I] * 8 + [P]
    lib.tvs_create_ex.restype = P
    lib.tvs_create_ex.argtypes = [I] * 8 + [P, P]
    lib.tvs_destroy.restype = None
    lib.tvs_destroy.argtypes = [P]
    lib.tvs_scale_batch.restype = I
    lib.tvs_scale_batch.argtypes = [P, P, P, P, P, P, P, I, P]
    lib.tvs_scale_batch_static.restype = I
    lib.tvs_scale_batch_static.argtypes = [P, P, P, P, P, P, P, I, P]
    lib.tvs_scale_host.restype = I
    lib.tvs_scale_host.argtypes = [P, P, P, P, P, P, P, I]
    lib.tvs_filter_table.restype = I
    lib.tvs_filter_table.argtypes = [I] * 6 + [P, Z, P, Z]
    lib.tvs_context_table.restype = I
    lib.tvs_context_table.argtypes = [P, I, P, Z, P, Z]
    lib.tvs_info.restype = ctypes.c_longlong
    lib.tvs_info.argtypes = [P, I]
    lib.tvs_set_path.restype = I
    lib.tvs_set_path.argtypes = [P, I]
    lib.tvs_host_alloc.restype = P
    lib.tvs_host_alloc.argtypes = [Z]
    lib.tvs_host_free.restype = None
    lib.tvs_host_free.argtypes = [P]
    lib.tvs_bwdif_batch.restype = I
    lib.tvs_bwdif_batch.argtypes = [I, I, I, I, P, P, P, P, P, P, I, I, I, P]
    lib.tvs_bwdif_host.restype = I
    lib.tvs_bwdif_host.argtypes = [I, I, I, I, P, P, I, I, I]
    lib.tvs_last_error.restype = ctypes.c_char_p
    lib.tvs_version.restype = ctypes.c_char_p
    _lib = lib
    return lib


def last_error() -> str:
    return load().tvs_last_error().decode()


def check(code: int) -> int:
    if code < 0:
        raise TvscaleError(code, last_error())
    return code
