"""Synthetic frames and plane geometry for the scale / pixel-format stage (4:2:0 formats, plus packed rgb24 as a
source: what tools/upscale_benchmark.py:159-182 feeds its lanczos stage).

The reference never owns pixels itself (it hands files to ffmpeg, worker/tasks.py:1558-1586),
so the frame containers here follow ffmpeg's pixel formats: a frame is a list of numpy planes,
``[Y, U, V]`` for planar formats and ``[Y, UV]`` for the semi-planar ones, 10-bit samples as
little-endian uint16 (p010le keeps its 10 bits in the top of the word).

The generators implement SURVEY.md §8d / Appendix B.2 exactly, so hashes of scaled output can be
compared with the golden vectors in tests/golden/.
"""
from __future__ import annotations

import hashlib
from dataclasses import dataclass

import numpy as np

# Pixel-format codes of the C ABI (include/tvscale.h).
YUV420P, NV12, YUV420P10LE, P010LE, RGB24 = 0, 1, 2, 3, 4
RGB48LE, RGB48BE, BGR48LE, BGR48BE = 7, 8, 9, 10   # 16 bits per component (16-bit PNGs decode as rgb48be, OpenCV holds bgr48le)
BGR24 = 6   # (5 is the C ABI's yuvj420p, which the Python layer spells as yuv420p + full range)
FMT_BY_NAME = {"yuv420p": YUV420P, "nv12": NV12, "yuv420p10le": YUV420P10LE, "p010le": P010LE, "rgb24": RGB24, "bgr24": BGR24, "rgb48le": RGB48LE, "rgb48be": RGB48BE, "bgr48le": BGR48LE, "bgr48be": BGR48BE}
SOURCE_ONLY_FORMATS = {RGB24, BGR24, RGB48LE, RGB48BE, BGR48LE, BGR48BE}
NAME_BY_FMT = {v: k for k, v in FMT_BY_NAME.items()}

# Scaler algorithm flags, numerically equal to libswscale's SWS_* values.
SWS_BILINEAR, SWS_BICUBIC, SWS_AREA, SWS_LANCZOS = 2, 4, 0x20, 0x200
ALG_BY_NAME = {"bilinear": SWS_BILINEAR, "bicubic": SWS_BICUBIC, "area": SWS_AREA, "lanczos": SWS_LANCZOS}
SWS_ACCURATE_RND, SWS_BITEXACT = 0x40000, 0x80000


def fmt_code(fmt) -> int:
    return fmt if isinstance(fmt, int) else FMT_BY_NAME[fmt]


def is_semiplanar(fmt) -> bool:
    return fmt_code(fmt) in (NV12, P010LE)


def bit_depth(fmt) -> int:
    code = fmt_code(fmt)
    return 8 if code in (YUV420P, NV12, RGB24, BGR24) else 16 if code in (RGB48LE, RGB48BE, BGR48LE, BGR48BE) else 10


def sample_dtype(fmt):
    return np.uint8 if bit_depth(fmt) == 8 else np.uint16


def plane_shapes(fmt, w: int, h: int):
    """(rows, cols-in-samples) of every stored plane of a tightly packed frame."""
    cw, ch = (w + 1) // 2, (h + 1) // 2
    if fmt_code(fmt) in (RGB24, BGR24, RGB48LE, RGB48BE, BGR48LE, BGR48BE):
        return [(h, 3 * w)]          # one plane of R, G, B bytes
    if is_semiplanar(fmt):
        return [(h, w), (ch, 2 * cw)]
    return [(h, w), (ch, cw), (ch, cw)]


def frame_bytes(fmt, w: int, h: int) -> int:
    bps = 1 if bit_depth(fmt) == 8 else 2
    return sum(r * c for r, c in plane_shapes(fmt, w, h)) * bps


def alloc_frame(fmt, w: int, h: int):
    return [np.zeros(s, dtype=sample_dtype(fmt)) for s in plane_shapes(fmt, w, h)]


def random_frame(fmt, w: int, h: int, seed: int = 1234):
    """Uniform-noise frame: one generator per frame, planes drawn in storage order (Appendix B.2)."""
    rng = np.random.default_rng(seed)
    planes = []
    for shape in plane_shapes(fmt, w, h):
        if bit_depth(fmt) == 8:
            planes.append(rng.integers(0, 256, shape, dtype=np.uint8))
        else:
            p = rng.integers(0, 65536 if bit_depth(fmt) == 16 else 1024, shape, dtype=np.uint16)
            planes.append(p << 6 if fmt_code(fmt) == P010LE else p)
    return planes


PATTERNS = ("cols3", "checker", "rows3", "blocks", "white")


def pattern_frame(fmt, w: int, h: int, pattern: str):
    """Adversarial clamp patterns of Appendix B.2: each stored plane filled from its own (y, x) grid."""
    hi = 255 if bit_depth(fmt) == 8 else (1023 << 6 if fmt_code(fmt) == P010LE else 1023)
    planes = []
    for rows, cols in plane_shapes(fmt, w, h):
        y, x = np.mgrid[0:rows, 0:cols]
        if pattern == "cols3":
            m = x % 3 == 1
        elif pattern == "checker":
            m = (x + y) % 2 == 0
        elif pattern == "rows3":
            m = y % 3 == 1
        elif pattern == "blocks":
            m = (x // 3 + y // 3) % 2 == 0
        elif pattern == "white":
            m = np.ones((rows, cols), dtype=bool)
        else:
            raise ValueError(pattern)
        planes.append(np.where(m, hi, 0).astype(sample_dtype(fmt)))
    return planes


def frame_sha256(planes) -> str:
    h = hashlib.sha256()
    for p in planes:
        h.update(np.ascontiguousarray(p).tobytes())
    return h.hexdigest()


def scaled_width(src_w: int, src_h: int, dst_h: int) -> int:
    """ffmpeg's ``scale=-2:H``: keep the aspect ratio, round the width to a multiple of two.

    Matches the shapes the reference produces with ``scale=-2:{H}`` (worker/tasks.py:62-65,448):
    720x480 -> 1620x1080, 1920x1080 -> 1280x720.
    """
    w = int(src_w * dst_h / (src_h * 2.0) + 0.5) * 2
    return max(w, 2)


@dataclass(frozen=True)
class ScaleConfig:
    """One named workload of BASELINE.json / SURVEY.md §8."""

    name: str
    src_fmt: str
    src_w: int
    src_h: int
    dst_fmt: str
    dst_w: int
    dst_h: int
    alg: str
    n_frames: int

    @property
    def src_bytes(self) -> int:
        return frame_bytes(self.src_fmt, self.src_w, self.src_h)

    @property
    def dst_bytes(self) -> int:
        return frame_bytes(self.dst_fmt, self.dst_w, self.dst_h)

    @property
    def algorithmic_bytes(self) -> int:
        """Bytes per frame of SURVEY.md §8d: source read once + destination written once."""
        return self.src_bytes + self.dst_bytes


CONFIGS = {
    "C1": ScaleConfig("C1", "yuv420p", 720, 480, "yuv420p", 1920, 1080, "lanczos", 300),
    "C2": ScaleConfig("C2", "yuv420p", 1920, 1080, "yuv420p", 3840, 2160, "bicubic", 64),
    "C3": ScaleConfig("C3", "nv12", 3840, 2160, "nv12", 1920, 1080, "area", 64),
    "C4": ScaleConfig("C4", "yuv420p10le", 3840, 2160, "yuv420p10le", 1920, 1080, "lanczos", 64),
    "C5": ScaleConfig("C5", "yuv420p", 1920, 1080, "yuv420p", 3840, 2160, "lanczos", 240),
}
# Shapes the reference's worker produces by default (worker/tasks.py:62-65) and planar / 8-bit variants of the
# named configs, for profiling; not BASELINE.json bench lines.
EXTRA_CONFIGS = {
    "W1080": ScaleConfig("W1080", "yuv420p", 720, 480, "nv12", 1620, 1080, "bicubic", 300),
    "W720": ScaleConfig("W720", "yuv420p", 1920, 1080, "nv12", 1280, 720, "bicubic", 128),
    "W720P": ScaleConfig("W720P", "yuv420p", 1920, 1080, "yuv420p", 1280, 720, "bicubic", 128),
    "W4K": ScaleConfig("W4K", "yuv420p", 3840, 2160, "nv12", 1920, 1080, "bicubic", 64),
    "DVD480": ScaleConfig("DVD480", "yuv420p", 720, 480, "nv12", 720, 480, "bicubic", 1000),      # SCALE_FILTER_480 on a DVD source
    "HDR8": ScaleConfig("HDR8", "yuv420p10le", 1920, 1080, "nv12", 1920, 1080, "bicubic", 128),   # same-size 10 -> 8 bit
    "C3P": ScaleConfig("C3P", "yuv420p", 3840, 2160, "yuv420p", 1920, 1080, "area", 64),
    "C4S": ScaleConfig("C4S", "p010le", 3840, 2160, "p010le", 1920, 1080, "lanczos", 64),      # C4 on NVDEC-style P010 surfaces
    "C4N": ScaleConfig("C4N", "p010le", 3840, 2160, "nv12", 1920, 1080, "bicubic", 64),        # 10-bit 4K source to the 8-bit 1080p target
    "RGB1": ScaleConfig("RGB1", "rgb24", 720, 480, "yuv420p", 1920, 1080, "lanczos", 120),   # upscale_benchmark.py's real input
    "RGB2": ScaleConfig("RGB2", "rgb48be", 720, 480, "yuv420p", 1920, 1080, "lanczos", 120),  # the same from 16-bit PNG frames
    "C4P8": ScaleConfig("C4P8", "yuv420p", 3840, 2160, "yuv420p", 1920, 1080, "lanczos", 64),
}
CONFIGS.update(EXTRA_CONFIGS)
