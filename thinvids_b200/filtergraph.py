"""Host-side mirror of the reference's scale / pixel-format operator interface.

The reference never calls a scaler directly: its `encode` task (worker/tasks.py:1354) and
tools/upscale_benchmark.py (`encode_output`, :143) build an ffmpeg ``-vf`` string and let ffmpeg's
``scale`` / ``format`` filters do the work.  That string IS the operator API of this path, so this module

* rebuilds the same strings from the same inputs (`software_scale_filter`, `vaapi_scale_filter`,
  `reset_segment_video_pts_filter`, target-height policy) so the worker keeps choosing shapes exactly as before
  (worker/tasks.py:56-68, 426-500), and
* parses such a string (`parse_filtergraph`, `ScaleStage`) into the geometry, algorithm and pixel format the
  CUDA scaler needs, with ffmpeg's semantics: ``scale=-2:H`` keeps the aspect ratio and rounds the width to
  a multiple of two, no ``flags=`` means bicubic, ``format=nv12`` folds the repack into the scaler.

Elements that are not pixel work of this stage are accepted and ignored (``setpts=PTS-STARTPTS`` only rewrites
timestamps, ``hwupload`` is the hand-off to the encoder).  ``bwdif`` is a different per-pixel stage that this
package does not implement yet: a graph that contains it is refused unless the caller
states that frames arrive already de-interlaced.
"""
from __future__ import annotations

import math
import os
import re
from dataclasses import dataclass, field
from typing import Mapping

from . import frames as F

ENV = os.environ.get

# --- the reference's knobs (worker/tasks.py:56-65), same names, same defaults, same env overrides -------------
ALLOWED_TARGET_HEIGHTS = {480, 576, 720, 1080}
try:
    DEFAULT_TARGET_HEIGHT = int(ENV("VEM_DEFAULT_TARGET_HEIGHT", "1080"))
except Exception:
    DEFAULT_TARGET_HEIGHT = 1080
_BWDIF = "bwdif=mode=send_frame:parity=auto:deint=all"
SCALE_FILTER_480 = ENV("VEM_SCALE_FILTER_480", f"{_BWDIF},scale=-2:480,format=nv12,hwupload")
SCALE_FILTER_576 = ENV("VEM_SCALE_FILTER_576", f"{_BWDIF},scale=-2:576,format=nv12,hwupload")
SCALE_FILTER_720 = ENV("VEM_SCALE_FILTER_720", "scale=-2:720,format=nv12,hwupload")
SCALE_FILTER_1080 = ENV("VEM_SCALE_FILTER_1080", "scale=-2:1080,format=nv12,hwupload")


def normalize_target_height(value) -> int:
    """worker/tasks.py:426 — anything that is not one of the allowed heights becomes the default."""
    try:
        h = int(value)
    except Exception:
        h = DEFAULT_TARGET_HEIGHT
    return h if h in ALLOWED_TARGET_HEIGHTS else DEFAULT_TARGET_HEIGHT


def vaapi_scale_filter(target_height: int) -> str:
    """worker/tasks.py:436 — the hardware-encode branch's filter text."""
    return {480: SCALE_FILTER_480, 576: SCALE_FILTER_576, 1080: SCALE_FILTER_1080}.get(target_height, SCALE_FILTER_720)


def software_scale_filter(target_height: int, *, deinterlace: bool) -> str:
    """worker/tasks.py:444 — the libx264 branch's filter text."""
    parts = [_BWDIF] if deinterlace else []
    parts.append(f"scale=-2:{target_height}")
    return ",".join(parts)


def reset_segment_video_pts_filter(filtergraph: str) -> str:
    """worker/tasks.py:452 — every segment restarts its video timeline at its first decoded frame."""
    filtergraph = (filtergraph or "").strip()
    return f"setpts=PTS-STARTPTS,{filtergraph}" if filtergraph else "setpts=PTS-STARTPTS"


def source_dimensions_for_job(job: Mapping) -> tuple[int, int]:
    """worker/tasks.py:464 — "WxH" from the job hash, (0, 0) when absent or malformed."""
    m = re.match(r"^\s*(\d+)\s*x\s*(\d+)\s*$", str(job.get("source_resolution") or "").strip().lower())
    return (int(m.group(1)), int(m.group(2))) if m else (0, 0)


def dvd_native_target_height(job: Mapping):
    """worker/tasks.py:475 — SD MPEG-2 under a dvd/ or movies/ path stays at its native 480 / 576 lines."""
    filename = str(job.get("filename") or "").strip().lower()
    codec = str(job.get("source_codec") or "").strip().lower()
    w, h = source_dimensions_for_job(job)
    on_dvd_path = any(filename.startswith(p) or ("/" + p) in filename for p in ("dvd/", "movies/"))
    if not on_dvd_path or not (0 < w <= 720 and 0 < h <= 576):
        return None
    if codec and codec != "mpeg2video":
        return None
    return 480 if h <= 480 else 576


def effective_target_height_for_job(job: Mapping) -> tuple[int, bool]:
    """worker/tasks.py:495 — (height, force_deinterlace); takes the job hash itself instead of a Redis key."""
    dvd = dvd_native_target_height(job)
    if dvd:
        return dvd, True
    return normalize_target_height(job.get("target_height")), False


def encode_filtergraph(job: Mapping) -> str:
    """The ``-vf`` argument `encode` passes to ffmpeg for this job (worker/tasks.py:1534-1586)."""
    height, deinterlace = effective_target_height_for_job(job)
    if int(job.get("software_encode") or 0):
        return reset_segment_video_pts_filter(software_scale_filter(height, deinterlace=deinterlace))
    return reset_segment_video_pts_filter(vaapi_scale_filter(height))


# --- parsing ---------------------------------------------------------------------------------------------------
class FilterGraphError(ValueError):
    pass


@dataclass
class FilterSpec:
    name: str
    args: list[str] = field(default_factory=list)        # positional
    kwargs: dict[str, str] = field(default_factory=dict)  # key=value


def parse_filtergraph(text: str) -> list[FilterSpec]:
    """Linear chains only (what the reference emits): ``f1=a:b:k=v,f2,...``."""
    out = []
    for chunk in filter(None, (c.strip() for c in (text or "").split(","))):
        if any(ch in chunk for ch in "[];"):
            raise FilterGraphError(f"only linear filter chains are supported, got {chunk!r}")
        name, _, rest = chunk.partition("=")
        spec = FilterSpec(name.strip())
        for item in filter(None, rest.split(":")) if rest else []:
            k, eq, v = item.partition("=")
            if eq:
                spec.kwargs[k.strip()] = v.strip()
            else:
                spec.args.append(k.strip())
        out.append(spec)
    return out


_SCALE_FLAG_NAMES = {"bilinear": F.SWS_BILINEAR, "bicubic": F.SWS_BICUBIC, "area": F.SWS_AREA, "lanczos": F.SWS_LANCZOS,
                     "accurate_rnd": F.SWS_ACCURATE_RND, "bitexact": F.SWS_BITEXACT}
_PASSTHROUGH = {"setpts", "hwupload", "null", "fps"}   # no pixel work in this stage
# scale's in_range / out_range values (libavfilter/vf_scale.c): None = keep what the frames carry
_RANGE_NAMES = {"auto": None, "unknown": None, "full": "full", "jpeg": "full", "pc": "full",
                "limited": "limited", "mpeg": "limited", "tv": "limited"}


def _eval_dim(expr: str, other: int, src_this: int, src_other: int) -> int:
    """ffmpeg scale dimension: a positive integer, or -N = keep aspect, rounded to a multiple of N."""
    try:
        v = int(expr)
    except ValueError:
        raise FilterGraphError(f"scale dimension {expr!r}: only integers and -N are supported") from None
    if v > 0:
        return v
    if v == 0:
        return src_this
    n = -v
    # ffmpeg: w = rint(h * a / n) * n with a = iw / ih
    return max(int(math.floor(other * src_this / (src_other * n) + 0.5)) * n, n)   # av_rescale rounds half away from zero


@dataclass(frozen=True)
class ScaleStage:
    """What a reference filter string asks of the scale / pixel-format stage for one source geometry."""

    src_w: int
    src_h: int
    src_fmt: str
    dst_w: int
    dst_h: int
    dst_fmt: str
    flags: int
    deinterlace: bool       # the graph contains bwdif (must happen upstream)
    hwupload: bool          # the graph hands frames to a hardware encoder
    # Sample ranges, "limited" (MPEG) or "full" (JPEG).  ffmpeg's decoders announce full-range 4:2:0 as yuvj420p (MJPEG, most
    # phone and camera H.264); the scale filter keeps the input's range unless out_range= says otherwise, so the usual
    # full-range part passes through with no conversion and comes out full range.
    src_range: str = "limited"
    dst_range: str = "limited"
    params: tuple = (None, None)   # scale's param0 / param1 (bicubic B and C, lanczos lobes); None = the algorithm's default

    @classmethod
    def from_filtergraph(cls, text: str, src_w: int, src_h: int, src_fmt: str = "yuv420p", *,
                         deinterlaced_upstream: bool = False, default_dst_fmt: str | None = None) -> "ScaleStage":
        dst_w, dst_h, flags = src_w, src_h, F.SWS_BICUBIC
        src_range, dst_range = "limited", None
        params = [None, None]
        if src_fmt == "yuvj420p":
            src_fmt, src_range = "yuv420p", "full"
        dst_fmt = default_dst_fmt or src_fmt
        deint = hwup = False
        seen_scale = False
        for f in parse_filtergraph(text):
            if f.name == "scale":
                if seen_scale:
                    raise FilterGraphError("more than one scale filter")
                seen_scale = True
                w_expr = f.kwargs.get("w", f.kwargs.get("width", f.args[0] if f.args else "0"))
                h_expr = f.kwargs.get("h", f.kwargs.get("height", f.args[1] if len(f.args) > 1 else "0"))
                if str(h_expr).lstrip().startswith("-") and str(w_expr).lstrip().startswith("-"):
                    raise FilterGraphError("scale: width and height cannot both be derived")
                if str(w_expr).lstrip().startswith("-"):
                    dst_h = _eval_dim(h_expr, 0, src_h, src_w)
                    dst_w = _eval_dim(w_expr, dst_h, src_w, src_h)
                else:
                    dst_w = _eval_dim(w_expr, 0, src_w, src_h)
                    dst_h = _eval_dim(h_expr, dst_w, src_h, src_w)
                if "flags" in f.kwargs or "sws_flags" in f.kwargs:
                    flags = 0
                    for name in filter(None, f.kwargs.get("flags", f.kwargs.get("sws_flags", "")).split("+")):
                        if name not in _SCALE_FLAG_NAMES:
                            raise FilterGraphError(f"scale flag {name!r} is not supported")
                        flags |= _SCALE_FLAG_NAMES[name]
                    if not flags & (F.SWS_BILINEAR | F.SWS_BICUBIC | F.SWS_AREA | F.SWS_LANCZOS):
                        flags |= F.SWS_BICUBIC
                for k, key in enumerate(("param0", "param1")):
                    if key in f.kwargs:
                        try:
                            params[k] = float(f.kwargs[key])
                        except ValueError:
                            raise FilterGraphError(f"scale {key}={f.kwargs[key]!r} is not a number") from None
                for key in ("in_range", "out_range"):
                    val = f.kwargs.get(key, "auto")
                    if val not in _RANGE_NAMES:
                        raise FilterGraphError(f"scale {key}={val!r} is not supported")
                    if _RANGE_NAMES[val] and key == "in_range":
                        src_range = _RANGE_NAMES[val]
                    elif _RANGE_NAMES[val]:
                        dst_range = _RANGE_NAMES[val]
            elif f.name == "format":
                fmt = f.kwargs.get("pix_fmts", f.args[0] if f.args else "")
                if fmt == "yuvj420p":
                    fmt, dst_range = "yuv420p", "full"
                if fmt not in F.FMT_BY_NAME or F.FMT_BY_NAME[fmt] in F.SOURCE_ONLY_FORMATS:
                    ok = sorted(n for n, c in F.FMT_BY_NAME.items() if c not in F.SOURCE_ONLY_FORMATS)
                    raise FilterGraphError(f"format={fmt!r}: only {ok} are supported")
                dst_fmt = fmt
            elif f.name == "bwdif":
                deint = True
            elif f.name == "hwupload":
                hwup = True
            elif f.name not in _PASSTHROUGH:
                raise FilterGraphError(f"filter {f.name!r} is not part of the scale / pixel-format stage")
        if deint and not deinterlaced_upstream:
            raise FilterGraphError("the graph de-interlaces with bwdif, which this stage does not implement; "
                                   "de-interlace upstream and pass deinterlaced_upstream=True")
        if src_fmt not in F.FMT_BY_NAME:
            raise FilterGraphError(f"source format {src_fmt!r} is not supported")
        if F.FMT_BY_NAME[dst_fmt] in F.SOURCE_ONLY_FORMATS:
            # tools/upscale_benchmark.py:163,182 always names the output format after its rgb24 input
            raise FilterGraphError(f"a {src_fmt} source needs an explicit format= (yuv420p, nv12, ...) after scale")
        if src_range != (dst_range or src_range) and F.FMT_BY_NAME[src_fmt] in F.SOURCE_ONLY_FORMATS:
            raise FilterGraphError("range conversion is not supported for rgb24 sources")
        return cls(src_w, src_h, src_fmt, dst_w, dst_h, dst_fmt, flags, deint, hwup, src_range, dst_range or src_range, tuple(params))

    @property
    def is_identity(self) -> bool:
        return (self.src_w, self.src_h, self.src_fmt, self.src_range) == (self.dst_w, self.dst_h, self.dst_fmt, self.dst_range)

    @staticmethod
    def _raw_args(fmt: str, rng: str) -> list:
        if rng != "full":
            return ["-pix_fmt", fmt]
        # full range: 8-bit planar 4:2:0 has its own name in ffmpeg; every other layout carries the range as an option
        return ["-pix_fmt", "yuvj420p"] if fmt == "yuv420p" else ["-color_range", "pc", "-pix_fmt", fmt]

    @property
    def src_raw_args(self) -> list:
        """ffmpeg arguments that describe the raw frames entering the stage (decode child's output)."""
        return self._raw_args(self.src_fmt, self.src_range)

    @property
    def dst_raw_args(self) -> list:
        """ffmpeg arguments that describe the raw frames leaving the stage (encode child's input)."""
        return self._raw_args(self.dst_fmt, self.dst_range)

    def make_scaler(self, device: int = 0):
        """The CUDA context for this stage (thinvids_b200.swscale.Scaler)."""
        from .swscale import Scaler
        return Scaler(self.src_w, self.src_h, self.src_fmt, self.dst_w, self.dst_h, self.dst_fmt, self.flags, device=device,
                      src_range=self.src_range, dst_range=self.dst_range, params=self.params)
