"""TEST INFRASTRUCTURE: numpy restatement of ffmpeg's bwdif de-interlacer as the reference asks for it
(``bwdif=mode=send_frame:parity=auto:deint=all``: worker/tasks.py:62-63,446-447, tools/upscale_benchmark.py:128-132).

PARITY UNPINNED.  The arithmetic lives in FFmpeg's libavfilter (vf_bwdif.c / bwdifdsp.c / yadif_common.c), which is neither
under /root/reference nor in this image (the bundled FFmpeg libraries are avcodec / avformat / avutil / swscale /
swresample; there is no libavfilter and no ffmpeg binary), so there is no binary and no golden vector to check this
restatement against.  It follows the published algorithm:

* one output frame per input frame (send_frame); with parity=auto and a top-field-first frame the top field (even lines)
  is kept and the odd lines are interpolated, for bottom-field-first the other way round; kept lines are copied;
* the first and the last frame of a stream have no temporal neighbour on one side and use the intra filter
  (YADIF_FIELD_END): ``(5077 * (above + below) - 981 * (above3 + below3)) >> 13`` with mirrored rows at the borders;
* every other interpolated line: temporal prediction ``d = (prev2 + next2) >> 1`` (for send_frame prev2 = the previous
  frame and next2 = the current one), the largest of three temporal differences as the allowed deviation, tightened by
  a spatial check on the rows two above / below, and a spatial interpolation - with the high-frequency temporal term
  (coefficients 5570 / 3801 / 1016 over rows 0, +-2, +-4 of prev2 / next2 and 4309 / 213 over rows +-1, +-3 of the current
  frame) when the vertical step |c - e| exceeds the temporal difference, else the plain 5077 / 981 kernel - clamped to
  ``d +- diff`` and to the sample range.  Rows closer than 4 to the top or 5 to the bottom use ``(c + e) >> 1`` instead
  (the spatial check only from row 2 to row h - 4).

Only tests import this.  The product's kernel (thinvids_b200/csrc/kernels_bwdif.cu) is compared with it bit for bit, which
proves the two agree with each other - not that either agrees with ffmpeg.
"""
from __future__ import annotations

import numpy as np

COEF_LF = (4309, 213)
COEF_HF = (5570, 3801, 1016)
COEF_SP = (5077, 981)


def _line(plane: np.ndarray, y: int) -> np.ndarray:
    return plane[y].astype(np.int64)


def _filter_plane(prev, cur, nxt, tff: int, intra: bool, depth: int) -> np.ndarray:
    h, w = cur.shape
    clip_max = (1 << depth) - 1
    df = (depth + 7) // 8                      # bytes per sample; libavfilter uses it in its row tests
    parity = tff ^ 1                           # send_frame: the frame's first field is kept
    out = cur.copy()
    prev2, next2 = prev, cur                   # filter_line is called with parity ^ tff = 1
    for y in range(h):
        if not ((y ^ parity) & 1):
            continue                           # a row of the kept field
        if intra:
            p1 = y + 1 if (y + df) < h else y - 1
            m1 = y - 1 if y > (df - 1) else y + 1
            p3 = y + 3 if (y + 3 * df) < h else y - 3
            m3 = y - 3 if y > (3 * df - 1) else y + 3
            # rows the mirrored offsets may still push outside very small planes are clamped
            p1, m1, p3, m3 = (min(max(v, 0), h - 1) for v in (p1, m1, p3, m3))
            interpol = (COEF_SP[0] * (_line(cur, m1) + _line(cur, p1)) - COEF_SP[1] * (_line(cur, m3) + _line(cur, p3))) >> 13
            out[y] = np.clip(interpol, 0, clip_max).astype(cur.dtype)
            continue
        edge = (y < 4) or ((y + 5) > h)
        p1 = y + 1 if (y + df) < h else y - 1
        m1 = y - 1 if y > (df - 1) else y + 1
        if not edge:
            p1, m1 = y + 1, y - 1
        p1, m1 = min(max(p1, 0), h - 1), min(max(m1, 0), h - 1)
        c, e = _line(cur, m1), _line(cur, p1)
        d = (_line(prev2, y) + _line(next2, y)) >> 1
        td0 = np.abs(_line(prev2, y) - _line(next2, y))
        td1 = (np.abs(_line(prev, m1) - c) + np.abs(_line(prev, p1) - e)) >> 1
        td2 = (np.abs(_line(nxt, m1) - c) + np.abs(_line(nxt, p1) - e)) >> 1
        diff = np.maximum(np.maximum(td0 >> 1, td1), td2)
        still = diff == 0                          # decided before the spatial check, as in FILTER1
        spat = (not edge) or not ((y < 2) or ((y + 3) > h))
        if spat:
            b = ((_line(prev2, y - 2) + _line(next2, y - 2)) >> 1) - c
            f = ((_line(prev2, y + 2) + _line(next2, y + 2)) >> 1) - e
            dc, de = d - c, d - e
            mx = np.maximum(np.maximum(de, dc), np.minimum(b, f))
            mn = np.minimum(np.minimum(de, dc), np.maximum(b, f))
            diff = np.maximum(np.maximum(diff, mn), -mx)
        if edge:
            interpol = (c + e) >> 1
        else:
            c3, e3 = _line(cur, y - 3), _line(cur, y + 3)
            hf = (COEF_HF[0] * (_line(prev2, y) + _line(next2, y))
                  - COEF_HF[1] * (_line(prev2, y - 2) + _line(next2, y - 2) + _line(prev2, y + 2) + _line(next2, y + 2))
                  + COEF_HF[2] * (_line(prev2, y - 4) + _line(next2, y - 4) + _line(prev2, y + 4) + _line(next2, y + 4))) >> 2
            full = (hf + COEF_LF[0] * (c + e) - COEF_LF[1] * (c3 + e3)) >> 13
            plain = (COEF_SP[0] * (c + e) - COEF_SP[1] * (c3 + e3)) >> 13
            interpol = np.where(np.abs(c - e) > td0, full, plain)
        interpol = np.where(interpol > d + diff, d + diff, np.where(interpol < d - diff, d - diff, interpol))
        res = np.clip(interpol, 0, clip_max)
        out[y] = np.where(still, d, res).astype(cur.dtype)
    return out


def bwdif_stream(frames, tff: int = 1, depth: int = 8, stream_start: bool = True, stream_end: bool = True):
    """frames: list of frames, each a list of planes (numpy, uint8 or uint16).  Returns the de-interlaced frames.
    With stream_start / stream_end False the first / last frame of the list is context only (it is not output)."""
    n = len(frames)
    out = []
    first = 0 if stream_start else 1
    last = n - 1 if stream_end else n - 2
    for i in range(first, last + 1):
        intra = (i == 0 and stream_start) or (i == n - 1 and stream_end)
        prev = frames[i - 1] if i > 0 else frames[i]
        nxt = frames[i + 1] if i + 1 < n else frames[i]
        out.append([_filter_plane(prev[p], frames[i][p], nxt[p], tff, intra, depth) for p in range(len(frames[i]))])
    return out
