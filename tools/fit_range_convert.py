"""How libswscale 9.1.100 converts between full (JPEG) and limited (MPEG) range on its 15-bit intermediate: fitted
against the genuine library, because the constants changed in FFmpeg 7.1/8 and the source is not in this image.

Model: after the horizontal pass t' = min((t * coeff + offset) >> 14, 32767) for luma and for chroma (own constants).
With an unscaled vertical axis the output is (t' + 64) >> 7, so every observed (t, output) pair bounds offset for a
given coeff; the script scans coeff, intersects the bounds over a dense set of t and prints what is left.
Needs oracle/_ref (or the opencv wheel); CPU only.      python tools/fit_range_convert.py
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from oracle import port, ref_swscale  # noqa: E402
from thinvids_b200 import frames as F  # noqa: E402


def intermediates(tab, plane):
    coef, pos = tab
    coef = np.asarray(coef, dtype=np.int64)
    pos = np.asarray(pos, dtype=np.int64)
    taps = coef.shape[1]
    idx = np.minimum(pos[:, None] + np.arange(taps)[None, :], plane.shape[1] - 1)
    acc = (plane.astype(np.int64)[:, idx] * coef[None, :, :]).sum(axis=2)
    return np.minimum(acc >> 7, 32767)


def fit(t, out, lo_c, hi_c):
    t = t.ravel().astype(np.int64)
    out = out.ravel().astype(np.int64)
    keep = (out > 0) & (out < 255)                   # clipped outputs bound one side only: leave them out
    t, out = t[keep], out[keep]
    # per distinct t one output
    order = np.argsort(t, kind="stable")
    t, out = t[order], out[order]
    first = np.concatenate([[True], t[1:] != t[:-1]])
    t, out = t[first], out[first]
    lo = (128 * out - 64) << 14
    hi = ((128 * out + 64) << 14) - 1
    found = []
    for c in range(lo_c, hi_c):
        a = int((lo - t * c).max())
        b = int((hi - t * c).min())
        if a <= b:
            found.append((c, a, b))
    return len(t), found


def main():
    sw, sh, dw = 300, 256, 4001
    y = np.add.outer(np.arange(sh), np.arange(sw)) % 256
    cw, ch = (sw + 1) // 2, (sh + 1) // 2
    u = np.add.outer(2 * np.arange(ch), np.arange(cw)) % 256
    v = (u + 77) % 256
    src = [y.astype(np.uint8), u.astype(np.uint8), v.astype(np.uint8)]
    ps = port.PortScaler(sw, sh, 0, dw, sh, 0, F.SWS_BILINEAR)
    t_lum = intermediates(ps.table(0), src[0])
    t_chr = [intermediates(ps.table(1), src[1]), intermediates(ps.table(1), src[2])]
    plain = ref_swscale.RefScaler(sw, sh, 0, dw, sh, 0, F.SWS_BILINEAR).scale(src)
    assert np.array_equal(plain[0], np.clip((t_lum + 64) >> 7, 0, 255)), "intermediate model wrong"
    for name, ranges, (cl, cc) in (("full -> limited", (1, 0), (14071, 14392)), ("limited -> full", (0, 1), (19077, 18651))):
        got = ref_swscale.RefScaler(sw, sh, 0, dw, sh, 0, F.SWS_BILINEAR, ranges=ranges).scale(src)
        n, f = fit(t_lum, got[0], cl - 40, cl + 40)
        print(name, "luma  :", n, "distinct t;", f)
        n, f = fit(np.concatenate([t_chr[0].ravel(), t_chr[1].ravel()]), np.concatenate([got[1].ravel(), got[2].ravel()]), cc - 40, cc + 40)
        print(name, "chroma:", n, "distinct t;", f)


if __name__ == "__main__":
    main()
