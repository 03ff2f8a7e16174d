"""Randomised parity campaign: random sizes (small to full HD), formats, algorithms and input statistics through every
kernel path, compared bit for bit with the oracle port.  Meant for a GPU box:  python tools/gpu_fuzz.py [seconds] [seed]

Inputs alternate between uniform noise, two-level noise (only 0 and the format's maximum: the worst case for clamps and
for the float sums' error bounds) and smooth ramps."""
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

from oracle import port  # noqa: E402
from thinvids_b200 import frames as F  # noqa: E402
from thinvids_b200 import lib as tv_lib  # noqa: E402
from thinvids_b200.swscale import Scaler  # noqa: E402


def make_input(fmt, w, h, kind, seed):
    rng = np.random.default_rng(seed)
    hi = 255 if F.bit_depth(fmt) == 8 else 1023
    planes = []
    for rows, cols in F.plane_shapes(fmt, w, h):
        if kind == 0:
            p = rng.integers(0, hi + 1, (rows, cols))
        elif kind == 1:
            p = rng.integers(0, 2, (rows, cols)) * hi
        else:
            p = (np.add.outer(np.arange(rows) * 3, np.arange(cols) * 2) + seed) % (hi + 1)
        p = p.astype(F.sample_dtype(fmt))
        planes.append(p << 6 if F.fmt_code(fmt) == F.P010LE else p)
    return planes


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    rnd = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
    algs = [F.SWS_BILINEAR, F.SWS_BICUBIC, F.SWS_AREA, F.SWS_LANCZOS]
    t0, n, bad = time.time(), 0, 0
    while time.time() - t0 < budget:
        sf = rnd.choice([0, 0, 1, 2, 3, 4])
        df = rnd.randrange(4)
        big = rnd.random() < 0.25
        sw, sh = (rnd.randrange(200, 1930), rnd.randrange(100, 1090)) if big else (rnd.randrange(2, 400), rnd.randrange(2, 300))
        if sf == 4:
            sw += sw & 1
        mode = rnd.random()
        if mode < 0.15:
            dw, dh = sw, sh
        elif mode < 0.3:
            dw, dh = max(2, sw // 2), max(2, sh // 2)
        elif mode < 0.45:
            dw, dh = sw * 2, sh * 2
        else:
            dw, dh = (rnd.randrange(100, 2200), rnd.randrange(60, 1300)) if big else (rnd.randrange(2, 500), rnd.randrange(2, 400))
        alg = rnd.choice(algs)
        kind = rnd.randrange(3)
        if n < int(os.environ.get("FUZZ_SKIP", "0")):   # fast-forward the case stream
            n += 1
            continue
        src = make_input(sf, sw, sh, kind, n)
        try:
            ref = port.PortScaler(sw, sh, sf, dw, dh, df, alg)
            want = [np.full_like(p, 0x5a) for p in F.alloc_frame(df, dw, dh)]
            ref.scale_into(src, want)
        except Exception as e:  # shapes the oracle refuses are refused by the product too
            continue
        for path in ("auto", "two-pass", "fused", "stream"):
            if os.environ.get("FUZZ_VERBOSE"):
                print("case", n, (sf, sw, sh, df, dw, dh, hex(alg)), path, flush=True)
            try:
                with Scaler(sw, sh, sf, dw, dh, df, alg) as s:
                    try:
                        s.set_path(path)
                    except tv_lib.TvscaleError:
                        continue
                    got = [np.full_like(p, 0x5a) for p in F.alloc_frame(df, dw, dh)]
                    s.scale(src, got)
                    used = s.path
            except tv_lib.TvscaleError as e:
                print("refused", (sf, sw, sh, df, dw, dh, hex(alg)), str(e)[:80])
                break
            for k, (a, b) in enumerate(zip(got, want)):
                if k == 1 and df == F.P010LE and sf in (F.YUV420P, F.YUV420P10LE) and sw % 2 and (sw, sh) == (dw, dh):
                    a, b = a[:, :-2], b[:, :-2]   # the one pair libswscale's same-size planar -> P010 converter leaves unwritten
                if not np.array_equal(a, b):
                    bad += 1
                    print("MISMATCH", (sf, sw, sh, df, dw, dh, hex(alg)), path, used, "plane", k, int((a != b).sum()))
        n += 1
        if n % 50 == 0:
            print("progress", n, "cases", bad, "mismatches", round(time.time() - t0), "s", flush=True)
    print("cases:", n, "mismatches:", bad, "seconds:", round(time.time() - t0, 1), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
