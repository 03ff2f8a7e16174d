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


def make_input(fmt, w, h, kind, seed, dirty_padding=True):
    rng = np.random.default_rng(seed)
    hi = {8: 255, 10: 1023, 16: 65535}[F.bit_depth(fmt)]
    planes = []
    for rows, cols in F.plane_shapes(fmt, w, h):
        if kind == 0:
            p = rng.integers(0, hi + 1, (rows, cols))
        elif kind == 1:
            p = rng.integers(0, 2, (rows, cols)) * hi
        else:
            p = (np.add.outer(np.arange(rows) * 3, np.arange(cols) * 2) + seed) % (hi + 1)
        p = p.astype(F.sample_dtype(fmt))
        if F.fmt_code(fmt) == F.P010LE:
            p = p << 6
            if seed % 2 and dirty_padding:      # decoders do not promise zeros in P010's six padding bits: libswscale shifts them away, so must we
                p = p | rng.integers(0, 64, p.shape).astype(p.dtype)
        planes.append(p)
    return planes


def surfaces_case(rnd, sf, sw, sh, df, dw, dh, alg, seed):
    """The same shape through tvs_scale_batch on pitched device surfaces: random pitch padding, random base misalignment,
    a few frames, guard bytes around everything.  Returns a list of problems."""
    import torch
    n = rnd.randrange(1, 4)
    frames = [make_input(sf, sw, sh, rnd.randrange(3), seed + i) for i in range(n)]
    ref = port.PortScaler(sw, sh, sf, dw, dh, df, alg)
    want = [ref.scale(fr) for fr in frames]

    def surfaces(fmt, w, h, fill_from=None):
        bufs, ptrs, pitches, fstrides, metas = [], [], [], [], []
        bps = 1 if F.bit_depth(fmt) == 8 else 2
        for k, (rows, cols) in enumerate(F.plane_shapes(fmt, w, h)):
            row_bytes = cols * bps
            pitch = row_bytes + rnd.choice([0, 0, 4, 16, 1, 3, 6, 64, 10]) * (bps if bps == 2 else 1)
            if bps == 2:
                pitch += pitch & 1
            lead = rnd.choice([0, 0, 16, 4, 2, 1, 8, 6]) * (1 if bps == 1 else 2) // (1 if bps == 1 else 1)
            if bps == 2:
                lead += lead & 1
            fstride = (rows + 1) * pitch
            host = np.full(lead + n * fstride + 32, 0xEE, dtype=np.uint8)
            if fill_from is not None:
                for i in range(n):
                    view = host[lead + i * fstride: lead + i * fstride + rows * pitch].reshape(rows, pitch)
                    view[:, :row_bytes] = np.ascontiguousarray(fill_from[i][k]).view(np.uint8).reshape(rows, row_bytes)
            t = torch.from_numpy(host).cuda()
            bufs.append(t); ptrs.append(t.data_ptr() + lead); pitches.append(pitch); fstrides.append(fstride)
            metas.append((lead, rows, row_bytes, pitch, fstride))
        return bufs, ptrs, pitches, fstrides, metas

    sb, sp, spi, sfs, _ = surfaces(sf, sw, sh, frames)
    db, dp, dpi, dfs, dmeta = surfaces(df, dw, dh)
    problems = []
    with Scaler(sw, sh, sf, dw, dh, df, alg) as s:
        s.scale_device_planes(sp, spi, sfs, dp, dpi, dfs, n, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        used = s.path
    for k, (lead, rows, row_bytes, pitch, fstride) in enumerate(dmeta):
        host = db[k].cpu().numpy()
        if not (host[:lead] == 0xEE).all() or not (host[lead + n * fstride:] == 0xEE).all():
            problems.append(("guard before/after", k, used))
        for i in range(n):
            frame = host[lead + i * fstride: lead + (i + 1) * fstride].reshape(rows + 1, pitch)
            exp = np.ascontiguousarray(want[i][k]).view(np.uint8).reshape(rows, row_bytes)
            got = frame[:rows, :row_bytes]
            if k == 1 and df == F.P010LE and sf in (F.YUV420P, F.YUV420P10LE) and sw % 2 and (sw, sh) == (dw, dh):
                got, exp = got[:, :-4], exp[:, :-4]
            if not np.array_equal(got, exp):
                problems.append(("pixels", k, i, used, int((got != exp).sum())))
            if not (frame[:rows, row_bytes:] == 0xEE).all() or not (frame[rows] == 0xEE).all():
                problems.append(("padding written", k, i, used))
    return problems


def main():
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 120.0
    rnd = random.Random(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
    algs = [F.SWS_BILINEAR, F.SWS_BICUBIC, F.SWS_AREA, F.SWS_LANCZOS]
    t0, n, bad = time.time(), 0, 0
    while time.time() - t0 < budget:
        sf = rnd.choice([0, 0, 1, 2, 3, 4])   # (FUZZ_SF=6 for bgr24, 7 / 8 for rgb48le / rgb48be)
        if os.environ.get("FUZZ_SF"):   # one source format only, e.g. 4 = rgb24
            sf = int(os.environ["FUZZ_SF"])
        df = rnd.randrange(4)
        huge = bool(os.environ.get("FUZZ_HUGE"))   # up to 4K frames, every case large
        big = huge or rnd.random() < 0.25
        sw, sh = (rnd.randrange(200, 3850 if huge else 1930), rnd.randrange(100, 2170 if huge else 1090)) if big \
            else (rnd.randrange(2, 400), rnd.randrange(2, 300))
        if sf in (4, 6, 7, 8, 9, 10):
            sw += sw & 1
        mode = rnd.random()
        if mode < 0.15:
            dw, dh = sw, sh
        elif mode < 0.3:
            dw, dh = max(2, sw // 2), max(2, sh // 2)
        elif mode < 0.45:
            dw, dh = sw * 2, sh * 2
        else:
            dw, dh = (rnd.randrange(100, 4200 if huge else 2200), rnd.randrange(60, 2400 if huge else 1300)) if big \
                else (rnd.randrange(2, 500), rnd.randrange(2, 400))
        alg = rnd.choice(algs)
        chr_pos = rnd.choice([None, None, None, (0, -513, 0, -513), (0, 128, 0, 128), (0, 0, 0, 0), (128, 128, 128, 128), (64, 192, 0, 64)]) \
            if os.environ.get("FUZZ_CHRPOS") else None
        kind = rnd.randrange(3)
        if n < int(os.environ.get("FUZZ_SKIP", "0")):   # fast-forward the case stream
            n += 1
            continue
        src = make_input(sf, sw, sh, kind, n)
        try:
            ref = port.PortScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=chr_pos)
            want = [np.full_like(p, 0x5a) for p in F.alloc_frame(df, dw, dh)]
            ref.scale_into(src, want)
        except Exception as e:  # shapes the oracle refuses are refused by the product too
            continue
        for path in ("auto", "two-pass", "fused", "stream", "slab"):
            if os.environ.get("FUZZ_VERBOSE"):
                print("case", n, (sf, sw, sh, df, dw, dh, hex(alg)), path, flush=True)
            try:
                with Scaler(sw, sh, sf, dw, dh, df, alg, chr_pos=chr_pos) as s:
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
                    print("MISMATCH", (sf, sw, sh, df, dw, dh, hex(alg)), chr_pos, path, used, "plane", k, int((a != b).sum()))
        if os.environ.get("FUZZ_SURFACES") and not (sf == 4 and False):
            try:
                for pr in surfaces_case(rnd, sf, sw, sh, df, dw, dh, alg, 7000 + n):
                    bad += 1
                    print("SURFACE MISMATCH", (sf, sw, sh, df, dw, dh, hex(alg)), pr, flush=True)
            except tv_lib.TvscaleError as e:
                print("surfaces refused", (sf, sw, sh, df, dw, dh, hex(alg)), str(e)[:80])
        n += 1
        if n % 50 == 0:
            print("progress", n, "cases", bad, "mismatches", round(time.time() - t0), "s", flush=True)
    print("cases:", n, "mismatches:", bad, "seconds:", round(time.time() - t0, 1), flush=True)
    sys.exit(1 if bad else 0)


if __name__ == "__main__":
    main()
