"""Pins the oracle (CPU checker): C restatement vs golden vectors vs the genuine libswscale.

No GPU, no product code: this file is what lets the `-m gpu` parity tests trust oracle/.
"""
import hashlib
import json
import os
import random

import numpy as np
import pytest

from oracle import port, ref_swscale
from thinvids_b200 import frames as F

GOLD = os.path.join(os.path.dirname(__file__), "golden")
HASHES = json.load(open(os.path.join(GOLD, "swscale_hashes.json")))["cases"]
SMALL = json.load(open(os.path.join(GOLD, "small_cases.json")))
TABLES = json.load(open(os.path.join(GOLD, "survey_tables.json")))


def _make_input(spec, fmt, w, h):
    kind, arg = spec.split(":")
    return F.random_frame(fmt, w, h, int(arg)) if kind == "seed" else F.pattern_frame(fmt, w, h, arg)


def _port(c):
    return port.PortScaler(c["src_w"], c["src_h"], F.FMT_BY_NAME[c["src_fmt"]], c["dst_w"], c["dst_h"],
                           F.FMT_BY_NAME[c["dst_fmt"]], F.ALG_BY_NAME[c["alg"]])


def _case_id(c):
    return "%s_%dx%d-%s_%dx%d-%s-%s" % (c["src_fmt"], c["src_w"], c["src_h"], c["dst_fmt"], c["dst_w"], c["dst_h"],
                                        c["alg"], c.get("input", c.get("seed")))


@pytest.mark.parametrize("name", sorted(TABLES["configs"]))
def test_port_tables_match_survey_hashes(name):
    ent = TABLES["configs"][name]
    sf, sw, sh, df, dw, dh, alg = ent["args"]
    p = port.PortScaler(sw, sh, F.FMT_BY_NAME[sf], dw, dh, F.FMT_BY_NAME[df], F.ALG_BY_NAME[alg])
    for which in range(4):
        coef, pos = p.table(which)
        assert coef.shape[1] == ent["taps"][which]
        digest = hashlib.sha256(coef.astype(np.int16).tobytes() + pos.astype(np.int32).tobytes()).hexdigest()
        assert digest == ent["sha256"][which], (name, which)


def test_port_table_spot_values():
    p = port.PortScaler(720, 480, 0, 1920, 1080, 0, F.SWS_LANCZOS)
    h, hpos = p.table(0)
    v, vpos = p.table(2)
    s = TABLES["spot"]
    assert h[:10].tolist() == s["C1_hLum_first10"] and h[-3:].tolist() == s["C1_hLum_last3"]
    assert v[:10].tolist() == s["C1_vLum_first10"] and v[-3:].tolist() == s["C1_vLum_last3"]
    assert (h.sum(axis=1) == 16384).all() and (v.sum(axis=1) == 4096).all()
    assert hpos[-1] == 714 and vpos[-1] == 474
    p2 = port.PortScaler(1920, 1080, 0, 3840, 2160, 0, F.SWS_BICUBIC)
    h2, _ = p2.table(0)
    assert h2[:6].tolist() == s["C2_hLum_first6"] and h2[-2:].tolist() == s["C2_hLum_last2"]


@pytest.mark.parametrize("case", HASHES, ids=_case_id)
def test_port_matches_golden_hashes(case):
    src = _make_input(case["input"], case["src_fmt"], case["src_w"], case["src_h"])
    assert F.frame_sha256(_port(case).scale(src)) == case["sha256"]


def test_port_matches_small_golden_outputs():
    blobs = np.load(os.path.join(GOLD, "small_cases.npz"))
    for i, c in enumerate(SMALL):
        src = F.random_frame(c["src_fmt"], c["src_w"], c["src_h"], c["seed"])
        got = _port(c).scale(src)
        for k in range(c["n_dst_planes"]):
            assert np.array_equal(got[k], blobs[f"case{i}_dst{k}"]), (_case_id(c), k)


needs_ref = pytest.mark.skipif(not ref_swscale.available(), reason="genuine libswscale not loadable here")


@needs_ref
def test_reference_library_is_the_pinned_build():
    assert ref_swscale.version().startswith("libswscale 9.1.100")


@needs_ref
def test_port_equals_genuine_library_on_random_shapes():
    rnd = random.Random(11)
    algs = [F.SWS_BILINEAR, F.SWS_BICUBIC, F.SWS_AREA, F.SWS_LANCZOS]
    for i in range(80):
        sf, df = rnd.randrange(4), rnd.randrange(4)
        sw, sh = rnd.randrange(8, 160), rnd.randrange(8, 100)
        dw = sw if rnd.random() < 0.2 else rnd.randrange(8, 240)
        dh = sh if rnd.random() < 0.2 else rnd.randrange(8, 160)
        alg = rnd.choice(algs)
        src = F.random_frame(sf, sw, sh, i)
        a = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        b = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        ref_swscale.RefScaler(sw, sh, sf, dw, dh, df, alg).scale_into(src, a)
        port.PortScaler(sw, sh, sf, dw, dh, df, alg).scale_into(src, b)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (sf, df, sw, sh, dw, dh, hex(alg))


@needs_ref
def test_port_equals_genuine_library_rgb24_sources():
    """Packed RGB in (libswscale's input stage: rgb24ToY_c, rgb24ToUV_c / _half_c, then hScale16To15 with shift 13):
    both chroma input widths, every destination format, same-size included.  Even widths only -- with an odd width the
    library itself reads past the row."""
    rnd = random.Random(23)
    algs = [F.SWS_BILINEAR, F.SWS_BICUBIC, F.SWS_AREA, F.SWS_LANCZOS]
    for i in range(60):
        sw, sh = 2 * rnd.randrange(1, 80), rnd.randrange(2, 100)
        dw, dh = (sw, sh) if rnd.random() < 0.15 else (rnd.randrange(2, 240), rnd.randrange(2, 160))
        alg, df = rnd.choice(algs), rnd.randrange(4)
        cp = rnd.choice([None, None, (0, -513, 0, -513), (0, 128, 0, 128), (64, 192, 0, 64)])   # explicit chroma siting too
        # bgr24 (OpenCV's order): the same input stage with R and B swapped in memory; rgb48le / rgb48be (16-bit PNGs): 16-bit
        # lines, shift 15, ordered dither on 8-bit output
        sf = (F.RGB24, F.RGB48LE, F.BGR24, F.BGR48LE, F.RGB48BE, F.BGR48BE)[i % 6]
        src = F.random_frame(sf, sw, sh, 500 + i)
        want = ref_swscale.RefScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=cp).scale(src)
        got = port.PortScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=cp).scale(src)
        for x, y in zip(got, want):
            assert np.array_equal(x, y), (sf, sw, sh, dw, dh, df, hex(alg))


@needs_ref
def test_port_equals_genuine_library_unscaled_all_format_pairs():
    for sf in range(4):
        for df in range(4):
            for w, h in [(64, 48), (67, 35), (76, 20), (2, 2)]:
                src = F.random_frame(sf, w, h, 7)
                a = [np.full_like(p, 77) for p in F.alloc_frame(df, w, h)]
                b = [np.full_like(p, 77) for p in F.alloc_frame(df, w, h)]
                ref_swscale.RefScaler(w, h, sf, w, h, df, F.SWS_BICUBIC).scale_into(src, a)
                port.PortScaler(w, h, sf, w, h, df, F.SWS_BICUBIC).scale_into(src, b)
                for x, y in zip(a, b):
                    assert np.array_equal(x, y), (sf, df, w, h)


@needs_ref
def test_port_equals_genuine_library_with_explicit_chroma_siting():
    for cp in [(0, -513, 0, -513), (0, 128, 0, 128), (0, 0, 0, 0), (0, 128, 128, 128)]:
        src = F.random_frame(0, 320, 240, 5)
        a = ref_swscale.RefScaler(320, 240, 0, 480, 270, 1, F.SWS_BICUBIC, chr_pos=cp).scale(src)
        b = port.PortScaler(320, 240, 0, 480, 270, 1, F.SWS_BICUBIC, chr_pos=cp).scale(src)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), cp


@needs_ref
def test_port_cascade_equals_genuine_library():
    """libswscale's cascade (filters that reach 170 taps after padding): boundary cases on both axes, every format,
    explicit chroma sitings (which the inner contexts ignore), and the refusal when the source is under 4x the destination."""
    rnd = random.Random(5)
    cases = [(0, 0, 399, 133, 9, 133, 4, None), (0, 3, 399, 133, 9, 133, 4, None), (0, 0, 16, 399, 16, 9, 4, None),
             (2, 1, 800, 64, 12, 40, 0x200, None), (1, 0, 1000, 50, 20, 60, 4, None), (0, 0, 380, 16, 9, 16, 4, None),
             (0, 0, 390, 16, 9, 16, 4, None), (0, 0, 16, 380, 16, 9, 4, None), (0, 0, 16, 384, 16, 9, 4, None), (0, 0, 16, 388, 16, 9, 4, None),
             (0, 0, 16, 392, 16, 9, 4, None), (3, 2, 2000, 40, 16, 50, 0x200, None), (0, 0, 4000, 8, 8, 8, 0x200, None),
             (0, 0, 399, 133, 9, 133, 4, (0, 128, 0, 128)), (0, 1, 399, 133, 9, 133, 4, (256, 128, 0, 0)),
             # rgb24 sources cascade like any other (found by tools/gpu_fuzz.py: the port used to skip the check for them)
             (4, 1, 306, 172, 114, 3, 2, None), (4, 3, 134, 199, 122, 2, 4, None), (4, 3, 234, 148, 6, 395, 0x200, None),
             (4, 2, 52, 192, 100, 5, 4, None), (4, 0, 332, 222, 3, 216, 4, None), (4, 0, 700, 16, 4, 800, 4, None)]
    for _ in range(20):
        sf, df = rnd.randrange(4), rnd.randrange(4)
        sw, dw, sh, dh = rnd.randrange(300, 1500), rnd.randrange(2, 16), rnd.randrange(8, 40), rnd.randrange(8, 60)
        if rnd.random() < 0.5:
            sw, dw, sh, dh = sh, dh, sw, dw
        cases.append((sf, df, sw, sh, dw, dh, rnd.choice([4, 0x200, 2, 0x20]), None))
    compared = 0
    for sf, df, sw, sh, dw, dh, alg, cp in cases:
        src = F.random_frame(sf, sw, sh, 91)
        try:
            ref = ref_swscale.RefScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=cp)
        except Exception:
            with pytest.raises(ValueError):
                port.PortScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=cp)
            continue
        a = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        b = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        if sf == 4:
            a = ref.scale(src)       # padded buffers for the library's rgb24 input stage
        else:
            ref.scale_into(src, a)
        port.PortScaler(sw, sh, sf, dw, dh, df, alg, chr_pos=cp).scale_into(src, b)
        compared += 1
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (sf, df, sw, sh, dw, dh, hex(alg), cp)
    assert compared >= 25
    with pytest.raises(Exception):
        ref_swscale.RefScaler(700, 16, 0, 4, 800, 0, 4)
    with pytest.raises(ValueError):
        port.PortScaler(700, 16, 0, 4, 800, 0, 4)


@needs_ref
def test_port_range_conversion_equals_genuine_library():
    """Full <-> limited range (yuvj420p sources, ffmpeg's in_range / out_range): the conversion of the 15-bit intermediate
    with libswscale 9.1's constants, for every format pair (the constants follow the destination depth), scaled and
    same-size, three input statistics.  tools/fit_range_convert.py shows how the form and the constants were found."""
    rnd = random.Random(3)
    cases = []
    for sf in range(4):
        for df in range(4):
            for rg in ((1, 0), (0, 1)):
                cases += [(sf, df, 64, 48, 64, 48, 4, rg), (sf, df, 67, 35, 67, 35, 4, rg), (sf, df, 64, 48, 96, 48, 0x200, rg),
                          (sf, df, 64, 48, 64, 30, 2, rg)]
    for _ in range(60):
        cases.append((rnd.randrange(4), rnd.randrange(4), rnd.randrange(2, 300), rnd.randrange(2, 200), rnd.randrange(2, 400),
                      rnd.randrange(2, 300), rnd.choice([2, 4, 0x20, 0x200]), rnd.choice([(1, 0), (0, 1)])))
    cases.append((0, 0, 64, 48, 128, 96, 4, (1, 1)))     # equal ranges: nothing to convert
    for n, (sf, df, sw, sh, dw, dh, alg, rg) in enumerate(cases):
        hi = 255 if sf < 2 else 1023
        rng = np.random.default_rng(n)
        src = []
        for rows, cols in F.plane_shapes(sf, sw, sh):
            kind = n % 3
            pl = (rng.integers(0, hi + 1, (rows, cols)) if kind == 0 else rng.integers(0, 2, (rows, cols)) * hi if kind == 1
                  else np.add.outer(np.arange(rows) * 3, np.arange(cols) * 2) % (hi + 1)).astype(F.sample_dtype(sf))
            src.append(pl << 6 if sf == 3 else pl)
        a = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        b = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        ref_swscale.RefScaler(sw, sh, sf, dw, dh, df, alg, ranges=rg).scale_into(src, a)
        port.PortScaler(sw, sh, sf, dw, dh, df, alg, ranges=rg).scale_into(src, b)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (sf, df, sw, sh, dw, dh, hex(alg), rg)
    # a cascaded context converts no range (its inner contexts are created with the defaults)
    src = F.random_frame(0, 399, 133, 1)
    plain = ref_swscale.RefScaler(399, 133, 0, 9, 133, 0, 4).scale(src)
    for rg in ((1, 0), (0, 1)):
        for got in (ref_swscale.RefScaler(399, 133, 0, 9, 133, 0, 4, ranges=rg).scale(src), port.PortScaler(399, 133, 0, 9, 133, 0, 4, ranges=rg).scale(src)):
            for x, y in zip(got, plain):
                assert np.array_equal(x, y), rg


@needs_ref
def test_port_scaler_params_equal_genuine_library():
    """ffmpeg's scale=...:param0=..:param1=.. (sws_getContext's `param`): bicubic B / C and the lanczos lobe count, which
    also sets the filter's support; a cascade hands them to its inner contexts."""
    rnd = random.Random(4)
    cases = []
    for _ in range(40):
        alg = rnd.choice([4, 0x200])
        pr = rnd.choice([(0, 0.5), (1 / 3, 1 / 3), (1, 0), (0, 0.75), (0.2, None), (None, 0.3), (0, 1)]) if alg == 4 else \
            rnd.choice([(2, None), (4, None), (2.5, None), (5, None), (1, None), (3, None)])
        cases.append((rnd.randrange(4), rnd.randrange(4), rnd.randrange(8, 300), rnd.randrange(8, 200), rnd.randrange(8, 400),
                      rnd.randrange(8, 300), alg, pr))
    cases.append((0, 0, 700, 64, 12, 40, 0x200, (5, None)))      # cascades
    cases.append((0, 1, 64, 48, 128, 96, 2, (0.7, 0.1)))          # bilinear ignores them
    for n, (sf, df, sw, sh, dw, dh, alg, pr) in enumerate(cases):
        src = F.random_frame(sf, sw, sh, n)
        a = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        b = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
        ref_swscale.RefScaler(sw, sh, sf, dw, dh, df, alg, params=pr).scale_into(src, a)
        port.PortScaler(sw, sh, sf, dw, dh, df, alg, params=pr).scale_into(src, b)
        for x, y in zip(a, b):
            assert np.array_equal(x, y), (sf, df, sw, sh, dw, dh, hex(alg), pr)
    src = F.random_frame(0, 64, 48, 1)
    plain = port.PortScaler(64, 48, 0, 128, 96, 0, 4).scale(src)
    assert not np.array_equal(plain[0], port.PortScaler(64, 48, 0, 128, 96, 0, 4, params=(1, 0)).scale(src)[0])
    assert np.array_equal(plain[0], port.PortScaler(64, 48, 0, 128, 96, 0, 4, params=(None, None)).scale(src)[0])


@needs_ref
def test_port_equals_genuine_library_on_p010_words_with_noise_in_the_padding():
    """P010 keeps its 10 bits in the top of the word; decoders do not promise zeros below.  libswscale shifts the padding away
    everywhere except in the same-size p010le -> p010le copy, which keeps it - apart from the first word of every group of four
    and the row tail (restated in copy_plane of the port)."""
    rng = np.random.default_rng(1)
    for df in range(4):
        for sw, sh, dw, dh in [(79, 97, 79, 97), (80, 96, 80, 96), (5, 3, 5, 3), (80, 96, 120, 100), (81, 47, 40, 50)]:
            clean = F.random_frame(F.P010LE, sw, sh, 2)
            dirty = [p | rng.integers(0, 64, p.shape, dtype=np.uint16) for p in clean]
            a = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
            b = [np.full_like(p, 77) for p in F.alloc_frame(df, dw, dh)]
            ref_swscale.RefScaler(sw, sh, F.P010LE, dw, dh, df, 4).scale_into(dirty, a)
            port.PortScaler(sw, sh, F.P010LE, dw, dh, df, 4).scale_into(dirty, b)
            for x, y in zip(a, b):
                assert np.array_equal(x, y), (df, sw, sh, dw, dh)
