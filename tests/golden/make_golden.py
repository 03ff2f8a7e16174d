"""Regenerate tests/golden/*.json|*.npz from the GENUINE libswscale in this image.

    python tests/golden/make_golden.py

The reference has no tests and pins nothing at this boundary, so the golden
vectors are outputs of the library the reference's ffmpeg child runs (libswscale 9.1.100, flags
alg|accurate_rnd|bitexact) on the seeded inputs of SURVEY.md Appendix B.2.  The hashes SURVEY.md
lists are embedded below and asserted while generating, so a drift in either direction is caught.
The files written here are committed; tests never need the library to be present.
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_swscale  # noqa: E402
from thinvids_b200 import frames as F  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

# (src_fmt, sw, sh, dst_fmt, dw, dh, alg, input, sha256 listed in SURVEY.md Appendix B.2 or None)
SURVEY_CASES = [
    ("yuv420p", 720, 480, "yuv420p", 1920, 1080, "lanczos", "seed:1234", "e7c1f77518590fa3b4041cd9292c144a6c63506541aaa7560a0c688c70785e1d"),
    ("yuv420p", 1920, 1080, "yuv420p", 3840, 2160, "bicubic", "seed:1234", "e37e726823681b019aa96bc477b96501befec5a5b597e8ac767e4d66ca66d5a5"),
    ("nv12", 3840, 2160, "nv12", 1920, 1080, "area", "seed:1234", "63a303db8002dd6f51c333e48f7d8e1b618b08193f60056262c2c6c2fbd5ed7d"),
    ("yuv420p10le", 3840, 2160, "yuv420p10le", 1920, 1080, "lanczos", "seed:1234", "db46a4ef92b3f2122a5053718fa48cafa40f0b86c3418ae2becd3d1624cec49e"),
    ("yuv420p", 1920, 1080, "yuv420p", 3840, 2160, "lanczos", "seed:1234", "d5d85c6cbc30ae28d65e4b3dc249c6d19de3843fec6705404071aa1c08d18e2b"),
    ("yuv420p", 720, 480, "yuv420p", 1620, 1080, "lanczos", "seed:1234", "b1d5f4b037d8c869bad6a266da8332fd09c860c243d850c69ebdf7f1d1808843"),
    ("yuv420p", 720, 480, "nv12", 1620, 1080, "bicubic", "seed:1234", "49fcb762e7671f29952b8001af8e55db0858ede88574d3ff5d1a57e757b7bc1f"),
    ("yuv420p", 853, 481, "yuv420p", 1279, 719, "lanczos", "seed:1234", "180eac7b82a8c2bc9699b1c8d75c64a5c2e0f1d4b8c3e861561f62bf4d3b5f92"),
    ("yuv420p", 1920, 1080, "yuv420p", 1280, 720, "lanczos", "seed:1234", "cc7c5f45977a5245ce6f9d2acb07cc958f57c6eae2288128e26f55c75c71b8fd"),
    ("yuv420p", 1920, 1080, "nv12", 1280, 720, "bicubic", "seed:1234", "6965a0497fd498f27b5c0d4cdb4fa5badad540d6c8fa0d7f1526e33016c363ca"),
    ("yuv420p", 1920, 1080, "yuv420p", 854, 480, "bicubic", "seed:1234", "124189fc41333ddb660e6146fd21c12b6110a5b0c954c4cc41825540238381d3"),
    ("yuv420p", 720, 576, "yuv420p", 1350, 1080, "lanczos", "seed:1234", "26638b86e2936f41ea551809baad02446834b26f5254d2bef9fbbc6de4ee90a2"),
    ("yuv420p", 1440, 1080, "yuv420p", 1920, 1080, "lanczos", "seed:1234", "2ddbc76e6feb9a699207668ee72df6084ba5c38bd6657dd2b7f191b5f43da32c"),
    ("yuv420p", 1920, 1080, "yuv420p", 1920, 540, "area", "seed:1234", "65c6152b318361dabd5e2ac5f942511ef6da1e412d00e0e0a20924dbeb386233"),
    ("nv12", 1920, 1080, "nv12", 3840, 2160, "lanczos", "seed:1234", "632faf291cbd4f4415c54c3519581641dda48d9d7c07ba257567091f7b207f0b"),
    ("nv12", 1920, 1080, "yuv420p", 3840, 2160, "area", "seed:1234", "b9a7dee824b7bed685e255042d86c7c864163c8e2af9019c8ee72c906fd52080"),
    ("p010le", 3840, 2160, "p010le", 1920, 1080, "lanczos", "seed:1234", "a0c7590005e7584d24d5ae2983a0c0b2de557a6917f7f0d3be4a55d0e306543e"),
    ("p010le", 1920, 1080, "yuv420p10le", 1280, 720, "bicubic", "seed:1234", "0f5484570b6e08c053aa97f96ed25b1b11d154fc4a7117bdcac61347302c56ed"),
    ("yuv420p10le", 1920, 1080, "yuv420p", 1280, 720, "lanczos", "seed:1234", "7aacb3007f99e504391fdfcfb8aa6403ac1eb88375bbb96fac2b48f58419db06"),
    ("yuv420p10le", 1920, 1080, "nv12", 1280, 720, "bicubic", "seed:1234", "18871a8e33cb78706ad5ee3cae0ef0384f71c3067d3791cdb2536f683ad606eb"),
    ("yuv420p", 3840, 2160, "yuv420p", 1920, 1080, "area", "seed:1234", "71bdcce4195a7a23453ca692ec6bddf92b716663656d24eefee9772b441ec7f8"),
    ("yuv420p", 3840, 2160, "yuv420p", 1280, 720, "area", "seed:1234", "08eb6a961ae732ca1ff8317abb0e52f5254c095903eda8fa383e7167af57c7a8"),
    ("yuv420p", 3840, 2160, "yuv420p", 640, 360, "bicubic", "seed:1234", "85fb6a1cdf89d5ca2774fb70c1a899d1013f1e881ef95ee9a71f3c144056236b"),
    ("yuv420p", 720, 480, "nv12", 720, 480, "bicubic", "seed:1234", "553a785bf318c9a9686c25848ab6a51998633add13a8fdd24faa75a68ffa049d"),
    ("yuv420p10le", 1920, 1080, "yuv420p", 1920, 1080, "bicubic", "seed:1234", "f17d45f8cebe372be0524008719d018ea915d2334556df1eca8392f13bbbf7c8"),
    ("yuv420p10le", 1920, 1080, "nv12", 1920, 1080, "bicubic", "seed:1234", "ca0a0a2c497b55a3a40287cc57dde665baf1f59f3cb8f86383e7e9a61a3fc069"),
    ("yuv420p", 720, 480, "yuv420p", 1920, 1080, "lanczos", "seed:1235", "68cc699196a9121cc80244b78285c08fe814a1dafc8eb8f9d3758dc3353b5ad3"),
]
_C1 = ("yuv420p", 720, 480, "yuv420p", 1920, 1080, "lanczos")
_C4 = ("yuv420p10le", 3840, 2160, "yuv420p10le", 1920, 1080, "lanczos")
_C3 = ("nv12", 3840, 2160, "nv12", 1920, 1080, "area")
_ADV = {
    _C1: ["67c001205c7a86bf2000bf8c1c03c49d3a53c38d3e44373db53dcd82c2b6712f", "2127dcaad72f780791e8cdc34fa6259cebdcf7549aade50d382563081f55d45c",
          "e55913d5b98b8634a9c9088b0da1efe0695f0f178eeab183ee4d48aa661ab588", "96a0e9e706bf9394a5f0c7c1cc30c43813a29a1858ce06b5104375515e5ab80f",
          "28f91e132843ba2f714e634f8103d5c9acb57b662f3b648efe540c3a1ddecf50"],
    _C4: ["caf2de99cac6a8710d8770589d96fefbdd101e1a35cdc97d9f7a42765831ab62", "08c2e09feec8991116db0b2154f70a87c2bac57624a6bc11c71ab407db84806b",
          "18069b0363c5eb0340553d1c7748575cd803f5c64cc783be1fd67b95f1d7ee33", "fb8a965156866e3b4e5d049e33707a292bdbf8ace7be62a6dbcdf54cf663e900",
          "5998fc46e64415c445edafc0b8ce21637d65a4410c91996eefbe779e4046684e"],
    _C3: ["d06243806eba1e0c13ab561e9c7d98a613d8234caeb95bb75f9d6d314512ea85", "00d2ed0a6c04c0d97f3a8085194d956a6892c87f1657ae422307dbd924f68040",
          "3d0dadb2aa8c061d1b087947ba58e596cea67d04b89b59472d1eebdd5ebf9b64", "5d1a495345e9d24277fa52523a528914d239dadc590da6d41550853001e1ffb6",
          "28f91e132843ba2f714e634f8103d5c9acb57b662f3b648efe540c3a1ddecf50"],
}
for cfg, hashes in _ADV.items():
    for pat, hsh in zip(F.PATTERNS, hashes):
        SURVEY_CASES.append(cfg + ("pattern:" + pat, hsh))

# Additional vectors not listed in SURVEY.md (format pairs, odd sizes, library quirks found while
# pinning the oracle): hash recorded from the genuine library, no prior expectation.
EXTRA_CASES = [
    ("p010le", 1920, 1080, "nv12", 1920, 1080, "bicubic", "seed:1234", None),     # stored-plane dither copy
    ("p010le", 1620, 1080, "nv12", 1620, 1080, "bicubic", "seed:1234", None),     # ... with the cols%8 tail quirk
    ("p010le", 1920, 1080, "nv12", 1280, 720, "lanczos", "seed:1234", None),
    ("p010le", 1920, 1080, "yuv420p", 1280, 720, "bicubic", "seed:1234", None),
    ("nv12", 1280, 720, "p010le", 1920, 1080, "bicubic", "seed:1234", None),
    ("yuv420p", 1280, 720, "yuv420p10le", 1920, 1080, "lanczos", "seed:1234", None),
    ("yuv420p", 853, 481, "p010le", 853, 481, "bicubic", "seed:1234", None),      # odd width: last chroma pair untouched (dst prefilled 0)
    ("yuv420p", 1920, 1080, "yuv420p", 1920, 1080, "bicubic", "seed:1234", None),
    ("yuv420p", 720, 480, "yuv420p", 1620, 1080, "bilinear", "seed:1234", None),
    ("yuv420p", 1920, 1080, "yuv420p", 1280, 720, "bilinear", "seed:1234", None),
    ("yuv420p", 720, 480, "yuv420p", 1620, 1080, "area", "seed:1234", None),
    ("nv12", 721, 479, "yuv420p", 333, 777, "lanczos", "seed:1234", None),
    ("yuv420p10le", 641, 361, "p010le", 1281, 723, "bicubic", "seed:1234", None),
    # packed RGB sources: what tools/upscale_benchmark.py:159-182 feeds its lanczos stage
    ("rgb24", 720, 480, "yuv420p", 1920, 1080, "lanczos", "seed:1234", None),
    ("rgb24", 720, 480, "nv12", 1620, 1080, "lanczos", "seed:1234", None),
    ("rgb24", 1920, 1080, "yuv420p", 1280, 720, "lanczos", "seed:1234", None),    # half-width chroma input
    ("rgb24", 1920, 1080, "nv12", 1280, 720, "bicubic", "seed:1234", None),
    ("rgb24", 640, 360, "yuv420p10le", 1280, 720, "bicubic", "seed:1234", None),
    ("rgb24", 720, 480, "yuv420p", 720, 480, "bicubic", "seed:1234", None),       # same size still runs the scaler
    ("rgb24", 722, 479, "p010le", 333, 777, "area", "seed:1234", None),
    ("rgb24", 720, 480, "yuv420p", 1920, 1080, "lanczos", "pattern:checker", None),
    ("rgb24", 720, 480, "yuv420p", 1920, 1080, "lanczos", "pattern:white", None),
    # bgr24 (OpenCV's order) and 16-bit RGB (rgb48be is what a 16-bit PNG decodes to): round 2
    ("bgr24", 720, 480, "nv12", 1620, 1080, "bicubic", "seed:1234", None),
    ("bgr24", 1920, 1080, "yuv420p", 1280, 720, "lanczos", "seed:1234", None),
    ("rgb48be", 720, 480, "yuv420p", 1920, 1080, "lanczos", "seed:1234", None),
    ("rgb48be", 720, 480, "yuv420p", 1920, 1080, "lanczos", "seed:1235", None),
    ("rgb48le", 1920, 1080, "nv12", 1280, 720, "bicubic", "seed:1234", None),
    ("rgb48le", 640, 360, "yuv420p10le", 1280, 720, "bicubic", "seed:1234", None),
    ("rgb48be", 722, 479, "p010le", 333, 777, "area", "seed:1234", None),
]

# Small cases whose full output is stored (tests/golden/small_cases.npz).
SMALL_CASES = [
    ("yuv420p", 64, 48, "yuv420p", 160, 108, "lanczos"),
    ("yuv420p", 67, 35, "yuv420p", 101, 77, "bicubic"),
    ("yuv420p", 160, 120, "nv12", 64, 48, "area"),
    ("nv12", 96, 64, "nv12", 48, 32, "lanczos"),
    ("nv12", 48, 32, "yuv420p", 96, 64, "area"),
    ("yuv420p10le", 128, 96, "yuv420p10le", 64, 48, "lanczos"),
    ("yuv420p10le", 64, 48, "yuv420p", 96, 72, "bicubic"),
    ("p010le", 64, 48, "nv12", 64, 48, "bicubic"),
    ("p010le", 76, 20, "nv12", 76, 20, "bicubic"),
    ("p010le", 64, 48, "p010le", 128, 96, "bilinear"),
    ("yuv420p", 64, 48, "nv12", 64, 48, "bicubic"),
    ("yuv420p10le", 67, 35, "yuv420p", 67, 35, "bicubic"),
    ("yuv420p", 16, 16, "yuv420p", 8, 8, "bicubic"),
    ("yuv420p", 9, 7, "yuv420p", 31, 23, "lanczos"),
    ("rgb24", 64, 48, "yuv420p", 160, 108, "lanczos"),
    ("rgb24", 96, 64, "nv12", 48, 32, "bicubic"),
    ("rgb24", 32, 18, "yuv420p", 32, 18, "bicubic"),
    ("rgb24", 40, 30, "p010le", 50, 44, "bilinear"),
]


def make_input(spec: str, fmt: str, w: int, h: int):
    kind, arg = spec.split(":")
    return F.random_frame(fmt, w, h, int(arg)) if kind == "seed" else F.pattern_frame(fmt, w, h, arg)


def main() -> None:
    print(ref_swscale.version())
    out = []
    for sf, sw, sh, df, dw, dh, alg, spec, expect in SURVEY_CASES + EXTRA_CASES:
        src = make_input(spec, sf, sw, sh)
        r = ref_swscale.RefScaler(sw, sh, F.FMT_BY_NAME[sf], dw, dh, F.FMT_BY_NAME[df], F.ALG_BY_NAME[alg])
        got = F.frame_sha256(r.scale(src))
        r.close()
        if expect is not None and got != expect:
            raise SystemExit(f"hash drift vs SURVEY.md for {(sf, sw, sh, df, dw, dh, alg, spec)}: {got}")
        out.append({"src_fmt": sf, "src_w": sw, "src_h": sh, "dst_fmt": df, "dst_w": dw, "dst_h": dh, "alg": alg,
                    "input": spec, "sha256": got, "in_survey": expect is not None})
        print(sf, sw, sh, "->", df, dw, dh, alg, spec, got[:16], "survey-listed" if expect else "new")
    with open(os.path.join(HERE, "swscale_hashes.json"), "w") as f:
        json.dump({"library": ref_swscale.version(), "flags": "alg|SWS_ACCURATE_RND|SWS_BITEXACT",
                   "dst_prefill": 0, "cases": out}, f, indent=1)

    blobs = {}
    meta = []
    for i, (sf, sw, sh, df, dw, dh, alg) in enumerate(SMALL_CASES):
        src = F.random_frame(sf, sw, sh, 100 + i)
        r = ref_swscale.RefScaler(sw, sh, F.FMT_BY_NAME[sf], dw, dh, F.FMT_BY_NAME[df], F.ALG_BY_NAME[alg])
        dst = r.scale(src)
        r.close()
        for k, p in enumerate(dst):
            blobs[f"case{i}_dst{k}"] = p
        meta.append({"src_fmt": sf, "src_w": sw, "src_h": sh, "dst_fmt": df, "dst_w": dw, "dst_h": dh, "alg": alg,
                     "seed": 100 + i, "n_dst_planes": len(dst)})
    np.savez_compressed(os.path.join(HERE, "small_cases.npz"), **blobs)
    with open(os.path.join(HERE, "small_cases.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print("wrote", len(out), "hash cases and", len(meta), "small cases")


if __name__ == "__main__":
    main()
