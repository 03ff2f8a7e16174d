"""Small parity sweep meant to run under `compute-sanitizer --tool memcheck`: every kernel path on odd sizes, odd
pitches and all format pairs, few frames.  Prints OK/FAIL per case; the sanitizer reports out-of-bounds accesses."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from oracle import port
from thinvids_b200 import frames as F
from thinvids_b200 import lib as tv_lib
from thinvids_b200.swscale import Scaler

cases = []
for sf in range(4):
    for df in range(4):
        cases.append((sf, 97, 53, df, 131, 77, F.SWS_LANCZOS))
        cases.append((sf, 211, 119, df, 67, 41, F.SWS_BICUBIC))
cases += [(0, 720, 480, 0, 1920, 1080, F.SWS_LANCZOS), (1, 640, 360, 1, 320, 180, F.SWS_AREA), (0, 300, 200, 1, 300, 200, F.SWS_BICUBIC),
          (3, 257, 131, 1, 257, 131, F.SWS_BICUBIC), (0, 1810, 64, 0, 905, 32, F.SWS_BICUBIC)]
# widths whose planes keep 4-byte row alignment: the raw-byte horizontal pass of the streaming kernel
for sf in (0, 1):
    for df in (0, 1):
        cases += [(sf, 216, 120, df, 68, 40, F.SWS_LANCZOS), (sf, 1928, 136, df, 1284, 90, F.SWS_BICUBIC), (sf, 648, 360, df, 220, 122, F.SWS_AREA),
                  (sf, 1000, 96, df, 334, 46, F.SWS_BILINEAR), (sf, 3840, 64, df, 1280, 22, F.SWS_LANCZOS)]
bad = 0
for sf, sw, sh, df, dw, dh, alg in cases:
    src = [F.random_frame(sf, sw, sh, 9), F.random_frame(sf, sw, sh, 10)]
    want = [port.PortScaler(sw, sh, sf, dw, dh, df, alg).scale(fr) for fr in src]
    for path in ("auto", "two-pass", "fused", "stream", "slab"):
        with Scaler(sw, sh, sf, dw, dh, df, alg) as s:
            try:
                s.set_path(path)
            except tv_lib.TvscaleError:
                continue
            got = [s.scale(fr) for fr in src]
            import torch
            lay_s, lay_d = s.src_layout, s.dst_layout
            d_src = torch.from_numpy(lay_s.pack(src)).cuda()
            d_dst = torch.zeros((2, lay_d.frame_bytes), dtype=torch.uint8, device="cuda")
            s.scale_packed_device(d_src, d_dst)
            torch.cuda.synchronize()
            got2 = lay_d.unpack(d_dst.cpu().numpy())
        ok = all(np.array_equal(a, b) for g, w in zip(got, want) for a, b in zip(g, w)) and \
             all(np.array_equal(a, b) for g, w in zip(got2, want) for a, b in zip(g, w))
        bad += not ok
        print("OK  " if ok else "FAIL", sf, (sw, sh), "->", df, (dw, dh), hex(alg), path)
print("failures:", bad)
sys.exit(1 if bad else 0)
