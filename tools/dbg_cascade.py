import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from thinvids_b200 import frames as F
from thinvids_b200.swscale import Scaler
from oracle import ref_swscale, port
mk = ref_swscale.RefScaler if ref_swscale.available() else port.PortScaler
cases = [(4, 1, 306, 172, 114, 3, F.SWS_BILINEAR), (4, 3, 134, 199, 122, 2, F.SWS_BICUBIC), (4, 3, 234, 148, 6, 395, F.SWS_LANCZOS),
         (4, 2, 52, 192, 100, 5, F.SWS_BICUBIC), (4, 0, 332, 222, 3, 216, F.SWS_BICUBIC)]
for c in cases:
    sf, df, sw, sh, dw, dh, alg = c
    print("case", c, flush=True)
    src = F.random_frame(sf, sw, sh, 5)
    want = mk(sw, sh, sf, dw, dh, df, alg).scale(src)
    with Scaler(sw, sh, sf, dw, dh, df, alg) as s:
        print("  path", s.path, flush=True)
        got = s.scale(src)
    print("  diffs", [int((a != b).sum()) for a, b in zip(got, want)], flush=True)
