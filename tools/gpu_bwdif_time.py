"""Device-resident timing of tvs_bwdif_batch (CUDA events): python tools/gpu_bwdif_time.py [w h n_frames fmt]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from thinvids_b200 import frames as F  # noqa: E402
from thinvids_b200 import lib as tv_lib  # noqa: E402

w, h, n = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (720, 480, 600)
fmt = F.fmt_code(sys.argv[4]) if len(sys.argv) > 4 else F.YUV420P
lib = tv_lib.load()
fb = F.frame_bytes(fmt, w, h)
bps = 1 if F.bit_depth(fmt) == 8 else 2
src = torch.randint(0, 256, (n, fb), dtype=torch.uint8, device="cuda")
if bps == 2:
    src &= torch.tensor([255, 3], dtype=torch.uint8, device="cuda").repeat(fb // 2)
dst = torch.empty_like(src)
cw, ch = (w + 1) // 2, (h + 1) // 2
offs = [0, w * h * bps, w * h * bps + cw * ch * bps]
P3, D3 = ctypes.c_void_p * 3, ctypes.c_ssize_t * 3
sp = P3(*[src.data_ptr() + o for o in offs]); dp = P3(*[dst.data_ptr() + o for o in offs])
pitch = D3(w * bps, cw * bps, cw * bps); fs = D3(fb, fb, fb)


def run():
    rc = lib.tvs_bwdif_batch(0, fmt, w, h, sp, pitch, fs, dp, pitch, fs, n, 1, 3, None)
    assert rc >= 0, tv_lib.last_error()


for _ in range(3):
    run()
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
reps = 10
for _ in range(reps):
    run()
b.record()
torch.cuda.synchronize()
ms = a.elapsed_time(b) / reps
print(f"bwdif {w}x{h} {F.NAME_BY_FMT[fmt]} {n} frames: {ms:.3f} ms per batch, {n / ms * 1e3:,.0f} frames/s, "
      f"{2 * n * fb / ms / 1e6:,.0f} GB/s of frame-in + frame-out")
