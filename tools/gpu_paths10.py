"""slab vs stream on 10-bit planar down-scales (device-resident, CUDA events)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from thinvids_b200 import lib as tv_lib  # noqa: E402
from thinvids_b200.swscale import Scaler  # noqa: E402

shapes = [(1920, 1080, 1280, 720), (1280, 720, 854, 480), (3840, 2160, 1280, 720), (1920, 1080, 854, 480), (1440, 1080, 960, 720),
          (3840, 2160, 1920, 1080), (2560, 1440, 1920, 1080), (2560, 1440, 1280, 720), (1920, 800, 1152, 480)]
for sw, sh, dw, dh in shapes:
    for dfmt in ("nv12", "yuv420p10le"):
        out = []
        for path in ("slab", "stream"):
            with Scaler(sw, sh, "yuv420p10le", dw, dh, dfmt, "bicubic") as s:
                try:
                    s.set_path(path)
                except tv_lib.TvscaleError:
                    out.append((path, None))
                    continue
                n = max(8, min(96, int(3e8 / (sw * sh * 3))))
                src = torch.randint(0, 256, (n, s.src_layout.frame_bytes), dtype=torch.uint8, device="cuda")
                src &= torch.tensor([255, 3], dtype=torch.uint8, device="cuda").repeat(s.src_layout.frame_bytes // 2)
                dst = torch.empty((n, s.dst_layout.frame_bytes), dtype=torch.uint8, device="cuda")
                for _ in range(2):
                    s.scale_packed_device(src, dst)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(5):
                    s.scale_packed_device(src, dst)
                b.record()
                torch.cuda.synchronize()
                out.append((path, round(n / (a.elapsed_time(b) / 5) * 1e3)))
        print(f"{sw}x{sh} -> {dw}x{dh} {dfmt}", out, flush=True)
