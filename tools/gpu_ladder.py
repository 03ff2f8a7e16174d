"""Which kernel the worker's ladder shapes get, and how fast it is (device-resident, CUDA events):
every source size x target height {480, 720, 1080} of worker/tasks.py:62-65 (scale=-2:H), nv12 (VAAPI) and yuv420p (x264)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import json  # noqa: E402

import torch  # noqa: E402

from thinvids_b200 import frames as F  # noqa: E402
from thinvids_b200.filtergraph import ScaleStage  # noqa: E402
from thinvids_b200.swscale import Scaler  # noqa: E402

PEAK = 6494.3
try:
    PEAK = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_copy_gbs"]["burst"]
except Exception:
    pass
sources = [(640, 360), (854, 480), (720, 480), (720, 576), (1280, 720), (1440, 1080), (1920, 1080), (2560, 1440), (3840, 2160), (1920, 800), (4096, 2160)]
rows = []
for sw, sh in sources:
    for th in (480, 720, 1080):
        for vf, sfmt in ((f"scale=-2:{th},format=nv12,hwupload", "yuv420p"), (f"scale=-2:{th}", "yuv420p"), (f"scale=-2:{th},format=nv12,hwupload", "yuv420p10le")):
            st = ScaleStage.from_filtergraph(vf, sw, sh, sfmt, default_dst_fmt="yuv420p" if "format" not in vf else None)
            if st.is_identity:
                continue
            n = max(8, min(128, int(3e8 / (sw * sh * 1.5 + st.dst_w * st.dst_h * 1.5))))
            with Scaler(st.src_w, st.src_h, st.src_fmt, st.dst_w, st.dst_h, st.dst_fmt, st.flags) as s:
                src = torch.randint(0, 256, (n, s.src_layout.frame_bytes), dtype=torch.uint8, device="cuda")
                if sfmt.endswith("10le"):
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
                ms = a.elapsed_time(b) / 5
                gbs = n * (s.src_layout.frame_bytes + s.dst_layout.frame_bytes) / ms / 1e6
                rows.append((f"{sw}x{sh} {sfmt}", f"{st.dst_w}x{st.dst_h} {st.dst_fmt}", s.path, round(n / ms * 1e3), round(gbs / PEAK, 3)))
                print(*rows[-1], flush=True)
slow = [r for r in rows if r[2] == "two-pass" or r[4] < 0.08]
print("slow or two-pass:", slow)
