"""tvs_scale_batch_static: the entry point for callers that cannot allocate device memory while scaling.

Some users of the library run where a cudaMalloc at scale time is not an option: the GPU's memory was handed out in
full up front (decoder surface pools, encoder buffers, another framework's arena), or the frame loop cannot take the
device-wide stall an allocation implies.  For them the two-pass kernels keep their intermediate in a fixed array
inside the library.  This file checks, through the C ABI, that the result is byte-identical to tvs_scale_batch on
two calls in a row (the array is shared and keeps the previous call's contents), and that a request larger than the
array is refused with a clear message.  It is named to be collected last.
"""
import ctypes

import numpy as np
import pytest

from thinvids_b200 import frames as F

pytestmark = pytest.mark.gpu

CASES = [
    ("yuv420p", 720, 480, "yuv420p", 1920, 1080, "lanczos", 3),     # C1 shape, 3 frames
    ("yuv420p10le", 320, 180, "nv12", 214, 120, "bicubic", 5),      # 10-bit in, dithered NV12 out
    ("nv12", 128, 96, "p010le", 128, 96, "bicubic", 2),             # same size, depth change
]


def _run_pair(lib, sc, d_src, n):
    import torch
    from thinvids_b200 import lib as tv_lib
    ref = torch.full((n, sc.dst_layout.frame_bytes), 0xA5, dtype=torch.uint8, device="cuda")
    got = torch.full((n, sc.dst_layout.frame_bytes), 0x5A, dtype=torch.uint8, device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    sp, ss, sf = sc.src_layout.pointer_args(d_src.data_ptr())
    for entry, out in ((lib.tvs_scale_batch, ref), (lib.tvs_scale_batch_static, got)):
        dp, ds, df = sc.dst_layout.pointer_args(out.data_ptr())
        rc = entry(sc._ctx, sp, ss, sf, dp, ds, df, n, st)
        assert rc == sc.dst_h, tv_lib.last_error()
    torch.cuda.synchronize()
    return ref, got


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%s_%dx%d-%s_%dx%d-%s" % c[:7])
def test_static_scratch_entry_matches_allocating_entry_twice(case):
    import torch
    from thinvids_b200 import lib as tv_lib
    from thinvids_b200.swscale import Scaler
    sf, sw, sh, df, dw, dh, alg, n = case
    lib = tv_lib.load()
    with Scaler(sw, sh, sf, dw, dh, df, alg) as sc:
        for call in range(2):   # second call: different frames, the array still holds the first call's intermediate
            frames = [F.random_frame(sf, sw, sh, 4000 + 17 * call + i) for i in range(n)]
            d_src = torch.from_numpy(sc.src_layout.pack(frames)).cuda()
            ref, got = _run_pair(lib, sc, d_src, n)
            assert torch.equal(ref, got), f"call {call}: {int((ref != got).sum())} bytes differ"
            assert not bool((got == 0x5A).all())


def test_static_scratch_entry_rejects_oversized_batches():
    import torch
    from thinvids_b200 import lib as tv_lib
    from thinvids_b200.swscale import Scaler
    lib = tv_lib.load()
    with Scaler(1920, 1080, "yuv420p", 3840, 2160, "yuv420p", "bicubic") as sc:
        n = 3   # 3 x 12.4 MB of intermediate > 32 MiB; 2 frames fit
        d_src = torch.zeros((n, sc.src_layout.frame_bytes), dtype=torch.uint8, device="cuda")
        d_dst = torch.zeros((n, sc.dst_layout.frame_bytes), dtype=torch.uint8, device="cuda")
        with pytest.raises(tv_lib.TvscaleError) as e:
            sc.scale_packed_device_static(d_src, d_dst)
        assert e.value.code == -7 and "at most 2 frame" in str(e.value)   # -E2BIG
        sc.scale_packed_device_static(d_src[:2], d_dst[:2])
        torch.cuda.synchronize()


def test_static_scratch_entry_takes_no_device_memory():
    """The fixed array is brought onto the device when the context is created, so a scale call through the
    no-allocation entry point leaves the device's free memory exactly where it was (cudaMemGetInfo)."""
    import torch
    from thinvids_b200 import lib as tv_lib
    from thinvids_b200.swscale import Scaler
    lib = tv_lib.load()
    sf, sw, sh, df, dw, dh, alg, n = CASES[0]
    with Scaler(sw, sh, sf, dw, dh, df, alg) as sc:
        frames = [F.random_frame(sf, sw, sh, 4100 + i) for i in range(n)]
        d_src = torch.from_numpy(sc.src_layout.pack(frames)).cuda()
        out = torch.zeros((n, sc.dst_layout.frame_bytes), dtype=torch.uint8, device="cuda")
        st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
        sp, ss, sfs = sc.src_layout.pointer_args(d_src.data_ptr())
        dp, ds, dfs = sc.dst_layout.pointer_args(out.data_ptr())
        # load the two-pass kernels' code first (lazy module loading), through the allocating entry point
        assert lib.tvs_set_path(sc._ctx, 1) == 0
        assert lib.tvs_scale_batch(sc._ctx, sp, ss, sfs, dp, ds, dfs, n, st) == sc.dst_h, tv_lib.last_error()
        torch.cuda.synchronize()
        free_before, _ = torch.cuda.mem_get_info()
        assert lib.tvs_scale_batch_static(sc._ctx, sp, ss, sfs, dp, ds, dfs, n, st) == sc.dst_h, tv_lib.last_error()
        torch.cuda.synchronize()
        free_after, _ = torch.cuda.mem_get_info()
        assert free_after == free_before, f"free device memory changed by {free_before - free_after} bytes across the call"
