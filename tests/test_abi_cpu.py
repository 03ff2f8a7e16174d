"""CPU-side checks of the product: the C-ABI library loads and exports what include/tvscale.h declares,
and its host-side table generator matches the golden table hashes (no GPU, no compute calls)."""
import ctypes
import hashlib
import json
import os
import re

import numpy as np
import pytest

from thinvids_b200 import build as tv_build
from thinvids_b200 import frames as F
from thinvids_b200 import lib as tv_lib
from thinvids_b200 import swscale

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TABLES = json.load(open(os.path.join(ROOT, "tests", "golden", "survey_tables.json")))


@pytest.fixture(scope="module")
def lib():
    tv_build.build()
    return tv_lib.load()


def test_header_and_library_agree_on_exports(lib):
    header = open(os.path.join(ROOT, "include", "tvscale.h")).read()
    declared = set(re.findall(r"TVS_API[^;]*?\b(tvs_[a-z_]+)\s*\(", header))
    assert declared == set(tv_lib.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name


def test_version_and_error_strings(lib):
    assert lib.tvs_version().decode().startswith("tvscale")
    assert isinstance(tv_lib.last_error(), str)


@pytest.mark.parametrize("name", sorted(TABLES["configs"]))
def test_host_tables_match_golden_hashes(lib, name):
    ent = TABLES["configs"][name]
    _, sw, sh, _, dw, dh, alg = ent["args"]
    axes = [(sw, dw, 1 << 14), ((sw + 1) // 2, (dw + 1) // 2, 1 << 14), (sh, dh, 1 << 12), ((sh + 1) // 2, (dh + 1) // 2, 1 << 12)]
    for which, (s, d, one) in enumerate(axes):
        coef, pos = swscale.filter_table(s, d, alg, one)
        assert coef.shape == (d, ent["taps"][which])
        digest = hashlib.sha256(coef.tobytes() + pos.tobytes()).hexdigest()
        assert digest == ent["sha256"][which], (name, which)


def test_table_rows_sum_to_one_and_stay_in_range(lib):
    for s, d, alg in [(720, 1620, "bicubic"), (1920, 1280, "lanczos"), (3840, 640, "bicubic"), (853, 1279, "bilinear"),
                      (1080, 1080, "area"), (7, 31, "lanczos")]:
        for one in (1 << 14, 1 << 12):
            coef, pos = swscale.filter_table(s, d, alg, one)
            assert (coef.astype(np.int64).sum(axis=1) == one).all()
            assert pos.min() >= 0 and (np.diff(pos) >= 0).all()
            # taps beyond the source must carry zero weight
            over = pos[:, None] + np.arange(coef.shape[1])[None, :] >= s
            assert (coef[over] == 0).all()


def test_table_errors(lib):
    with pytest.raises(tv_lib.TvscaleError):
        swscale.filter_table(4000, 8, "lanczos", 1 << 14)   # > 256 taps: libswscale would cascade
    with pytest.raises(tv_lib.TvscaleError):
        swscale.filter_table(100, 50, 0x400, 1 << 14)       # unknown algorithm
    coef = np.empty(4, dtype=np.int16)
    pos = np.empty(4, dtype=np.int32)
    rc = lib.tvs_filter_table(100, 50, 4, 1 << 14, 128, 128, coef.ctypes.data, coef.size, pos.ctypes.data, pos.size)
    assert rc == -28  # -ENOSPC


def test_create_without_gpu_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(tv_lib.TvscaleError) as e:
        swscale.Scaler(720, 480, "yuv420p", 1920, 1080, "yuv420p", "lanczos")
    assert "no CPU path" in str(e.value)


def test_packed_layout_round_trip():
    for fmt in ("yuv420p", "nv12", "yuv420p10le", "p010le"):
        lay = swscale.PackedLayout(F.fmt_code(fmt), 37, 21)
        frames = [F.random_frame(fmt, 37, 21, s) for s in (1, 2)]
        packed = lay.pack(frames)
        assert packed.shape == (2, F.frame_bytes(fmt, 37, 21))
        back = lay.unpack(packed)
        for a, b in zip(frames, back):
            for x, y in zip(a, b):
                assert np.array_equal(x, y)


def test_scaled_width_matches_reference_shapes():
    # scale=-2:H of worker/tasks.py:62-65: 720x480 -> 1620x1080, 1920x1080 -> 1280x720
    assert F.scaled_width(720, 480, 1080) == 1620
    assert F.scaled_width(1920, 1080, 720) == 1280
    assert F.scaled_width(720, 576, 1080) == 1350
    assert F.CONFIGS["C1"].algorithmic_bytes == 3628800
    assert F.CONFIGS["C2"].algorithmic_bytes == 15552000
    assert F.CONFIGS["C4"].algorithmic_bytes == 31104000


def test_libswscale_api_shim_loads_and_exports_the_trio():
    """thinvids_b200/libswscale_tvs.so: sws_getContext / sws_scale / sws_freeContext on top of libtvscale (no compute here)."""
    import ctypes
    import os
    from thinvids_b200 import build as tv_build
    tv_build.build()
    if not os.path.exists(tv_build.SHIM_LIB):
        tv_build.build_shim()
    shim = ctypes.CDLL(tv_build.SHIM_LIB)
    for name in ("sws_getContext", "sws_scale", "sws_freeContext", "tvs_shim_map_pix_fmt"):
        assert hasattr(shim, name), name
    shim.sws_getContext.restype = ctypes.c_void_p
    shim.sws_getContext.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p] * 3
    assert not shim.sws_getContext(64, 48, 9999, 32, 24, 0, 4, None, None, None)      # unknown pixel format: NULL, like libswscale
    assert not shim.sws_getContext(64, 48, 0, 32, 24, 0, 4, ctypes.c_void_p(1), None, None)   # source filters are not supported
    assert shim.tvs_shim_map_pix_fmt(9999, 0) == 0
    shim.sws_freeContext.argtypes = [ctypes.c_void_p]
    shim.sws_freeContext(None)


def test_python_constants_match_the_header():
    """Format codes, flag bits and info keys are spelled twice (include/tvscale.h, thinvids_b200/{frames,lib}.py): keep them equal."""
    import os
    import re
    from thinvids_b200 import frames as F
    from thinvids_b200 import lib as tv_lib
    text = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "tvscale.h")).read()
    vals = {m.group(1): int(m.group(2), 0) for m in re.finditer(r"\b(TVS_[A-Z0-9_]+)\s*=\s*(0x[0-9a-fA-F]+|\d+)", text)}
    for name, code in (("YUV420P", F.YUV420P), ("NV12", F.NV12), ("YUV420P10LE", F.YUV420P10LE), ("P010LE", F.P010LE), ("RGB24", F.RGB24),
                       ("YUVJ420P", tv_lib.FMT_YUVJ420P), ("BGR24", F.BGR24), ("RGB48LE", F.RGB48LE), ("RGB48BE", F.RGB48BE),
                       ("BGR48LE", F.BGR48LE), ("BGR48BE", F.BGR48BE)):
        assert vals["TVS_FMT_" + name] == code, name
    assert (vals["TVS_BILINEAR"], vals["TVS_BICUBIC"], vals["TVS_AREA"], vals["TVS_LANCZOS"]) == (F.SWS_BILINEAR, F.SWS_BICUBIC, F.SWS_AREA, F.SWS_LANCZOS)
    assert (vals["TVS_ACCURATE_RND"], vals["TVS_BITEXACT"]) == (F.SWS_ACCURATE_RND, F.SWS_BITEXACT)
    assert (vals["TVS_SRC_FULL_RANGE"], vals["TVS_DST_FULL_RANGE"]) == (tv_lib.SRC_FULL_RANGE, tv_lib.DST_FULL_RANGE)
    assert [vals[k] for k in ("TVS_INFO_KERNEL_LAUNCHES", "TVS_INFO_PATH", "TVS_INFO_SCRATCH_BYTES", "TVS_INFO_SRC_FRAME_BYTES",
                              "TVS_INFO_DST_FRAME_BYTES")] == [tv_lib.INFO_KERNEL_LAUNCHES, tv_lib.INFO_PATH, tv_lib.INFO_SCRATCH_BYTES,
                                                               tv_lib.INFO_SRC_FRAME_BYTES, tv_lib.INFO_DST_FRAME_BYTES]
    assert float(re.search(r"#define TVS_PARAM_DEFAULT\s+([0-9.]+)", text).group(1)) == tv_lib.PARAM_DEFAULT
