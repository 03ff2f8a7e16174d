"""Host-side logic that needs no GPU: the mirror of the reference's filter-string interface, segment sharding,
and the multi-process (gloo, world_size 2) plumbing bench.py uses for its barrier / max-over-ranks timing."""
import os
import socket
import subprocess
import sys
import textwrap

import pytest

from thinvids_b200 import farm
from thinvids_b200 import filtergraph as fg
from thinvids_b200 import frames as F

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_filter_strings_match_reference_text():
    # worker/tasks.py:62-65, 444-449, 452-461
    assert fg.software_scale_filter(720, deinterlace=False) == "scale=-2:720"
    assert fg.software_scale_filter(480, deinterlace=True) == "bwdif=mode=send_frame:parity=auto:deint=all,scale=-2:480"
    assert fg.vaapi_scale_filter(1080) == "scale=-2:1080,format=nv12,hwupload"
    assert fg.vaapi_scale_filter(999) == fg.SCALE_FILTER_720
    assert fg.vaapi_scale_filter(576).startswith("bwdif=")
    assert fg.reset_segment_video_pts_filter("scale=-2:720") == "setpts=PTS-STARTPTS,scale=-2:720"
    assert fg.reset_segment_video_pts_filter("  ") == "setpts=PTS-STARTPTS"


def test_target_height_policy():
    assert fg.normalize_target_height("720") == 720
    assert fg.normalize_target_height("721") == fg.DEFAULT_TARGET_HEIGHT
    assert fg.normalize_target_height(None) == fg.DEFAULT_TARGET_HEIGHT
    dvd = {"filename": "dvd/Some Film/title.mkv", "source_codec": "mpeg2video", "source_resolution": "720x480"}
    assert fg.dvd_native_target_height(dvd) == 480 and fg.effective_target_height_for_job(dvd) == (480, True)
    pal = dict(dvd, source_resolution="720 x 576", filename="/srv/Movies/x.vob")
    assert fg.dvd_native_target_height(pal) == 576
    assert fg.dvd_native_target_height(dict(dvd, source_codec="h264")) is None
    assert fg.dvd_native_target_height(dict(dvd, source_resolution="1920x1080")) is None
    assert fg.dvd_native_target_height(dict(dvd, filename="tv/show.mkv")) is None
    hd = {"filename": "tv/show.mkv", "target_height": "720", "source_resolution": "1920x1080"}
    assert fg.effective_target_height_for_job(hd) == (720, False)
    assert fg.encode_filtergraph(dict(hd, software_encode=1)) == "setpts=PTS-STARTPTS,scale=-2:720"
    assert fg.encode_filtergraph(hd) == "setpts=PTS-STARTPTS,scale=-2:720,format=nv12,hwupload"


def test_scale_stage_from_reference_graphs():
    st = fg.ScaleStage.from_filtergraph("setpts=PTS-STARTPTS,scale=-2:1080,format=nv12,hwupload", 720, 480)
    assert (st.dst_w, st.dst_h, st.dst_fmt, st.flags, st.hwupload) == (1620, 1080, "nv12", F.SWS_BICUBIC, True)
    st = fg.ScaleStage.from_filtergraph("setpts=PTS-STARTPTS,scale=-2:720", 1920, 1080)
    assert (st.dst_w, st.dst_h, st.dst_fmt) == (1280, 720, "yuv420p")
    # tools/upscale_benchmark.py:163,182
    st = fg.ScaleStage.from_filtergraph("scale=-2:1080:flags=lanczos,format=yuv420p", 1440, 960)
    assert (st.dst_w, st.dst_h, st.flags) == (1620, 1080, F.SWS_LANCZOS)
    st = fg.ScaleStage.from_filtergraph("scale=1920:1080:flags=lanczos+accurate_rnd+bitexact", 720, 480)
    assert st.flags == F.SWS_LANCZOS | F.SWS_ACCURATE_RND | F.SWS_BITEXACT and (st.dst_w, st.dst_h) == (1920, 1080)
    st = fg.ScaleStage.from_filtergraph("scale=w=640:h=-2", 1920, 1080)
    assert (st.dst_w, st.dst_h) == (640, 360)
    assert fg.ScaleStage.from_filtergraph("setpts=PTS-STARTPTS", 720, 480).is_identity


def test_upscale_benchmark_filter_text_on_rgb_frames():
    """tools/upscale_benchmark.py:163,182 -- PNG (rgb24) frames into `scale=-2:H:flags=lanczos,format=...`."""
    st = fg.ScaleStage.from_filtergraph("scale=-2:1080:flags=lanczos,format=yuv420p", 720, 480, "rgb24")
    assert (st.src_fmt, st.dst_w, st.dst_h, st.dst_fmt, st.flags) == ("rgb24", 1620, 1080, "yuv420p", F.SWS_LANCZOS)
    st = fg.ScaleStage.from_filtergraph("scale=-2:720:flags=lanczos,format=nv12,hwupload", 1920, 1080, "rgb24")
    assert (st.dst_w, st.dst_h, st.dst_fmt, st.hwupload) == (1280, 720, "nv12", True)
    with pytest.raises(fg.FilterGraphError, match="explicit format"):
        fg.ScaleStage.from_filtergraph("scale=-2:720", 1920, 1080, "rgb24")


def test_scale_stage_rejects_what_it_cannot_do():
    with pytest.raises(fg.FilterGraphError, match="bwdif"):
        fg.ScaleStage.from_filtergraph(fg.SCALE_FILTER_480, 720, 480)
    st = fg.ScaleStage.from_filtergraph(fg.SCALE_FILTER_480, 720, 480, deinterlaced_upstream=True)
    assert st.deinterlace and (st.dst_w, st.dst_h, st.dst_fmt) == (720, 480, "nv12")
    for bad in ("scale=-2:720,drawtext=text=x", "scale=-2:-2", "scale=iw/2:720", "scale=-2:720:flags=spline",
                "format=rgb24", "[in]scale=2:2[out]", "scale=1:1,scale=2:2"):
        with pytest.raises(fg.FilterGraphError):
            fg.ScaleStage.from_filtergraph(bad, 1920, 1080)


def test_scale_stage_params():
    st = fg.ScaleStage.from_filtergraph("scale=-2:720:flags=bicubic:param0=0.333:param1=0.25", 1920, 1080)
    assert st.params == (0.333, 0.25) and st.flags == F.SWS_BICUBIC
    assert fg.ScaleStage.from_filtergraph("scale=-2:720:flags=lanczos:param0=4", 1920, 1080).params == (4.0, None)
    assert fg.ScaleStage.from_filtergraph("scale=-2:720", 1920, 1080).params == (None, None)
    with pytest.raises(fg.FilterGraphError):
        fg.ScaleStage.from_filtergraph("scale=-2:720:param0=abc", 1920, 1080)


def test_scale_stage_ranges():
    """yuvj420p sources are full-range yuv420p; scale keeps the range unless out_range= changes it (vf_scale)."""
    st = fg.ScaleStage.from_filtergraph("setpts=PTS-STARTPTS,scale=-2:1080,format=nv12,hwupload", 720, 480, "yuvj420p")
    assert (st.src_fmt, st.src_range, st.dst_fmt, st.dst_range) == ("yuv420p", "full", "nv12", "full")
    assert st.src_raw_args == ["-pix_fmt", "yuvj420p"] and st.dst_raw_args == ["-color_range", "pc", "-pix_fmt", "nv12"]
    st = fg.ScaleStage.from_filtergraph("scale=-2:720:out_range=tv", 1920, 1080, "yuvj420p")
    assert (st.src_range, st.dst_range, st.dst_fmt) == ("full", "limited", "yuv420p") and st.dst_raw_args == ["-pix_fmt", "yuv420p"]
    st = fg.ScaleStage.from_filtergraph("scale=-2:720:in_range=limited:out_range=full,format=yuvj420p", 1920, 1080)
    assert (st.src_range, st.dst_range, st.dst_fmt) == ("limited", "full", "yuv420p")
    assert not fg.ScaleStage.from_filtergraph("scale=out_range=pc", 64, 48).is_identity     # same size, but a range change
    assert fg.ScaleStage.from_filtergraph("setpts=PTS-STARTPTS", 64, 48, "yuvj420p").is_identity
    with pytest.raises(fg.FilterGraphError):
        fg.ScaleStage.from_filtergraph("scale=-2:720:out_range=bogus", 1920, 1080)
    with pytest.raises(fg.FilterGraphError):
        fg.ScaleStage.from_filtergraph("scale=-2:720:out_range=full,format=yuv420p", 1920, 1080, "rgb24")


def test_shard_segments_is_a_partition():
    for n in (0, 1, 7, 64):
        for world in (1, 2, 4, 8):
            parts = [farm.shard_segments(n, r, world) for r in range(world)]
            assert sorted(i for p in parts for i in p) == list(range(n))
            assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
    with pytest.raises(ValueError):
        farm.shard_segments(4, 2, 2)


def test_segment_farm_runs_every_segment_once():
    class FakeScaler:
        def __init__(self, dev):
            self.dev, self.closed = dev, False

        def close(self):
            self.closed = True

    made = []

    def make(dev):
        made.append(FakeScaler(dev))
        return made[-1]

    res = farm.SegmentFarm(make, [0, 1, 2]).run(10, lambda s, i: (240, (s.dev, i * i)))
    assert [r.index for r in res] == list(range(10)) and all(r.frames == 240 for r in res)
    assert all(r.output == (r.device, r.index ** 2) for r in res) and all(s.closed for s in made)
    with pytest.raises(RuntimeError):
        farm.SegmentFarm(make, [0, 1]).run(4, lambda s, i: (_ for _ in ()).throw(RuntimeError("boom")))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def test_two_rank_gloo_sharding_and_max_timing(tmp_path):
    """world_size 2 on CPU: each rank takes its stripe of segments, results are joined on the host, and the
    timing reduction is the MAX over ranks — the same plumbing bench.py runs over NCCL."""
    script = tmp_path / "rank.py"
    script.write_text(textwrap.dedent(f"""
        import os, sys, json
        sys.path.insert(0, {ROOT!r})
        import torch, torch.distributed as dist
        from thinvids_b200.farm import shard_segments
        dist.init_process_group("gloo")
        rank, world = dist.get_rank(), dist.get_world_size()
        mine = shard_segments(9, rank, world)
        dist.barrier()
        elapsed = torch.tensor([0.25 + rank], dtype=torch.float64)     # rank 1 is the slow one
        dist.all_reduce(elapsed, op=dist.ReduceOp.MAX)
        counts = torch.tensor([len(mine) * 240], dtype=torch.int64)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
        gathered = [None] * world
        dist.all_gather_object(gathered, mine)
        if rank == 0:
            print(json.dumps({{"max_s": float(elapsed), "frames": int(counts), "segments": gathered}}))
        dist.destroy_process_group()
    """))
    port = _free_port()
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                          "--master-addr", "127.0.0.1", "--master-port", str(port), str(script)],
                         capture_output=True, text=True, timeout=180)
    assert out.returncode == 0, out.stderr[-2000:]
    import json
    line = [ln for ln in out.stdout.splitlines() if ln.startswith("{")][-1]
    got = json.loads(line)
    assert got["max_s"] == 1.25 and got["frames"] == 9 * 240
    assert got["segments"] == [[0, 2, 4, 6, 8], [1, 3, 5, 7]]
