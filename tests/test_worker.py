"""The pixel stage of the reference's `encode` task (thinvids_b200/worker.py): control flow on the CPU with a stand-in
scaler (status dicts, halts, heartbeats, filter text: worker/tasks.py:1532-1652), the real thing on the GPU against
the oracle."""
import io

import numpy as np
import pytest

from thinvids_b200 import frames as F
from thinvids_b200.swscale import PackedLayout
from thinvids_b200.worker import Cv2Encoder, EncodeHooks, cv2_decode_yuv420p, encode_part, read_frames


class _StandIn:
    """Records the stage it was built for and returns recognisable bytes of the right size."""

    made = []

    def __init__(self, stage, device):
        self.stage, self.device, self.batches, self.closed = stage, device, [], False
        self.out_bytes = F.frame_bytes(F.FMT_BY_NAME[stage.dst_fmt], stage.dst_w, stage.dst_h)
        _StandIn.made.append(self)

    def scale_packed_host(self, batch):
        self.batches.append(len(batch))
        return np.full((len(batch), self.out_bytes), 7, dtype=np.uint8)

    def close(self):
        self.closed = True


def _raw(fmt, w, h, n, seed=3):
    lay = PackedLayout(fmt, w, h)
    frames = [F.random_frame(fmt, w, h, seed + i) for i in range(n)]
    return frames, lay.pack(frames).tobytes()


def _run(job, data, w, h, **kw):
    _StandIn.made.clear()
    out = io.BytesIO()
    res = encode_part(job, 4, io.BytesIO(data), out, src_w=w, src_h=h, scaler_factory=_StandIn, **kw)
    return res, out.getvalue()


def test_hardware_branch_completes_with_the_references_filter_text():
    _, data = _raw(F.YUV420P, 720, 480, 5)
    beats = []
    res, out = _run({"target_height": "1080", "software_encode": "0"}, data, 720, 480, batch_frames=2,
                    hooks=EncodeHooks(heartbeat=beats.append))
    assert res == {"status": "COMPLETED", "idx": 4, "frames": 5, "filtergraph": "setpts=PTS-STARTPTS,scale=-2:1080,format=nv12,hwupload",
                   "dst_w": 1620, "dst_h": 1080, "dst_fmt": "nv12"}
    assert len(out) == 5 * 1620 * 1080 * 3 // 2 and set(out) == {7}
    sc = _StandIn.made[0]
    assert sc.batches == [2, 2, 1] and sc.closed and sc.device == 0
    assert beats == ["encode_start", "encode", "encode", "encode"]


def test_software_branch_yields_yuv420p_bicubic():
    _, data = _raw(F.YUV420P, 1920, 1080, 1)
    res, out = _run({"target_height": 720, "software_encode": 1}, data, 1920, 1080)
    assert res["status"] == "COMPLETED" and res["filtergraph"] == "setpts=PTS-STARTPTS,scale=-2:720"
    assert (res["dst_w"], res["dst_h"], res["dst_fmt"]) == (1280, 720, "yuv420p")
    assert _StandIn.made[0].stage.flags == F.SWS_BICUBIC and len(out) == 1280 * 720 * 3 // 2


def test_halts_abort_the_part():
    _, data = _raw(F.YUV420P, 64, 48, 6)
    res, out = _run({"target_height": 720}, data, 64, 48, hooks=EncodeHooks(is_halted=lambda: True))
    assert res == {"status": "ABORTED"} and out == b"" and not _StandIn.made
    calls = {"n": 0}

    def halted_after_first_batch():
        calls["n"] += 1
        return calls["n"] > 2   # the two checks before the loop pass, the one after the first batch does not

    res, out = _run({"target_height": 720}, data, 64, 48, batch_frames=4, hooks=EncodeHooks(is_halted=halted_after_first_batch))
    assert res == {"status": "ABORTED"}
    assert _StandIn.made[0].batches == [4] and _StandIn.made[0].closed


def test_failures_are_reported_not_raised():
    _, data = _raw(F.YUV420P, 64, 48, 2)
    res, _ = _run({"target_height": 720}, data[:-10], 64, 48)
    assert res["status"] == "FAILED" and res["stage"] == "encode" and "ended inside a frame" in res["reason"]
    assert res["reason"].startswith("part 4 (encode): ")

    def no_device(stage, device):
        raise RuntimeError("no CUDA device")

    out = io.BytesIO()
    res = encode_part({"target_height": 720}, 1, io.BytesIO(data), out, src_w=64, src_h=48, scaler_factory=no_device)
    assert res["status"] == "FAILED" and "no CUDA device" in res["reason"]


def test_dvd_jobs_need_the_deinterlacer_upstream():
    job = {"filename": "dvd/title.vob", "source_codec": "mpeg2video", "source_resolution": "720x480", "target_height": 1080}
    _, data = _raw(F.YUV420P, 720, 480, 1)
    res, _ = _run(job, data, 720, 480)
    assert res["status"] == "FAILED" and "bwdif" in res["reason"]
    res, out = _run(job, data, 720, 480, deinterlaced_upstream=True)
    assert res["status"] == "COMPLETED" and (res["dst_w"], res["dst_h"], res["dst_fmt"]) == (720, 480, "nv12")
    assert "bwdif=mode=send_frame:parity=auto:deint=all,scale=-2:480,format=nv12,hwupload" in res["filtergraph"]


def test_identity_stage_passes_frames_through():
    _, data = _raw(F.YUV420P, 854, 480, 3)
    res, out = _run({"target_height": 480, "software_encode": 1}, data, 854, 480)
    assert res["status"] == "COMPLETED" and res["frames"] == 3 and out == data and not _StandIn.made


def test_read_frames_batches():
    data = bytes(range(12)) * 5
    got = list(read_frames(io.BytesIO(data), 12, 2))
    assert [len(b) for b in got] == [2, 2, 1] and b"".join(b.tobytes() for b in got) == data
    assert list(read_frames(io.BytesIO(b""), 12, 2)) == []


@pytest.mark.gpu
def test_encode_part_on_the_gpu_matches_the_oracle():
    from oracle import port
    for job, sw, sh in [({"target_height": 720, "software_encode": 0}, 320, 180), ({"target_height": 480, "software_encode": 1}, 1280, 720)]:
        frames, data = _raw(F.YUV420P, sw, sh, 5, seed=40)
        out = io.BytesIO()
        res = encode_part(job, 0, io.BytesIO(data), out, src_w=sw, src_h=sh, batch_frames=2)
        assert res["status"] == "COMPLETED", res
        dfmt = F.FMT_BY_NAME[res["dst_fmt"]]
        lay = PackedLayout(dfmt, res["dst_w"], res["dst_h"])
        got = lay.unpack(np.frombuffer(out.getvalue(), dtype=np.uint8).reshape(5, lay.frame_bytes))
        ref = port.PortScaler(sw, sh, F.YUV420P, res["dst_w"], res["dst_h"], dfmt, F.SWS_BICUBIC)
        for g, fr in zip(got, frames):
            for a, b in zip(g, ref.scale(fr)):
                assert np.array_equal(a, b)


@pytest.mark.gpu
def test_decode_scale_encode_smoke_with_opencv(tmp_path):
    """a real container in, a real container out, frames through the GPU scaler in between."""
    cv2 = pytest.importorskip("cv2")
    src_path, dst_path = str(tmp_path / "in.avi"), str(tmp_path / "out.avi")
    w, h, n = 320, 240, 6
    wr = cv2.VideoWriter(src_path, cv2.VideoWriter_fourcc(*"MJPG"), 24.0, (w, h))
    if not wr.isOpened():
        pytest.skip("OpenCV cannot write MJPG here")
    rng = np.random.default_rng(5)
    for i in range(n):
        img = np.zeros((h, w, 3), np.uint8)
        img[:, :, 0] = np.linspace(0, 255, w, dtype=np.uint8)[None, :]
        img[:, :, 1] = (np.arange(h, dtype=np.uint8) * (i + 1))[:, None]
        img[:, :, 2] = rng.integers(0, 256, (h, w), dtype=np.uint8)
        wr.write(img)
    wr.release()
    data = b"".join(f.tobytes() for f in cv2_decode_yuv420p(src_path))
    assert len(data) == n * w * h * 3 // 2
    sink = Cv2Encoder(dst_path, 960, 720, "nv12", fourcc="MJPG")
    res = encode_part({"target_height": 720}, 2, io.BytesIO(data), sink, src_w=w, src_h=h)
    sink.close()
    assert res["status"] == "COMPLETED" and (res["dst_w"], res["dst_h"], res["dst_fmt"]) == (960, 720, "nv12") and sink.frames == n
    cap = cv2.VideoCapture(dst_path)
    ok, frame = cap.read()
    cap.release()
    assert ok and frame.shape == (720, 960, 3)


# ---- the task with the reference's signature (worker/tasks.py:1354-1366): argv of both children, status dicts ----------
class _FakeProc:
    def __init__(self, argv, stdout=None, stdin=None, stderr=None, rc=0, on_wait=None):
        self.argv, self.stdout, self.stdin, self.stderr, self.rc, self.on_wait = argv, stdout, stdin, stderr, rc, on_wait
        self.terminated = False

    def wait(self, timeout=None):
        if self.on_wait:
            self.on_wait()
        return self.rc

    def terminate(self):
        self.terminated = True

    def kill(self):
        self.terminated = True


def _task_env(tmp_path, job, data, w, h, rc_enc=0, halted=lambda job_id: False, make_output=True, pix_fmt="yuv420p"):
    from thinvids_b200 import worker
    procs, beats, fields = [], [], {}
    sink = io.BytesIO()
    sink.close = lambda: None   # keep the encoder's stdin readable after the task closes it

    def popen(argv, **kw):
        if argv[-1] == "pipe:1":
            pr = _FakeProc(argv, stdout=io.BytesIO(data))
        else:
            def touch():
                if make_output:
                    open(argv[-1], "wb").write(b"mp4")
            pr = _FakeProc(argv, stdin=sink, stderr=io.BytesIO(b"frame=1\nprogress=continue\nprogress=end\n"), rc=rc_enc, on_wait=touch)
        procs.append(pr)
        return pr

    uploads = []
    env = worker.EncodeEnv(job_fields=lambda job_id: job, probe=lambda path: (w, h, pix_fmt, 24.0),
                           set_job_fields=lambda job_id, m: fields.update(m), is_halted=halted,
                           heartbeat=lambda job_id, stage, force=False, note="": beats.append(stage),
                           base_dir=lambda job_id: str(tmp_path / job_id), popen=popen, scaler_factory=_StandIn,
                           fetch_part=lambda job_id, idx, host, path: open(path, "wb").write(b"ts"),
                           upload_result=lambda job_id, idx, host, path: uploads.append((host, os.path.basename(path))))
    worker.configure(env)
    return env, procs, beats, fields, uploads, sink


import os  # noqa: E402


def test_encode_task_vaapi_branch_builds_the_references_commands(tmp_path):
    from thinvids_b200 import worker
    _, data = _raw(F.YUV420P, 720, 480, 3)
    env, procs, beats, fields, uploads, sink = _task_env(tmp_path, {"target_height": "1080", "software_encode": "0"}, data, 720, 480)
    res = worker.encode("job1", 7, "http://master:8000", stitch_host="http://stitch:8000")
    assert res == {"status": "COMPLETED", "idx": 7}
    base = str(tmp_path / "job1")
    dec, enc = procs[0].argv, procs[1].argv
    assert dec == ["ffmpeg", "-hide_banner", "-nostats", "-loglevel", "error", "-fflags", "+genpts", "-i", f"{base}/in_007.ts",
                   "-map", "0:v:0", "-an", "-sn", "-dn", "-vf", "setpts=PTS-STARTPTS", "-f", "rawvideo", "-pix_fmt", "yuv420p", "pipe:1"]
    assert enc == ["ffmpeg", "-hide_banner", "-nostats", "-loglevel", "error", "-y", "-progress", "pipe:2",
                   "-vaapi_device", "/dev/dri/renderD128", "-f", "rawvideo", "-pix_fmt", "nv12", "-s", "1620x1080", "-r", "24.000000",
                   "-i", "pipe:0", "-fflags", "+genpts", "-i", f"{base}/in_007.ts", "-map", "0:v:0", "-map", "1:a?", "-vf", "hwupload",
                   "-c:v", "h264_vaapi", "-rc_mode", "CQP", "-qp", "27", "-c:a", "aac", "-ac", "2", "-b:a", "192k", f"{base}/out_007.mp4"]
    assert len(sink.getvalue()) == 3 * 1620 * 1080 * 3 // 2
    assert fields == {"target_height_effective": 1080, "deinterlace_effective": "0"}
    assert uploads == [("http://stitch:8000", "out_007.mp4")]
    assert beats[0] == "encode_prepare" and "encode_start" in beats and beats[-1] == "upload_start"


def test_encode_task_software_dvd_branch_deinterlaces_in_the_decode_child(tmp_path):
    from thinvids_b200 import worker
    _, data = _raw(F.YUV420P, 720, 480, 2)
    job = {"target_height": "1080", "software_encode": "1", "filename": "/watch/dvd/movie.mkv", "source_codec": "mpeg2video",
           "source_resolution": "720x480"}
    env, procs, *_ = _task_env(tmp_path, job, data, 720, 480)
    res = worker.encode("job2", 0, "http://master:8000", v_sel=1, a_sel=2, source_path="/src/movie.mkv", source_start_s=12.5,
                        source_duration_s=10.0)
    assert res == {"status": "COMPLETED", "idx": 0}
    dec, enc = procs[0].argv, procs[1].argv
    vf = dec[dec.index("-vf") + 1]
    assert vf == "setpts=PTS-STARTPTS,bwdif=mode=send_frame:parity=auto:deint=all"     # DVD sources keep their height and de-interlace
    assert dec[5:11] == ["-ss", "12.500000", "-i", "/src/movie.mkv", "-t", "10.000000"] and dec[dec.index("-map") + 1] == "0:v:1"
    assert enc[enc.index("-s") + 1] == "720x480" and enc[enc.index("-pix_fmt") + 1] == "yuv420p"
    tail = enc[enc.index("pipe:0") + 1:]
    assert tail[:6] == ["-ss", "12.500000", "-i", "/src/movie.mkv", "-t", "10.000000"]
    assert tail[6:] == ["-map", "0:v:0", "-map", "1:a:2?", "-sn", "-dn", "-map_metadata", "-1", "-map_chapters", "-1",
                        "-c:v", "libx264", "-preset", "veryfast", "-crf", "23", "-c:a", "aac", "-ac", "2", "-b:a", "192k",
                        enc[-1]]


def test_encode_task_keeps_full_range_sources_full_range(tmp_path):
    """Phone / camera H.264 and MJPEG decode as yuvj420p.  The reference's scale filter keeps the range (no out_range=), so
    both children are told the frames are full range and the scaler converts nothing (worker/tasks.py:62-65,1565-1580)."""
    from thinvids_b200 import worker
    _, data = _raw(F.YUV420P, 720, 480, 2)
    for soft, want_dst in (("1", ["-pix_fmt", "yuvj420p"]), ("0", ["-color_range", "pc", "-pix_fmt", "nv12"])):
        env, procs, *_ = _task_env(tmp_path, {"target_height": "1080", "software_encode": soft}, data, 720, 480, pix_fmt="yuvj420p")
        res = worker.encode("job9", 1, "http://master:8000")
        assert res == {"status": "COMPLETED", "idx": 1}
        dec, enc = procs[0].argv, procs[1].argv
        assert dec[dec.index("rawvideo") + 1:] == ["-pix_fmt", "yuvj420p", "pipe:1"]
        at = enc.index("rawvideo") + 1
        assert enc[at:at + len(want_dst)] == want_dst and enc[at + len(want_dst)] == "-s"


def test_encode_task_status_dicts(tmp_path):
    from thinvids_b200 import worker
    _, data = _raw(F.YUV420P, 64, 48, 2)
    job = {"target_height": "720", "software_encode": "1"}
    env, procs, *_ = _task_env(tmp_path, job, data, 64, 48)
    env.token_is_current = lambda job_id, tok: tok == "current"
    assert worker.encode("j", 1, "m", run_token="old") == {"status": "STALE", "job_id": "j", "idx": 1}
    env, procs, *_ = _task_env(tmp_path, job, data, 64, 48, halted=lambda job_id: True)
    assert worker.encode("j", 1, "m") == {"status": "ABORTED"} and not procs
    env, procs, *_ = _task_env(tmp_path, job, data, 64, 48, rc_enc=1)
    res = worker.encode("j", 1, "m")
    assert res["status"] == "FAILED" and res["stage"] == "encode" and "ffmpeg rc=1" in res["reason"] and "progress=end" in res["reason"]
    env, procs, *_ = _task_env(tmp_path, job, data, 64, 48, make_output=False)
    assert worker.encode("j", 2, "m")["status"] == "FAILED"
    n = {"calls": 0}

    def halted_late(job_id):
        n["calls"] += 1
        return n["calls"] > 5
    env, procs, *_ = _task_env(tmp_path, job, data, 64, 48, halted=halted_late)
    env.batch_frames = 1
    assert worker.encode("j", 3, "m") == {"status": "ABORTED"} and procs[0].terminated and procs[1].terminated


def test_decoder_surface_descriptor():
    from thinvids_b200.worker import DecoderSurface
    s = DecoderSurface(dev_ptr=0x1000, pitch=2048, width=1920, height=1080, aligned_height=1088, frame_stride=2048 * 1088 * 3 // 2)
    ptrs, pitch, fstride = s.plane_args()
    assert s.fmt == "nv12" and ptrs == [0x1000, 0x1000 + 2048 * 1088] and pitch == [2048, 2048] and fstride == [s.frame_stride] * 2
    assert DecoderSurface(0, 4096, 1920, 1080, bit_depth=10).fmt == "p010le"
    assert DecoderSurface(0, 1024, 720, 479).chroma_offset == 1024 * 480


@pytest.mark.gpu
@pytest.mark.parametrize("depth,w,h,aligned_h,pitch,dw,dh", [(8, 720, 480, 480, 768, 1620, 1080), (10, 1920, 1080, 1088, 4096, 1280, 720),
                                                               (8, 1280, 718, 720, 1280, 1920, 1078)])
def test_decoder_surfaces_are_scaled_in_place(depth, w, h, aligned_h, pitch, dw, dh):
    """A batch of NVDEC-layout surfaces (luma, then interleaved chroma `pitch x aligned_height` further down, P016 for
    10-bit streams) goes through tvs_scale_batch without a copy or a repack and matches the oracle."""
    import torch
    from oracle import port
    from thinvids_b200.swscale import Scaler
    from thinvids_b200.worker import DecoderSurface, scale_surfaces
    fmt = "nv12" if depth == 8 else "p010le"
    n = 3
    frames = [F.random_frame(fmt, w, h, 900 + i) for i in range(n)]
    surface_bytes = pitch * aligned_h * 3 // 2
    host = np.full((n, surface_bytes), 0xEE, dtype=np.uint8)
    bps = 1 if depth == 8 else 2
    for i, (y, uv) in enumerate(frames):
        img = host[i].reshape(aligned_h * 3 // 2, pitch)
        img[:h, :w * bps] = y.view(np.uint8).reshape(h, w * bps)
        img[aligned_h:aligned_h + (h + 1) // 2, :uv.shape[1] * bps] = uv.view(np.uint8).reshape(uv.shape[0], uv.shape[1] * bps)
    d_src = torch.from_numpy(host).cuda()
    surf = DecoderSurface(dev_ptr=d_src.data_ptr(), pitch=pitch, width=w, height=h, aligned_height=aligned_h, bit_depth=depth,
                          frame_stride=surface_bytes)
    with Scaler(w, h, fmt, dw, dh, "nv12", "bicubic") as sc:
        out = torch.zeros((n, sc.dst_layout.frame_bytes), dtype=torch.uint8, device="cuda")
        dp, ds, df = sc.dst_layout.pointer_args(out.data_ptr())
        scale_surfaces(sc, surf, list(dp)[:2], list(ds)[:2], list(df)[:2], n, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = sc.dst_layout.unpack(out.cpu().numpy())
    o = port.PortScaler(w, h, F.fmt_code(fmt), dw, dh, F.NV12, F.SWS_BICUBIC)
    for fr, g in zip(frames, got):
        for a, b in zip(g, o.scale(fr)):
            assert np.array_equal(a, b)
