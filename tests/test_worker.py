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
