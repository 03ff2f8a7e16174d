"""bwdif (thinvids_b200/csrc/kernels_bwdif.cu, the reference's DVD-path de-interlacer: worker/tasks.py:62-63,446-447).

PARITY UNPINNED: there is no libavfilter in the image, so the checker is oracle/bwdif_port.py - a numpy restatement of the
published algorithm - plus properties any implementation of it must have.  These tests prove the kernel and the
restatement agree and behave; they do not prove agreement with ffmpeg."""
import numpy as np
import pytest

from thinvids_b200 import frames as F
from thinvids_b200.swscale import PackedLayout


def _interlaced_sequence(fmt, w, h, n, seed):
    """A moving pattern plus noise, field-interleaved: different content in the two fields of a frame."""
    rng = np.random.default_rng(seed)
    top = 255 if F.bit_depth(fmt) == 8 else 1023
    frames = []
    for i in range(n):
        fr = []
        for rows, cols in F.plane_shapes(fmt, w, h):
            y, x = np.mgrid[0:rows, 0:cols]
            shift = 3 * (2 * i + (y & 1))                           # each field is a later instant
            base = ((x + shift) * 7 + (y // 2) * 5) % (top + 1)
            noise = rng.integers(-6, 7, (rows, cols))
            fr.append(np.clip(base + noise, 0, top).astype(F.sample_dtype(fmt)))
        frames.append(fr)
    return frames


def test_oracle_restatement_invariants():
    from oracle import bwdif_port
    for fmt, depth in (("yuv420p", 8), ("yuv420p10le", 10)):
        frames = _interlaced_sequence(fmt, 48, 38, 5, 1)
        for tff in (1, 0):
            out = bwdif_port.bwdif_stream(frames, tff=tff, depth=depth)
            assert len(out) == 5
            keep = 0 if tff else 1
            for fi, fo in zip(frames, out):
                for a, b in zip(fi, fo):
                    assert np.array_equal(a[keep::2], b[keep::2])          # the kept field is copied
                    assert not np.array_equal(a[1 - keep::2], b[1 - keep::2])
        # a static scene: temporal differences are zero, every interpolated sample is the temporal prediction = itself
        still = [frames[0]] * 4
        out = bwdif_port.bwdif_stream(still, tff=1, depth=depth)
        for k in (1, 2):                                                    # frames with both neighbours (not intra)
            for a, b in zip(still[k], out[k]):
                assert np.array_equal(a, b)
        # batch-wise processing with one frame of context each side gives the same frames
        whole = bwdif_port.bwdif_stream(frames, tff=1, depth=depth)
        head = bwdif_port.bwdif_stream(frames[:3], tff=1, depth=depth, stream_end=False)
        tail = bwdif_port.bwdif_stream(frames[1:], tff=1, depth=depth, stream_start=False)
        for got, want in zip(head + tail, whole):
            for a, b in zip(got, want):
                assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt,w,h", [("yuv420p", 720, 480), ("yuv420p", 97, 53), ("yuv420p10le", 128, 72), ("yuv420p", 16, 7),
                                     ("yuv420p10le", 33, 12)])
def test_gpu_bwdif_equals_the_restatement(fmt, w, h):
    from oracle import bwdif_port
    from thinvids_b200.deinterlace import BwdifStream, bwdif_packed_host
    depth = F.bit_depth(fmt)
    frames = _interlaced_sequence(fmt, w, h, 6, 7)
    lay = PackedLayout(F.fmt_code(fmt), w, h)
    packed = lay.pack(frames)
    for tff in (True, False):
        want = bwdif_port.bwdif_stream(frames, tff=int(tff), depth=depth)
        got = lay.unpack(bwdif_packed_host(packed, w, h, fmt, tff=tff))
        assert len(got) == 6
        for k, (g, wnt) in enumerate(zip(got, want)):
            for a, b in zip(g, wnt):
                assert np.array_equal(a, b), (fmt, w, h, tff, k, int((a != b).sum()))
    # the same stream in batches of 1, 2 and 3 frames
    want = bwdif_port.bwdif_stream(frames, tff=1, depth=depth)
    st = BwdifStream(w, h, fmt, True)
    outs = [st.push(packed[0:1]), st.push(packed[1:3]), st.push(packed[3:6]), st.flush()]
    got = lay.unpack(np.concatenate([o for o in outs if o.size]))
    assert len(got) == 6
    for g, wnt in zip(got, want):
        for a, b in zip(g, wnt):
            assert np.array_equal(a, b)


@pytest.mark.gpu
def test_dvd_job_runs_end_to_end_with_gpu_deinterlace():
    """The reference's DVD job (480-line mpeg2 source under dvd/: SCALE_FILTER_480 = bwdif,scale=-2:480,format=nv12) through
    encode_part with nothing de-interlaced upstream: bwdif on the GPU, then the same-size NV12 repack."""
    import io
    from oracle import bwdif_port, port
    from thinvids_b200.worker import encode_part
    w, h, n = 720, 480, 5
    frames = _interlaced_sequence("yuv420p", w, h, n, 3)
    lay = PackedLayout(F.YUV420P, w, h)
    job = {"target_height": "1080", "software_encode": "0", "filename": "/watch/dvd/movie.mkv", "source_codec": "mpeg2video",
           "source_resolution": "720x480"}
    out = io.BytesIO()
    res = encode_part(job, 0, io.BytesIO(lay.pack(frames).tobytes()), out, src_w=w, src_h=h, gpu_deinterlace=True, tff=True, batch_frames=2)
    assert res["status"] == "COMPLETED" and res["frames"] == n and (res["dst_w"], res["dst_h"], res["dst_fmt"]) == (720, 480, "nv12")
    assert "bwdif" in res["filtergraph"]
    deint = bwdif_port.bwdif_stream(frames, tff=1, depth=8)
    o = port.PortScaler(w, h, F.YUV420P, w, h, F.NV12, F.SWS_BICUBIC)
    got = PackedLayout(F.NV12, w, h).unpack(np.frombuffer(out.getvalue(), dtype=np.uint8).reshape(n, -1))
    for fr, g in zip(deint, got):
        for a, b in zip(g, o.scale(fr)):
            assert np.array_equal(a, b)
    # without the switch the graph is refused, as before
    res = encode_part(job, 0, io.BytesIO(b""), io.BytesIO(), src_w=w, src_h=h)
    assert res["status"] == "FAILED" and "bwdif" in res["reason"]
