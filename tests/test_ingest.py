"""Ingest: RTSPStreamCaptureGPU over the frame ring.

CPU part: the capture loop's host logic against the reference's behaviour (/root/reference/src/
stream_capture_gpu_ffmpeg.py:212-322): queue-entry keys, monotonic frame ids, frames dropped when the queue is full
(their ids are NOT consumed), latest frame, reconnect after the decoder ends, the ffmpeg command line.  The decoder is a
synthetic raw-bgr24 writer (`command=`): there is no ffmpeg binary on this image.
GPU part: frames arrive as device tensors equal to what was written, and a detector fed the device frame returns what it
returns for the host frame."""
import queue
import sys
import time

import numpy as np
import pytest

H, W = 48, 64


def _writer(n, delay=0.0, h=H, w=W):
    prog = ("import sys,time,numpy as np\n"
            "for i in range(%d):\n"
            "  a=np.full((%d,%d,3),i%%251,np.uint8); a[0,0,0]=(7*i)%%256; a[-1,-1,2]=255-(i%%256)\n"
            "  sys.stdout.buffer.write(a.tobytes()); sys.stdout.buffer.flush(); time.sleep(%f)\n" % (n, h, w, delay))
    return [sys.executable, "-c", prog]


def _expect(i, h=H, w=W):
    a = np.full((h, w, 3), i % 251, np.uint8)
    a[0, 0, 0] = (7 * i) % 256
    a[-1, -1, 2] = 255 - (i % 256)
    return a


def _drain(q, n, timeout=20.0):
    out, t0 = [], time.time()
    while len(out) < n and time.time() - t0 < timeout:
        try:
            out.append(q.get(timeout=0.2))
        except queue.Empty:
            pass
    return out


def _cap(q, n, **kw):
    from telescope_cam_detection_b200.stream_capture_gpu_ffmpeg import RTSPStreamCaptureGPU
    kw.setdefault("keep_frames_on_gpu", False)
    kw.setdefault("device", None)
    return RTSPStreamCaptureGPU("rtsp://cam/1", q, W, H, camera_id="cam1", camera_name="Cam One",
                                command=_writer(n, kw.pop("delay", 0.0)), max_failures=kw.pop("max_failures", 2), **kw)


def test_queue_entries_ids_and_content():
    q = queue.Queue(maxsize=64)
    cap = _cap(q, 40)
    assert cap.start() is True
    items = _drain(q, 40)
    cap.stop()
    assert len(items) == 40
    assert sorted(items[0]) == ["camera_id", "camera_name", "frame", "frame_id", "is_gpu_tensor", "timestamp"]
    assert [it["frame_id"] for it in items] == list(range(40))
    assert all(it["camera_id"] == "cam1" and it["camera_name"] == "Cam One" and it["is_gpu_tensor"] is False for it in items)
    for i, it in enumerate(items):      # the ring is deeper than the queue: every queued frame is still intact
        assert np.array_equal(it["frame"], _expect(i)), i
    assert np.array_equal(cap.get_latest_frame_as_numpy(), _expect(39))
    assert not cap.capture_thread.is_alive()


def test_full_queue_drops_frames_without_consuming_ids():
    q = queue.Queue(maxsize=3)
    cap = _cap(q, 30, delay=0.002)
    cap.start()
    # nobody reads: 3 frames queued, 27 dropped.  Wait for the writer's LAST frame rather than a fixed time (the
    # decoder process needs anything from 0.3 s to well over a second to start on a loaded host).
    t0 = time.time()
    while time.time() - t0 < 30.0:
        latest = cap.get_latest_frame_as_numpy()
        if latest is not None and np.array_equal(latest, _expect(29)):
            break
        time.sleep(0.05)
    stats = cap.get_stats()
    items = _drain(q, 3, timeout=2.0)
    cap.stop()
    assert [it["frame_id"] for it in items] == [0, 1, 2]
    for i, it in enumerate(items):
        assert np.array_equal(it["frame"], _expect(i))
    assert stats["queue_size"] == 3 and set(stats) == {"is_connected", "fps", "total_frames", "dropped_frames", "queue_size"}
    assert cap.dropped_frames + cap.frame_count >= 27 or cap.fps > 0     # counters reset once per second, as upstream
    assert np.array_equal(cap.get_latest_frame_as_numpy(), _expect(29))  # the newest frame, although it was dropped


def test_reconnects_when_the_decoder_ends():
    q = queue.Queue(maxsize=64)
    cap = _cap(q, 5, max_failures=2)
    cap.start()
    items = _drain(q, 12, timeout=20.0)  # the writer exits after 5 frames; the loop restarts it (ids keep counting)
    cap.stop()
    assert len(items) >= 10
    assert [it["frame_id"] for it in items] == list(range(len(items)))
    assert np.array_equal(items[5]["frame"], _expect(0)) and np.array_equal(items[9]["frame"], _expect(4))


def test_ffmpeg_command_line_is_the_reference_one():
    from telescope_cam_detection_b200.stream_capture_gpu_ffmpeg import RTSPStreamCaptureGPU
    cap = RTSPStreamCaptureGPU("rtsp://u:p@10.0.0.2/s", queue.Queue(), 1920, 1080, use_tcp=True, device=None,
                               keep_frames_on_gpu=False)
    assert cap._build_ffmpeg_command() == ["ffmpeg", "-hwaccel", "cuda", "-c:v", "h264_cuvid", "-rtsp_transport", "tcp",
                                           "-i", "rtsp://u:p@10.0.0.2/s", "-f", "rawvideo", "-pix_fmt", "bgr24",
                                           "-s", "1920x1080", "pipe:1"]
    cap.use_tcp = False
    assert "-rtsp_transport" not in cap._build_ffmpeg_command()
    assert cap.connect() is False        # no ffmpeg binary here: "not connected", no exception (reference :157-159)


@pytest.mark.gpu
def test_device_frames_equal_the_stream_and_feed_the_detector():
    import torch
    from telescope_cam_detection_b200.stream_capture_gpu_ffmpeg import RTSPStreamCaptureGPU
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector

    h, w, n = 360, 640, 12
    q = queue.Queue(maxsize=16)
    cap = RTSPStreamCaptureGPU("rtsp://cam/2", q, w, h, camera_id="cam2", command=_writer(n, 0.0, h, w), max_failures=2,
                               keep_frames_on_gpu=True, device="cuda:0")
    cap.start()
    items = _drain(q, n)
    cap.stop()
    assert len(items) == n
    for i, it in enumerate(items):
        f = it["frame"]
        assert it["is_gpu_tensor"] is True and isinstance(f, torch.Tensor) and f.is_cuda and f.dtype == torch.uint8
        assert tuple(f.shape) == (h, w, 3)
        assert np.array_equal(f.cpu().numpy(), _expect(i, h, w)), i
    assert np.array_equal(cap.get_latest_frame_as_numpy(), _expect(n - 1, h, w))

    # a device frame from the ring and the same frame from host memory give the same detections
    from util import structured_frame, oracle_model
    frame = structured_frame(5, h, w)
    ring = cap.ring
    ring.host(0)[...] = frame
    ring.publish(0)
    ring.wait(0)
    det = YOLOXDetector(state_dict=oracle_model(0).state_dict(), wildlife_only=False, max_batch=2,
                        max_src=(h, w))
    assert det.load_model()
    a = det.detect(ring.device_tensor(0))
    b = det.detect(frame)
    assert a == b and len(a) > 0
