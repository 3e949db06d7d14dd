"""JPEG encode on the result path.

CPU: oracle/jpeg.py returns the same BYTES as cv2.imencode (the reference's encoder: web_server.py:923,
snapshot_saver.py:160) for sizes with and without whole MCUs and qualities 20-100.
GPU: the CUDA encoder returns the same bytes as cv2.imencode for the same cases, at 1080p, from device frames, in
batches of mixed sizes; the decoded image equals cv2's decode of cv2's file (implied by byte equality, checked once)."""
import numpy as np
import pytest

from util import structured_frame

CASES = [(40, 56, 95), (64, 64, 95), (33, 47, 95), (120, 200, 75), (97, 131, 50), (16, 16, 100), (100, 100, 20),
         (8, 8, 90), (1, 1, 95), (17, 250, 85), (250, 17, 60)]


def _image(h, w, q):
    if q == 50:
        return np.random.default_rng(h * 1000 + w).integers(0, 256, (h, w, 3), dtype=np.uint8)   # noise: long codes
    return structured_frame(h * 7 + w, h, w)


def _cv2(img, q):
    import cv2
    ok, buf = cv2.imencode(".jpg", img, [cv2.IMWRITE_JPEG_QUALITY, q])
    assert ok
    return buf.tobytes()


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%dx%d_q%d" % c)
def test_oracle_bytes_equal_cv2(case):
    from oracle import jpeg as J
    h, w, q = case
    img = _image(h, w, q)
    assert J.encode(img, q) == _cv2(img, q)


@pytest.mark.gpu
def test_cuda_encoder_bytes_equal_cv2():
    import cv2
    import torch
    from telescope_cam_detection_b200 import jpeg_encode as JE
    enc = JE.JpegEncoder((1080, 1920), max_batch=8)
    for h, w, q in CASES:
        img = _image(h, w, q)
        got = enc.encode_batch([img], q)[0].tobytes()
        ref = _cv2(img, q)
        assert got == ref, "%dx%d q%d: %d vs %d bytes, first difference at %d" % (
            h, w, q, len(got), len(ref), next((i for i in range(min(len(got), len(ref))) if got[i] != ref[i]), -1))
    # full frames: structured, noise (every AC coefficient coded), flat (all-EOB blocks), from host and from the device
    frames = [structured_frame(3), np.random.default_rng(4).integers(0, 256, (1080, 1920, 3), dtype=np.uint8),
              np.full((1080, 1920, 3), 200, np.uint8), structured_frame(5, 720, 1280)]
    for q in (95, 90, 60):
        refs = [_cv2(f, q) for f in frames]
        assert [b.tobytes() for b in enc.encode_batch(frames, q)] == refs                      # mixed sizes, one call
        dev = [torch.from_numpy(f).cuda() for f in frames]
        assert [b.tobytes() for b in enc.encode_batch(dev, q)] == refs
    # a view into a larger device frame (row stride of the parent)
    big = torch.from_numpy(structured_frame(6)).cuda()
    view = big[100:613, 200:905]
    assert enc.encode_batch([view], 95)[0].tobytes() == _cv2(np.ascontiguousarray(view.cpu().numpy()), 95)
    # cv2's call signature
    ok, buf = JE.imencode(".jpg", frames[0], [cv2.IMWRITE_JPEG_QUALITY, 90])
    assert ok and buf.dtype == np.uint8 and buf.tobytes() == _cv2(frames[0], 90)
    assert JE.imencode(".jpg", frames[3])[1].tobytes() == cv2.imencode(".jpg", frames[3])[1].tobytes()   # default quality 95
    dec = cv2.imdecode(buf, cv2.IMREAD_COLOR)
    assert dec.shape == frames[0].shape
    with pytest.raises(NotImplementedError):
        JE.imencode(".png", frames[0])
    # more frames than the encoder's batch: split, order kept
    many = [structured_frame(40 + i, 64, 96) for i in range(11)]
    assert [b.tobytes() for b in enc.encode_batch(many, 80)] == [_cv2(f, 80) for f in many]


def test_detection_message_and_mjpeg_part():
    from telescope_cam_detection_b200.jpeg_encode import mjpeg_part, prepare_detection_message
    r = {"camera_id": "cam3", "frame_id": 7, "timestamp": 1.5, "total_latency_ms": 12.0, "detections": [{"class_id": 14}],
         "detection_counts": {"bird": 1}, "total_detections": 1, "inference_time_ms": 3.0, "extra": "dropped"}
    m = prepare_detection_message(r)
    assert m == {"type": "detections", "camera_id": "cam3", "camera_name": "Default Camera", "frame_id": 7, "timestamp": 1.5,
                 "processing_timestamp": None, "latency_ms": 12.0, "inference_time_ms": 3.0, "detections": [{"class_id": 14}],
                 "detection_counts": {"bird": 1}, "total_detections": 1}
    assert prepare_detection_message({})["camera_id"] == "default" and prepare_detection_message({})["detections"] == []
    assert mjpeg_part(b"\xff\xd8x\xff\xd9") == b"--frame\r\nContent-Type: image/jpeg\r\n\r\n\xff\xd8x\xff\xd9\r\n"
