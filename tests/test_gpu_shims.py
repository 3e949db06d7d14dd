"""GPU: the reference-facing classes end to end (YOLOXDetector / SharedInferenceCoordinator /
InferenceEngine over libtcd) against the CPU oracle port of the reference path
(oracle/reference_detector.py).  Tolerances are north_star's: boxes 0.5 px (network pixels),
scores 1e-2, class ids exact; keep sets must agree wherever the score margin exceeds 2e-2."""
import queue
import time

import numpy as np
import pytest

from util import structured_frame, oracle_model

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def detector():
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector
    d = YOLOXDetector(state_dict=oracle_model(0).state_dict(), wildlife_only=True, max_batch=4)
    assert d.load_model() is True
    yield d
    d.model.engine.close()


def _match(ours, ref, rx, ry):
    """every confident reference detection has a counterpart; nothing confident is invented"""
    def close(a, b):
        bb, cc = a["bbox"], b["bbox"]
        return (a["class_id"] == b["class_id"] and abs(a["confidence"] - b["confidence"]) <= 1e-2 and
                abs(bb["x1"] - cc["x1"]) <= 0.5 * rx and abs(bb["x2"] - cc["x2"]) <= 0.5 * rx and
                abs(bb["y1"] - cc["y1"]) <= 0.5 * ry and abs(bb["y2"] - cc["y2"]) <= 0.5 * ry)
    strong_ref = [r for r in ref if r["confidence"] >= 0.27]
    strong_ours = [o for o in ours if o["confidence"] >= 0.27]
    miss = [r for r in strong_ref if not any(close(r, o) for o in ours)]
    extra = [o for o in strong_ours if not any(close(o, r) for r in ref)]
    return strong_ref, miss, extra


def test_detect_batch_dicts_match_reference_port(detector):
    from oracle.reference_detector import ReferenceYOLOXDetector
    frames = [structured_frame(51), structured_frame(52, 720, 1280), structured_frame(53), structured_frame(54),
              structured_frame(55), structured_frame(56)]                      # 6 > max_batch: two launches
    ref = ReferenceYOLOXDetector(oracle_model(0), 0.25, 0.45, (640, 640), True).detect_batch(frames)
    ours = detector.detect_batch(frames)
    assert len(ours) == len(frames)
    total = 0
    for f, o, r in zip(frames, ours, ref):
        assert all(set(d) == {"class_id", "class_name", "confidence", "bbox"} for d in o)
        assert all(set(d["bbox"]) == {"x1", "y1", "x2", "y2", "area"} and isinstance(d["bbox"]["area"], int) for d in o)
        assert all(d["class_id"] in (0, 14, 15, 16, 21) for d in o)            # wildlife filter
        conf = [d["confidence"] for d in o]
        assert conf == sorted(conf, reverse=True)
        strong, miss, extra = _match(o, r, f.shape[1] / 640, f.shape[0] / 640)
        # NMS decisions that sit within the bf16 tolerance of the IoU / score thresholds may flip
        assert len(miss) <= max(1, 0.03 * len(strong)) and len(extra) <= max(1, 0.03 * len(strong)), (
            "%d missed, %d extra of %d" % (len(miss), len(extra), len(strong)))
        total += len(strong)
    assert total >= 20, "synthetic weights produced too few wildlife detections (%d) to mean anything" % total
    assert detector.detect(frames[0]) == ours[0]
    assert detector.detect_batch([]) == []


def test_preprocess_method_is_the_reference_tensor(detector):
    import torch
    from oracle.reference_detector import preprocess
    f = structured_frame(61, 720, 1280)
    got = detector.preprocess(f)
    assert got.shape == (1, 3, 640, 640) and got.dtype == torch.float32 and got.is_cuda
    assert (got.cpu() - preprocess(f)).abs().max().item() <= 2e-3
    got2 = detector.preprocess(torch.from_numpy(f).cuda())
    assert torch.equal(got, got2)


def test_threshold_attributes_are_honoured(detector):
    f = structured_frame(51)
    base = detector.detect(f)
    detector.conf_threshold = 0.35
    fewer = detector.detect(f)
    detector.conf_threshold = 0.25
    assert len(fewer) < len(base) and all(d["confidence"] >= 0.35 for d in fewer)
    assert detector.detect(f) == base


def test_engine_in_coordinator_mode(detector):
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    from telescope_cam_detection_b200.shared_inference_coordinator import SharedInferenceCoordinator
    frames = [structured_frame(70 + i) for i in range(4)]
    expect = detector.detect_batch(frames)
    with SharedInferenceCoordinator(detector, max_batch_size=4, max_batch_wait_ms=20.0) as co:
        engines, outs = [], []
        for cam in range(2):
            qi, qo = queue.Queue(), queue.Queue()
            e = InferenceEngine(input_queue=qi, output_queue=qo, shared_coordinator=co, camera_id="cam%d" % cam,
                                max_det=300)
            assert e.start()
            engines.append(e)
            outs.append(qo)
        for i, f in enumerate(frames):
            engines[i % 2].input_queue.put({"frame": f, "timestamp": 100.0 + i, "frame_id": i})
        results = {}
        t0 = time.time()
        while len(results) < 4 and time.time() - t0 < 20:
            for qo in outs:
                try:
                    r = qo.get(timeout=0.05)
                    results[r["frame_id"]] = r
                except queue.Empty:
                    pass
        for e in engines:
            e.stop()
        st = co.get_stats()
    assert sorted(results) == [0, 1, 2, 3]
    for i in range(4):
        r = results[i]
        assert set(r) == {"frame_id", "timestamp", "inference_time", "detections", "frame_shape", "camera_id",
                          "camera_name"}
        assert r["camera_id"] == "cam%d" % (i % 2) and r["frame_shape"] == (1080, 1920, 3)
        got = [(d["class_id"], d["confidence"], d["bbox"]["x1"]) for d in r["detections"]]
        want = [(d["class_id"], d["confidence"], d["bbox"]["x1"]) for d in expect[i]][:300]   # max_det
        assert got == want                                                        # same kernels, same numbers
        assert all("width" in d["bbox"] for d in r["detections"])                 # ensure_valid_bbox applied
    assert st["total_frames"] == 4


def test_engine_standalone_with_empty_frame_gate():
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    e = InferenceEngine(state_dict=oracle_model(0).state_dict(), empty_frame_filter_config={"enabled": True})
    assert e.load_model()
    f = structured_frame(51)
    a = e.infer_single(f)
    b = e.infer_single(f)            # identical frame: the gate reports no motion -> []
    assert len(a) > 0 and b == []
    st = e.get_stats()
    assert st["empty_frame_filter"]["skipped_frames"] == 1 and st["memory"]["cuda_available"] is True
    e.detector.model.engine.close()
