"""GPU: the device-side per-camera class filter (K7) against the golden vectors the REFERENCE's own
InferenceEngine._post_process_detections produced (tests/golden/format_ref.json), and end to end through
YOLOXDetector / InferenceEngine / SharedInferenceCoordinator: the rows that cross PCIe must be exactly the ones
the reference's Python loop keeps, in the same order."""
import json
import os
import queue
import sys
import time

import numpy as np
import pytest

from util import structured_frame, oracle_model

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
sys.path.insert(0, GOLD)
SHIPPED_CONF = {"person": 0.75, "bird": 0.70, "bear": 0.95, "dog": 0.70, "cat": 0.70}
SHIPPED_SIZE = {"bird": {"max": 8000}, "bear": {"min": 5000}, "dog": {"min": 1000}, "cat": {"min": 800},
                "rabbit": {"min": 400, "max": 3000}}


def _cases():
    with open(os.path.join(GOLD, "format_ref.json")) as f:
        return json.load(f)["cases"]


@pytest.fixture(scope="module")
def detector():
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector
    d = YOLOXDetector(state_dict=oracle_model(0).state_dict(), wildlife_only=True, max_batch=4)
    assert d.load_model() is True
    yield d
    d.model.engine.close()


@pytest.mark.parametrize("case", _cases(), ids=lambda c: "seed%d" % c["seed"])
def test_device_filter_keeps_exactly_the_reference_rows(detector, case):
    from make_golden import seeded_rows
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    rows = seeded_rows(case["seed"])
    oh, ow = case["orig_hw"]
    detector.wildlife_only = case["wildlife_only"]
    detector._sync_settings()
    tid = detector.register_class_filter(0.25, dict(SHIPPED_CONF), 300, json.loads(json.dumps(SHIPPED_SIZE)), 20)
    try:
        keep = detector.model.engine.filter_rows(tid, rows, oh, ow)
        dets = detector._format_model_output_to_detections(rows[keep], oh, ow)
        eng = InferenceEngine(conf_threshold=0.25, min_box_area=300, max_det=20,
                              class_confidence_overrides=dict(SHIPPED_CONF),
                              class_size_constraints=json.loads(json.dumps(SHIPPED_SIZE)),
                              wildlife_only=case["wildlife_only"])
        kept = eng._post_process_detections(dets, np.zeros((oh, ow, 3), np.uint8))
        assert json.loads(json.dumps(kept)) == case["post_processed"]        # equal to the reference's own output
        assert int(keep.sum()) == len(case["post_processed"])                # and the host loop had nothing left to drop
        # own table (wildlife mask only) == the reference's formatting step
        own = detector.model.engine.filter_rows(detector._OWN_TABLE, rows, oh, ow)
        assert json.loads(json.dumps(detector._format_model_output_to_detections(rows[own], oh, ow))) == case["formatted"]
    finally:
        detector.release_class_filter(tid)
        detector.wildlife_only = True


def test_detect_batch_with_filter_ids_equals_host_filter(detector):
    """Two cameras with different tables in ONE batch; each frame's device-filtered dicts equal the host loop run
    over the unfiltered dicts of the same frame."""
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    frames = [structured_frame(80 + i) for i in range(4)]
    cams = [dict(conf_threshold=0.3, min_box_area=2000, max_det=7, class_confidence_overrides={"person": 0.4},
                 class_size_constraints={"dog": {"min": 9000}, "cat": {"max": 60000}}),
            dict(conf_threshold=0.25, min_box_area=0, max_det=300, class_confidence_overrides={},
                 class_size_constraints={})]
    engines = [InferenceEngine(**c) for c in cams]
    for e in engines:
        e.detector = detector            # what load_model() would set in standalone mode
    tids = [e._device_filter() for e in engines]
    assert None not in tids and tids[0] != tids[1]
    raw = detector.detect_batch(frames)
    got = detector.detect_batch(frames, [tids[i % 2] for i in range(4)])
    n_dropped = 0
    for i, f in enumerate(frames):
        want = engines[i % 2]._post_process_detections([dict(d, bbox=dict(d["bbox"])) for d in raw[i]], f)
        mine = engines[i % 2]._post_process_detections(got[i], f)
        assert len(mine) == len(got[i])                                   # nothing left for the host loop
        assert mine == want
        n_dropped += len(raw[i]) - len(want)
    assert n_dropped > 0, "the tables dropped nothing: test frames too quiet"
    # hot reload reaches the device table
    engines[0].update_settings(max_det=2)
    assert all(len(d) <= 2 for d in detector.detect_batch(frames[:1], [tids[0]]))
    for e in engines:
        detector.release_class_filter(e._filter_id)


def test_oversize_frame_through_engine_and_coordinator():
    """a 2560x1440 camera (the reference's shipped config has one) must not blank its batch."""
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    from telescope_cam_detection_b200.shared_inference_coordinator import SharedInferenceCoordinator
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector
    big, small = structured_frame(90, 1440, 2560), structured_frame(91)
    e = InferenceEngine(state_dict=oracle_model(0).state_dict())
    assert e.load_model()
    a = e.infer_single(big)
    assert len(a) > 0 and max(d["bbox"]["x2"] for d in a) > 1920         # boxes in the big frame's own pixels
    e.detector.model.engine.close()
    det = YOLOXDetector(state_dict=oracle_model(0).state_dict(), max_batch=4)
    assert det.load_model()
    want = det.detect_batch([small])[0]
    got = {}
    with SharedInferenceCoordinator(det, max_batch_size=4, max_batch_wait_ms=50.0) as co:
        co.infer_async(big, lambda d: got.__setitem__("big", d), "cam_big")
        co.infer_async(small, lambda d: got.__setitem__("small", d), "cam_small")
        t0 = time.time()
        while len(got) < 2 and time.time() - t0 < 20:
            time.sleep(0.01)
    assert len(got["big"]) > 0 and got["small"] == want
    det.model.engine.close()


def test_coordinator_uses_every_slot_and_keeps_order(detector):
    from telescope_cam_detection_b200.shared_inference_coordinator import SharedInferenceCoordinator
    frames = [structured_frame(100 + i % 5) for i in range(40)]
    want = detector.detect_batch(frames[:5])
    got = []
    with SharedInferenceCoordinator(detector, max_batch_size=2, max_batch_wait_ms=1.0, max_queue_depth=100) as co:
        for i, f in enumerate(frames):
            co.infer_async(f, lambda d, i=i: got.append((i, d)), "cam%d" % (i % 3))
        t0 = time.time()
        while len(got) < len(frames) and time.time() - t0 < 30:
            time.sleep(0.01)
        st = co.get_stats()
    assert [i for i, _ in got] == list(range(40))                          # callbacks in request order
    for i, d in got:
        assert d == want[i % 5]
    assert st["total_frames"] == 40
