"""CPU: host-side logic of the reference-facing classes against golden vectors produced by the
REFERENCE's own code (tests/golden/format_ref.json), coordinator semantics, and the C-ABI library
(loads, exports every symbol include/tcd.h declares; no compute calls without a GPU)."""
import json
import os
import re
import sys
import threading
import time

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
sys.path.insert(0, GOLD)

SHIPPED_CONF = {"person": 0.75, "bird": 0.70, "bear": 0.95, "dog": 0.70, "cat": 0.70}
SHIPPED_SIZE = {"bird": {"max": 8000}, "bear": {"min": 5000}, "dog": {"min": 1000}, "cat": {"min": 800},
                "rabbit": {"min": 400, "max": 3000}}


def _cases():
    with open(os.path.join(GOLD, "format_ref.json")) as f:
        return json.load(f)["cases"]


def _rows(seed):
    from make_golden import seeded_rows
    return seeded_rows(seed)


@pytest.mark.parametrize("case", _cases(), ids=lambda c: "seed%d" % c["seed"])
def test_format_and_per_class_filter_equal_the_reference(case):
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector
    from oracle.reference_detector import format_rows

    oh, ow = case["orig_hw"]
    det = YOLOXDetector(input_size=(640, 640), wildlife_only=case["wildlife_only"])
    dets = det._format_model_output_to_detections(_rows(case["seed"]), oh, ow)
    assert json.loads(json.dumps(dets)) == case["formatted"]           # A4, exact (floats included)
    assert format_rows(_rows(case["seed"]), oh, ow, (640, 640), case["wildlife_only"]) == case["formatted"]  # oracle
    eng = InferenceEngine(conf_threshold=0.25, min_box_area=300, max_det=20,
                          class_confidence_overrides=dict(SHIPPED_CONF),
                          class_size_constraints=json.loads(json.dumps(SHIPPED_SIZE)),
                          wildlife_only=case["wildlife_only"])
    kept = eng._post_process_detections(dets, np.zeros((oh, ow, 3), np.uint8))
    assert json.loads(json.dumps(kept)) == case["post_processed"]      # A6, exact, order preserved
    assert len(kept) <= 20


def test_bbox_utils():
    from telescope_cam_detection_b200.bbox_utils import ensure_valid_bbox, is_valid_bbox, validate_bbox_coords
    b = ensure_valid_bbox({"x1": 30.5, "y1": 2.0, "x2": 10.0, "y2": 2.25})
    assert b == {"x1": 10.0, "y1": 2.0, "x2": 30.5, "y2": 3.0, "width": 20.5, "height": 1.0, "area": 20.5}
    assert validate_bbox_coords(5, 5, 5, 9) == (5, 5, 6, 9)
    assert is_valid_bbox({"x1": 0, "y1": 0, "x2": 2, "y2": 2}) and not is_valid_bbox({"x1": -1, "y1": 0, "x2": 2, "y2": 2})
    assert not is_valid_bbox({"x1": 3, "y1": 0, "x2": 2, "y2": 2}) and not is_valid_bbox({"x1": 3})


class _FakeDetector:
    def __init__(self, fail_on=None, delay=0.0):
        self.batches, self.fail_on, self.delay = [], fail_on, delay

    def detect_batch(self, frames):
        self.batches.append(list(frames))
        time.sleep(self.delay)
        if self.fail_on is not None and self.fail_on in frames:
            raise RuntimeError("boom")
        return [[{"frame": f}] for f in frames]


class _FakeSplitDetector(_FakeDetector):
    """also offers the split-phase pair the pipelined loop uses"""

    def __init__(self, **kw):
        super().__init__(**kw)
        self.slots = {}

    def submit_batch(self, slot, frames):
        assert slot not in self.slots, "slot reused before finish"
        self.slots[slot] = list(frames)

    def finish_batch(self, slot):
        return self.detect_batch(self.slots.pop(slot))


@pytest.mark.parametrize("cls", [_FakeDetector, _FakeSplitDetector])
def test_coordinator_batches_in_order_and_reports_errors_as_empty(cls):
    from telescope_cam_detection_b200.shared_inference_coordinator import SharedInferenceCoordinator
    det = cls(fail_on=13, delay=0.002)
    got, lock = [], threading.Lock()
    with SharedInferenceCoordinator(det, max_batch_size=4, max_batch_wait_ms=5.0) as co:
        for i in range(20):
            co.infer_async(i, lambda d, i=i: (lock.acquire(), got.append((i, d)), lock.release()), camera_id="c%d" % i)
        t0 = time.time()
        while len(got) < 20 and time.time() - t0 < 5:
            time.sleep(0.01)
        stats = co.get_stats()
    assert [i for i, _ in got] == list(range(20))                      # callbacks fire in request order
    assert all(len(b) <= 4 for b in det.batches) and sum(len(b) for b in det.batches) == 20
    bad = next(b for b in det.batches if 13 in b)
    for i, d in got:
        assert d == ([] if i in bad else [{"frame": i}])                # a failing batch -> callback([]) for all of it
    assert stats["enabled"] and stats["total_frames"] == 20 - len(bad)
    assert set(stats) == {"enabled", "total_batches", "total_frames", "avg_batch_size", "avg_batch_time_ms",
                          "avg_wait_time_ms", "throughput_fps", "queue_depth"}


def test_coordinator_drops_oldest_when_full_and_refuses_when_stopped():
    from telescope_cam_detection_b200.shared_inference_coordinator import SharedInferenceCoordinator
    co = SharedInferenceCoordinator(_FakeDetector(), max_batch_size=2, max_queue_depth=3)
    with pytest.raises(RuntimeError):
        co.infer_async(0, lambda d: None)
    co.running = True  # queue without a consumer thread
    got = {}
    for i in range(5):
        co.infer_async(i, lambda d, i=i: got.setdefault(i, d))
    assert got == {0: [], 1: []} and co.dropped_frames == 2 and [r.frame for r in co.pending_queue] == [2, 3, 4]
    assert co.get_stats() == {"enabled": False, "total_batches": 0, "total_frames": 0}


def test_engine_surface_without_gpu():
    """constructor keywords, stats keys, sparse keyframes and hot reload need no device"""
    from telescope_cam_detection_b200.inference_engine_yolox import InferenceEngine
    eng = InferenceEngine(sparse_detection_config={"enabled": True, "keyframe_interval": 3}, some_legacy_key=1)
    assert eng.infer_single(np.zeros((4, 4, 3), np.uint8)) == []        # not loaded -> [] like the reference
    with pytest.raises(ValueError):
        InferenceEngine(detector_type="dino")

    class Det:
        conf_threshold, nms_threshold, calls = 0.25, 0.45, 0

        def detect(self, frame):
            Det.calls += 1
            return [{"class_id": 14, "class_name": "bird", "confidence": 0.9,
                     "bbox": {"x1": 0.0, "y1": 0.0, "x2": 40.0, "y2": 30.0, "area": 1200}}]

    eng.detector, eng.is_loaded = Det(), True
    outs = [eng.infer_single(np.zeros((8, 8, 3), np.uint8)) for _ in range(6)]
    assert Det.calls == 2 and [len(o) for o in outs] == [0, 0, 1, 1, 1, 1]   # frames 3 and 6 are keyframes
    assert outs[3][0] is outs[2][0]                                     # shallow copy quirk kept (reference :586)
    eng.update_settings(conf_threshold=0.95, nms_threshold=0.5, max_det=7)
    assert eng.detector.conf_threshold == 0.95 and eng.detector.nms_threshold == 0.5 and eng.max_det == 7
    st = eng.get_stats()
    for k in ("device", "fps", "avg_inference_time", "avg_inference_time_ms", "total_inferences", "dropped_results",
              "coordinator_mode", "degradation_active", "drop_rate", "memory", "sparse_detection"):
        assert k in st
    assert st["sparse_detection"]["keyframes_processed"] == 2


def test_library_exports_every_declared_symbol():
    from telescope_cam_detection_b200 import _lib
    lib = _lib.load()
    with open(os.path.join(ROOT, "include", "tcd.h")) as f:
        declared = set(re.findall(r"\b(tcd_[a-z0-9_]+)\s*\(", f.read()))
    assert len(declared) >= 35
    for name in sorted(declared):
        assert hasattr(lib, name), "libtcd.so does not export %s" % name
        assert name in _lib.SIGNATURES, "no ctypes signature for %s" % name
    assert lib.tcd_version() >= 100 and lib.tcd_launch_count() == 0


def test_no_gpu_means_failure_not_fallback():
    """product path fails loudly without a device: status != 0 and a message, never a CPU result"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("box has a GPU")
    from telescope_cam_detection_b200 import _lib
    from telescope_cam_detection_b200.engine import Mog2Engine
    with pytest.raises(_lib.TcdError):
        Mog2Engine(48, 64)
    from telescope_cam_detection_b200.yolox_detector import YOLOXDetector
    d = YOLOXDetector(state_dict={})
    assert d.load_model(max_retries=1) is False and d.detect(np.zeros((8, 8, 3), np.uint8)) == []


def test_product_never_imports_or_links_the_oracle():
    """oracle/ is test infrastructure: no module of the package imports it, no native source includes it, and the
    package imports with oracle/ off sys.path.  (tests/, bench.py's cpu_baseline / reference arm and smoke() may.)"""
    import ast
    import glob
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    pkg = os.path.join(root, "telescope_cam_detection_b200")
    for path in glob.glob(os.path.join(pkg, "**", "*.py"), recursive=True):
        tree = ast.parse(open(path).read(), path)
        for node in ast.walk(tree):
            names = []
            if isinstance(node, ast.Import):
                names = [a.name for a in node.names]
            elif isinstance(node, ast.ImportFrom):
                names = [node.module or ""]
            elif isinstance(node, ast.Call) and getattr(node.func, "attr", getattr(node.func, "id", "")) in ("import_module", "__import__"):
                names = [a.value for a in node.args[:1] if isinstance(a, ast.Constant) and isinstance(a.value, str)]
            for n in names:
                assert n.split(".")[0] != "oracle", (path, n)
    for path in glob.glob(os.path.join(pkg, "csrc", "*")):
        for line in open(path, errors="replace"):
            if line.lstrip().startswith("#include"):
                assert "oracle" not in line, (path, line)
    # a fresh interpreter started outside the repo root: `import oracle` is impossible there, the package still imports
    code = ("import sys; sys.path = [p for p in sys.path if p not in ('', %r)]; sys.path.insert(0, %r); "
            "import importlib.util as u; sys.modules.pop('oracle', None); "
            "import telescope_cam_detection_b200 as t; "
            "import telescope_cam_detection_b200.species_classifier, telescope_cam_detection_b200.eva02_engine, "
            "telescope_cam_detection_b200.shared_inference_coordinator; "
            "assert not any(m == 'oracle' or m.startswith('oracle.') for m in sys.modules), 'oracle imported'" % (root, root))
    # the repo root must be importable for the package itself; oracle/ sits beside it, so check sys.modules instead
    r = subprocess.run([sys.executable, "-c", code], cwd="/", capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
