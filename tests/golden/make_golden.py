"""Generates tests/golden/*.npz|json by running the REFERENCE's own classes (imported from
/root/reference, read-only) in the build container.  /root/reference does not exist on the GPU
box, so the outputs are committed; rerun with `python tests/golden/make_golden.py`.

  motion_ref.npz   MotionFilter.filter_detections on tests/util.motion_sequence (per-detection
                   verdict + ratio, incl. the one-apply-per-detection quirk)
  empty_ref.npz    EmptyFrameFilter.has_motion verdicts on the same frames
  stage2_ref.json  TwoStageDetectionPipeline.process_detections over recording classifiers: the crop
                   windows the reference slices (read back from a coordinate-encoding frame), the
                   top_k it asks for and the species fields it writes, incl. the time-of-day re-rank
  format_ref.json  YOLOXDetector._format_model_output_to_detections and
                   InferenceEngine._post_process_detections on seeded postprocess rows, with the
                   shipped per-class tables (config/config.yaml:238-264).  These two import
                   src/yolox_detector.py, which needs a YOLOX checkout: a stub package exporting
                   the oracle restatement is put on sys.path.
  enhance_ref.npz  ImageEnhancer(method="clahe").enhance on seeded crops (tests/util.enhance_crops): colour crops of
                   several sizes (tile-divisible and not), a grey crop (final outputs only: the stages are checked
                   against cv2 itself, which is on both boxes)
Versions of the libraries that did the arithmetic are stored in every file.
"""
import json
import os
import shutil
import sys
import tempfile
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.dont_write_bytecode = True
REF = "/root/reference"


def versions():
    import cv2
    import torch
    import torchvision
    return {"cv2": cv2.__version__, "numpy": np.__version__, "torch": torch.__version__,
            "torchvision": torchvision.__version__, "reference": "filthyrake/telescope_cam_detection@249c75e"}


def motion_and_empty():
    sys.path.insert(0, REF)
    from src.motion_filter import MotionFilter
    from src.empty_frame_filter import EmptyFrameFilter
    from util import motion_sequence
    seq = motion_sequence()
    mf = MotionFilter()
    verdict, ratio, nper = [], [], []
    for frame, boxes in seq:
        dets = [{"bbox": dict(b), "idx": i} for i, b in enumerate(boxes)]
        # per detection: has_motion_in_bbox applies MOG2 once, exactly what filter_detections does
        for d in dets:
            ok, r = mf.has_motion_in_bbox(frame, d["bbox"])
            verdict.append(ok)
            ratio.append(r)
        nper.append(len(dets))
    np.savez_compressed(os.path.join(HERE, "motion_ref.npz"), verdict=np.array(verdict), ratio=np.array(ratio),
                        n_per_frame=np.array(nper), versions=json.dumps(versions()))
    ef = EmptyFrameFilter()
    ev = [ef.has_motion(f) for f, _ in seq]
    ef2 = EmptyFrameFilter(min_motion_area=5000, threshold=10, blur_size=9)
    ev2 = [ef2.has_motion(f) for f, _ in seq]
    np.savez_compressed(os.path.join(HERE, "empty_ref.npz"), default=np.array(ev), tuned=np.array(ev2),
                        versions=json.dumps(versions()))
    print("motion: %d verdicts (%d true); empty: %d/%d true" % (len(verdict), sum(verdict), sum(ev), len(ev)))


SHIPPED_CONF = {"person": 0.75, "bird": 0.70, "bear": 0.95, "dog": 0.70, "cat": 0.70}      # config.yaml:238-243
SHIPPED_SIZE = {"bird": {"max": 8000}, "bear": {"min": 5000}, "dog": {"min": 1000}, "cat": {"min": 800},
                "rabbit": {"min": 400, "max": 3000}}                                            # config.yaml:253-264


def seeded_rows(seed, n=160):
    """postprocess-shaped rows [x1,y1,x2,y2,obj,cls,class_id] (network pixels) covering wildlife and
    other classes, tiny/huge/inverted boxes and confidences around every shipped threshold."""
    rng = np.random.default_rng(seed)
    cx, cy = rng.uniform(0, 640, n), rng.uniform(0, 640, n)
    w, h = np.exp(rng.uniform(np.log(2), np.log(400), n)), np.exp(rng.uniform(np.log(2), np.log(400), n))
    rows = np.stack([cx - w / 2, cy - h / 2, cx + w / 2, cy + h / 2, rng.uniform(0.3, 1, n), rng.uniform(0.3, 1, n),
                     rng.choice([0, 14, 15, 16, 21, 2, 56, 17, 79], n).astype(np.float64)], 1).astype(np.float32)
    rows[::17, [0, 2]] = rows[::17, [2, 0]]  # inverted x
    order = np.argsort(-(rows[:, 4] * rows[:, 5]), kind="stable")
    return rows[order]


def stub_reference_tree():
    """Scratch copy of /root/reference/src beside a stub YOLOX package (existence check at
    src/yolox_detector.py:17-31) so the reference's own detector / engine modules import."""
    tmp = tempfile.mkdtemp(prefix="tcd_ref_")
    shutil.copytree(os.path.join(REF, "src"), os.path.join(tmp, "src"))
    for sub in ("YOLOX/yolox", "YOLOX/yolox/exp", "YOLOX/yolox/utils"):
        os.makedirs(os.path.join(tmp, sub))
    open(os.path.join(tmp, "YOLOX/yolox/__init__.py"), "w").close()
    with open(os.path.join(tmp, "YOLOX/yolox/exp/__init__.py"), "w") as f:
        f.write("from oracle.yolox_s import get_exp\n")
    with open(os.path.join(tmp, "YOLOX/yolox/utils/__init__.py"), "w") as f:
        f.write("from oracle.postprocess import postprocess\n")
    sys.path.insert(0, os.path.join(tmp, "src"))
    sys.path.insert(0, tmp)
    sys.modules.setdefault("timm", types.ModuleType("timm"))
    return tmp


def format_and_filter():
    tmp = stub_reference_tree()
    try:
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            del sys.modules[k]
        from src.yolox_detector import YOLOXDetector
        from src.inference_engine_yolox import InferenceEngine
        out = {"versions": versions(), "cases": []}
        for seed, (oh, ow), wildlife in [(0, (1080, 1920), True), (1, (720, 1280), True), (2, (1080, 1920), False),
                                         (3, (640, 640), True)]:
            rows = seeded_rows(seed)
            det = YOLOXDetector(device="cpu", input_size=(640, 640), wildlife_only=wildlife)
            dets = det._format_model_output_to_detections(rows, oh, ow)
            formatted = json.loads(json.dumps(dets))
            eng = InferenceEngine(device="cpu", conf_threshold=0.25, min_box_area=300, max_det=20,
                                  class_confidence_overrides=dict(SHIPPED_CONF),
                                  class_size_constraints=json.loads(json.dumps(SHIPPED_SIZE)), wildlife_only=wildlife)
            kept = eng._post_process_detections(dets, np.zeros((oh, ow, 3), np.uint8))
            out["cases"].append({"seed": seed, "orig_hw": [oh, ow], "wildlife_only": wildlife, "formatted": formatted,
                                 "post_processed": json.loads(json.dumps(kept))})
        with open(os.path.join(HERE, "format_ref.json"), "w") as f:
            json.dump(out, f)
        print("format: %s" % [(len(c["formatted"]), len(c["post_processed"])) for c in out["cases"]])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def index_frame(h=1080, w=1920):
    """frame whose pixel values encode their own coordinates, so a crop reveals its window"""
    ys, xs = np.mgrid[0:h, 0:w]
    return np.stack([xs % 256, ys % 256, (xs // 256) * 16 + (ys // 256)], -1).astype(np.uint8)


def stage2_detections(seed=5, n=60, h=1080, w=1920):
    rng = np.random.default_rng(seed)
    dets = []
    for i in range(n):
        bw, bh = np.exp(rng.uniform(np.log(20), np.log(700), 2))
        x1, y1 = rng.uniform(-60, w - 10), rng.uniform(-60, h - 10)
        box = {"x1": float(x1), "y1": float(y1), "x2": float(x1 + bw), "y2": float(y1 + bh)}
        if i % 11 == 0:
            box["x1"], box["x2"] = box["x2"], box["x1"]
        d = {"class_id": int(rng.choice([14, 15, 16, 21, 0, 2])), "class_name": "x", "confidence": 0.9, "bbox": box}
        if i % 3 == 0:
            d["time_of_day"] = "day"
        dets.append(d)
    return dets


CANNED = [{"species": "Great Horned Owl", "confidence": 0.62, "class_id": 7, "taxonomic_level": "species"},
          {"species": "Gambel's Quail", "confidence": 0.41, "class_id": 9, "taxonomic_level": "genus"},
          {"species": "Aves", "confidence": 0.35, "class_id": 3, "taxonomic_level": "class"}]


def stage2():
    tmp = stub_reference_tree()
    try:
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            del sys.modules[k]
        from src.two_stage_pipeline_yolox import TwoStageDetectionPipeline

        class Recorder:
            def __init__(self):
                self.calls = []

            def classify(self, crop, top_k=5):
                c = np.asarray(crop)
                x1 = int(c[0, 0, 0]) + 256 * (int(c[0, 0, 2]) // 16)
                y1 = int(c[0, 0, 1]) + 256 * (int(c[0, 0, 2]) % 16)
                self.calls.append([x1, y1, x1 + c.shape[1], y1 + c.shape[0], int(top_k)])
                return [dict(r) for r in CANNED[:top_k]]

        frame = index_frame()
        out = {"versions": versions(), "cases": []}
        for pad, min_crop in [(20, 64), (0, 32), (35, 100)]:
            pipe = TwoStageDetectionPipeline(crop_padding_percent=pad, min_crop_size=min_crop, device="cpu")
            rec_b, rec_m = Recorder(), Recorder()
            pipe.add_species_classifier("bird", rec_b)
            pipe.add_species_classifier("mammal", rec_m)
            dets = pipe.process_detections(frame, stage2_detections())
            out["cases"].append({"pad": pad, "min_crop": min_crop, "bird_calls": rec_b.calls, "mammal_calls": rec_m.calls,
                                 "detections": json.loads(json.dumps(dets))})
        with open(os.path.join(HERE, "stage2_ref.json"), "w") as f:
            json.dump(out, f)
        print("stage2: %s" % [(len(c["bird_calls"]), len(c["mammal_calls"])) for c in out["cases"]])
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


def enhance():
    sys.path.insert(0, REF)
    # image_enhancement.py imports torch and cv2 only; the Real-ESRGAN imports are inside _load_realesrgan
    from src.image_enhancement import ImageEnhancer
    from util import enhance_crops
    enh = ImageEnhancer(method="clahe")
    out = {"versions": json.dumps(versions())}
    for i, crop in enumerate(enhance_crops()):
        out["out%d" % i] = enh.enhance(crop)
    np.savez_compressed(os.path.join(HERE, "enhance_ref.npz"), **out)
    print("enhance: %d crops" % len(enhance_crops()))


if __name__ == "__main__":
    which = sys.argv[1:] or ["motion", "format", "stage2", "enhance"]
    if "enhance" in which:
        enhance()
    if "motion" in which:
        motion_and_empty()
    if "format" in which:
        format_and_filter()
    if "stage2" in which:
        stage2()
