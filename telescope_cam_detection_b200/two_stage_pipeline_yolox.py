"""TwoStageDetectionPipeline -- drop-in for the reference's src/two_stage_pipeline_yolox.py.

Same constructor, `add_species_classifier`, `classify_detection`, `process_detections`,
`get_stats`, same fields written on each detection.  The crop geometry is the reference's
integer arithmetic (int() truncation, padding = int(w*pct/100), clamps; :245-285), computed on
the host.  What changes: `process_detections` gathers ALL crops of a frame per category with one
batched CUDA kernel and one classifier forward (SpeciesClassifier.classify_crops) instead of one
slice + resize + forward + .item() sync per detection (:453-481).

Not rebuilt (out of scope / "next" N3): the Real-ESRGAN / CLAHE enhancer and
its LRU cache -- `enhancement_config` is accepted and ignored with a warning, which is also what
the reference does when the enhancer fails to load (:104-106).  The species activity table
(src/species_activity_patterns.py) is used if importable, else every species counts as active.
"""
import logging
import time
from collections import OrderedDict, deque
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

from .bbox_utils import ensure_valid_bbox
from .coco_constants import CLASS_ID_TO_CATEGORY

logger = logging.getLogger(__name__)


def _activity_fn():
    for mod in ("species_activity_patterns", "src.species_activity_patterns"):
        try:
            m = __import__(mod, fromlist=["is_species_likely_active"])
            return m.is_species_likely_active
        except Exception:
            continue
    return lambda species, time_of_day: True


class TwoStageDetectionPipeline:
    def __init__(self, enable_species_classification: bool = True, stage2_confidence_threshold: float = 0.3,
                 device: str = "cuda:0", enhancement_config: Optional[Dict[str, Any]] = None,
                 rejected_taxonomic_levels: Optional[List[str]] = None, crop_padding_percent: int = 20,
                 min_crop_size: int = 64, time_of_day_top_k: int = 5, time_of_day_penalty: float = 0.3,
                 enhancement_cache_size: int = 100):
        self.enable_species_classification = enable_species_classification
        self.stage2_confidence_threshold = stage2_confidence_threshold
        self.device = device
        self.rejected_taxonomic_levels = rejected_taxonomic_levels or ["order", "class"]
        self.crop_padding_percent = crop_padding_percent
        self.min_crop_size = min_crop_size
        self.time_of_day_top_k = time_of_day_top_k
        self.time_of_day_penalty = time_of_day_penalty
        self.species_classifiers: Dict[str, Any] = {}
        self.class_id_to_category = CLASS_ID_TO_CATEGORY
        self.enhancement_cache_size = enhancement_cache_size
        self.enhancement_cache: OrderedDict = OrderedDict()
        self.cache_hits = 0
        self.cache_misses = 0
        self.enhancer = None
        if enhancement_config and enhancement_config.get("enabled", False):
            logger.warning("image enhancement (%s) is not part of the B200 hot path; continuing without it",
                           enhancement_config.get("method", "none"))
        self.enhancement_times = deque(maxlen=1000)
        self.classification_times = deque(maxlen=1000)
        self._is_active = _activity_fn()

    def add_species_classifier(self, category: str, classifier: Any):
        self.species_classifiers[category] = classifier
        logger.info("Added species classifier for category: %s", category)

    @staticmethod
    def _set_detection_species_fields(detection, species, confidence, category, taxonomic_level):
        detection["species"] = species
        detection["species_confidence"] = float(confidence)
        detection["stage2_category"] = category
        detection["taxonomic_level"] = taxonomic_level

    # ------------------------------------------------------------------ integer crop geometry (A13)
    def _crop_box(self, frame_hw: Tuple[int, int], detection: Dict[str, Any]):
        """-> ('skip', None) leave the detection untouched; ('none', category) write empty species
        fields; ('crop', category, (x1,y1,x2,y2)) classify this window.  Reference :219-285."""
        bbox = detection.get("bbox", {})
        valid = ensure_valid_bbox(bbox)
        if valid != bbox:
            detection["bbox"] = valid
        bbox = valid
        category = self.class_id_to_category.get(detection.get("class_id"))
        if category not in self.species_classifiers:
            detection["species"] = None
            detection["species_confidence"] = 0.0
            return ("skip", None)
        x1, y1, x2, y2 = int(bbox["x1"]), int(bbox["y1"]), int(bbox["x2"]), int(bbox["y2"])
        cw, ch = x2 - x1, y2 - y1
        if cw < self.min_crop_size or ch < self.min_crop_size:
            return ("none", category)
        px = int(cw * self.crop_padding_percent / 100)
        py = int(ch * self.crop_padding_percent / 100)
        h, w = frame_hw
        xa = max(0, min(x1 - px, w - 1))
        ya = max(0, min(y1 - py, h - 1))
        xb = max(0, min(x2 + px, w))
        yb = max(0, min(y2 + py, h))
        if xb <= xa or yb <= ya:
            return ("none", category)
        return ("crop", category, (xa, ya, xb, yb))

    # ------------------------------------------------------------------ result logic (reference :372-446)
    def _apply_results(self, detection, category, results):
        time_of_day = detection.get("time_of_day")
        if not results:
            self._set_detection_species_fields(detection, None, 0.0, category, None)
            return
        if time_of_day:
            for r in results:
                r["confidence_original"] = r["confidence"]
                r["activity_boosted"] = bool(self._is_active(r["species"], time_of_day))
                if not r["activity_boosted"]:
                    r["confidence"] = r["confidence"] * self.time_of_day_penalty
            results = sorted(results, key=lambda r: r["confidence"], reverse=True)  # stable, like list.sort
        top = results[0]
        level = top.get("taxonomic_level", "species")
        if level in self.rejected_taxonomic_levels:
            self._set_detection_species_fields(detection, None, 0.0, category, None)
        else:
            self._set_detection_species_fields(detection, top["species"], top["confidence"], category, level)

    def classify_detection(self, frame, detection: Dict[str, Any]) -> Dict[str, Any]:
        """One detection (reference :203-451)."""
        if not self.enable_species_classification:
            return detection
        return self.process_detections(frame, [detection])[0]

    def process_detections(self, frame, detections: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        if not self.enable_species_classification:
            for det in detections:
                det["species"] = None
                det["species_confidence"] = 0.0
            return detections
        hw = (int(frame.shape[0]), int(frame.shape[1]))
        per_cat: Dict[str, List[Tuple[int, Tuple[int, int, int, int]]]] = {}
        for i, det in enumerate(detections):
            plan = self._crop_box(hw, det)
            if plan[0] == "none":
                self._set_detection_species_fields(det, None, 0.0, plan[1], None)
            elif plan[0] == "crop":
                per_cat.setdefault(plan[1], []).append((i, plan[2]))
        for category, items in per_cat.items():
            clf = self.species_classifiers[category]
            boxes = [b for _, b in items]
            # top-1 unless a time-of-day re-rank may reorder the candidates (:376-377); crops asking for
            # the same k share one gather + forward, so filtering happens on exactly the rows the
            # reference would have asked for
            top_ks = [self.time_of_day_top_k if detections[i].get("time_of_day") else 1 for i, _ in items]
            t0 = time.time()
            try:
                all_results: List[Any] = [None] * len(items)
                for k in sorted(set(top_ks)):
                    sel = [n for n, kk in enumerate(top_ks) if kk == k]
                    if hasattr(clf, "classify_crops"):
                        res = clf.classify_crops(frame, [boxes[n] for n in sel], top_k=k)
                    else:  # any classifier with the reference's per-crop API
                        res = [clf.classify(frame[boxes[n][1]:boxes[n][3], boxes[n][0]:boxes[n][2]], top_k=k)
                               for n in sel]
                    for n, r in zip(sel, res):
                        all_results[n] = r
            except Exception as e:
                logger.error("Species classification failed (category=%s): %s", category, e, exc_info=True)
                for i, _ in items:
                    detections[i]["species"] = None
                    detections[i]["species_confidence"] = 0.0
                    detections[i]["taxonomic_level"] = None
                continue
            dt = (time.time() - t0) * 1000 / max(len(items), 1)
            for (i, _), results in zip(items, all_results):
                self.classification_times.append(dt)
                self._apply_results(detections[i], category, results)
        return detections

    def get_stats(self) -> Dict[str, Any]:
        stats = {"stage2_enabled": self.enable_species_classification,
                 "classifiers_loaded": list(self.species_classifiers.keys()),
                 "num_classifiers": len(self.species_classifiers), "enhancement_enabled": self.enhancer is not None}
        if self.classification_times:
            stats["avg_classification_ms"] = float(np.mean(self.classification_times))
            stats["classification_count"] = len(self.classification_times)
        return stats
