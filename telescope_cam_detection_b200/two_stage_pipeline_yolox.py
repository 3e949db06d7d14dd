"""TwoStageDetectionPipeline -- drop-in for the reference's src/two_stage_pipeline_yolox.py.

Same constructor, `add_species_classifier`, `classify_detection`, `process_detections`,
`get_stats`, same fields written on each detection.  The crop geometry is the reference's
integer arithmetic (int() truncation, padding = int(w*pct/100), clamps; :245-285), computed on
the host.  What changes: `process_detections` gathers ALL crops of a frame per category with one
batched CUDA kernel and one classifier forward (SpeciesClassifier.classify_crops) instead of one
slice + resize + forward + .item() sync per detection (:453-481).

Enhancement (`enhancement_config`, reference :90-121, :312-362): method "clahe" runs ImageEnhancer's batched CUDA
path on all crops of a category at once; the LRU cache keeps the reference's key (MD5 of the 8x8 INTER_AREA grey
thumbnail, :142-178) and eviction, so a cache hit returns exactly what the reference would return, and
`enhancement_cache_size <= 0` skips hashing altogether (the batched kernels are cheaper than the thumbnail).  Real-ESRGAN
is a third-party network outside the hot path: the enhancer fails to load and the pipeline continues without it, which is
the reference's own behaviour for a load failure (:119-121).  The species activity table
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
            try:
                from .image_enhancement import ImageEnhancer
                params = {"method": enhancement_config.get("method", "none"),
                          "device": enhancement_config.get("device", self.device)}
                for group in ("realesrgan", "clahe", "bilateral"):
                    for key, value in (enhancement_config.get(group) or {}).items():
                        params["%s_%s" % (group, key)] = value
                self.enhancer = ImageEnhancer(**params)
                logger.info("Image enhancer loaded (cache size: %d)", self.enhancement_cache_size)
            except Exception as e:  # noqa: BLE001 -- reference :119-121
                logger.error("Failed to load image enhancer (method=%s): %s", enhancement_config.get("method", "none"), e)
                logger.warning("Continuing without image enhancement")
                self.enhancer = None
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

    # ------------------------------------------------------------------ enhancement (reference :142-178, :312-362)
    @staticmethod
    def _compute_crop_hash(crop) -> str:
        import hashlib
        import cv2
        if not isinstance(crop, np.ndarray):
            crop = crop.detach().cpu().numpy()
            if crop.ndim == 3 and crop.shape[0] in (1, 3):
                crop = crop.transpose(1, 2, 0)
        small = cv2.resize(np.ascontiguousarray(crop), (8, 8), interpolation=cv2.INTER_AREA)
        gray = cv2.cvtColor(small, cv2.COLOR_BGR2GRAY) if small.ndim == 3 else small
        return hashlib.md5(gray.tobytes()).hexdigest()

    def _enhance_crops(self, crops: List[Any]) -> List[Any]:
        """All crops of one category: cache look-ups first, ONE batched enhancement for the misses."""
        t0 = time.time()
        if self.enhancement_cache_size <= 0:
            out = self.enhancer.enhance_batch(crops) if self.enhancer.method == "clahe" else [self.enhancer.enhance(c) for c in crops]
            dt = (time.time() - t0) * 1000 / max(len(crops), 1)
            self.enhancement_times.extend([dt] * len(crops))
            return out
        out: List[Any] = [None] * len(crops)
        keys = [self._compute_crop_hash(c) for c in crops]
        miss: Dict[str, int] = {}
        for n, key in enumerate(keys):
            if key in self.enhancement_cache:
                out[n] = self.enhancement_cache[key]
                self.enhancement_cache.move_to_end(key)
                self.cache_hits += 1
            elif key in miss:       # the same thumbnail twice in one frame: the reference's second look-up is a hit
                self.cache_hits += 1
            else:
                miss[key] = n
                self.cache_misses += 1
        if miss:
            todo = [crops[n] for n in miss.values()]
            done = self.enhancer.enhance_batch(todo) if self.enhancer.method == "clahe" else [self.enhancer.enhance(c) for c in todo]
            dt = (time.time() - t0) * 1000 / len(todo)
            for key, img in zip(miss, done):
                self.enhancement_cache[key] = img
                self.enhancement_times.append(dt)
                if len(self.enhancement_cache) > self.enhancement_cache_size:
                    self.enhancement_cache.popitem(last=False)
        for n, key in enumerate(keys):
            if out[n] is None:
                out[n] = self.enhancement_cache.get(key)
                if out[n] is None:   # evicted within this very batch (cache smaller than the frame's crop count)
                    out[n] = done[list(miss).index(key)]
        return out

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
                enhanced = None
                if self.enhancer is not None and self.enhancer.method != "none":
                    try:
                        enhanced = self._enhance_crops([frame[b[1]:b[3], b[0]:b[2]] for b in boxes])
                    except Exception as e:  # noqa: BLE001 -- reference :364-366: classify the plain crop
                        logger.warning("Enhancement failed: %s, using original crop", e)
                for k in sorted(set(top_ks)):
                    sel = [n for n, kk in enumerate(top_ks) if kk == k]
                    if enhanced is not None:
                        if hasattr(clf, "classify_crops"):
                            res = [clf.classify_crops(enhanced[n], [(0, 0, int(enhanced[n].shape[1]), int(enhanced[n].shape[0]))],
                                                      top_k=k)[0] for n in sel]
                        else:
                            res = [clf.classify(enhanced[n], top_k=k) for n in sel]
                    elif hasattr(clf, "classify_crops"):
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
        if self.enhancer is not None:    # reference :493-503
            stats["enhancement_method"] = self.enhancer.method
            if self.enhancement_times:
                stats["avg_enhancement_ms"] = float(np.mean(self.enhancement_times))
                stats["enhancement_count"] = len(self.enhancement_times)
            total = self.cache_hits + self.cache_misses
            if total > 0:
                stats["enhancement_cache_hit_rate"] = float(self.cache_hits / total)
                stats["enhancement_cache_hits"] = self.cache_hits
                stats["enhancement_cache_misses"] = self.cache_misses
                stats["enhancement_cache_size"] = len(self.enhancement_cache)
                stats["enhancement_cache_max_size"] = self.enhancement_cache_size
        if self.classification_times:
            stats["avg_classification_ms"] = float(np.mean(self.classification_times))
            stats["classification_count"] = len(self.classification_times)
        return stats
