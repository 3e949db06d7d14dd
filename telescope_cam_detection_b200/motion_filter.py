"""MotionFilter / AdaptiveMotionFilter -- drop-ins for the reference's src/motion_filter.py.

The MOG2 model (5 Gaussians per pixel, fp32) lives on the GPU (csrc/motion.cu mog2_apply_kernel);
shadow threshold, 21x21 fixed-point Gaussian, second threshold and per-bbox countNonZero are one
smem-staged kernel plus a ROI reduction.  The reference updates the background model ONCE PER
DETECTION with the same frame and never on frames without detections (src/motion_filter.py:88,
155-162): detection i is judged on the mask after i+1 updates.  That order is
kept here, so verdicts and motion ratios equal the reference's.
"""
import logging
import threading
from datetime import datetime
from typing import Any, Dict, List, Tuple

import numpy as np

from .bbox_utils import ensure_valid_bbox
from .engine import Mog2Engine, device_index

logger = logging.getLogger(__name__)


class MotionFilter:
    def __init__(self, history: int = 500, var_threshold: int = 16, detect_shadows: bool = True,
                 min_motion_area: int = 100, motion_required: bool = True, motion_blur_size: int = 21,
                 min_motion_ratio: float = 0.05, device=0):
        self.history = history
        self.var_threshold = var_threshold
        self.detect_shadows = detect_shadows
        self.min_motion_area = min_motion_area
        self.motion_required = motion_required
        self.motion_blur_size = motion_blur_size if motion_blur_size % 2 == 1 else motion_blur_size + 1
        self.min_motion_ratio = min_motion_ratio
        self.device = device_index(device)   # ordinal, "cuda:N" or N
        self.bg_subtractor = None  # Mog2Engine, created at the first frame (needs the frame size)
        self._stale = True
        self.total_frames = 0
        self.total_detections_filtered = 0
        logger.info("MotionFilter initialized (history=%s, var_threshold=%s)", history, var_threshold)

    # -- model lifetime
    def _model(self, frame) -> Mog2Engine:
        h, w = frame.shape[0], frame.shape[1]
        m = self.bg_subtractor
        if m is None or self._stale or (m.h, m.w) != (h, w):
            if m is not None:
                m.close()
            m = Mog2Engine(h, w, self.history, self.var_threshold, self.detect_shadows, self.device)
            self.bg_subtractor = m
            self._stale = False
        return m

    def cleanup(self):
        if self.bg_subtractor is not None:
            self.bg_subtractor.close()
            self.bg_subtractor = None
            logger.debug("Motion filter resources released")

    def reset_background(self):
        self._stale = True
        logger.info("Background model reset")

    # -- data path
    def has_motion_in_bbox(self, frame, bbox: Dict[str, float], min_motion_pixels: int = 10) -> Tuple[bool, float]:
        """One model update with `frame`, mask clean-up, motion ratio inside the bbox (reference :68-127)."""
        m = self._model(frame)
        m.apply(frame, repeats=1)
        box = ensure_valid_bbox(bbox)
        h, w = frame.shape[0], frame.shape[1]
        x1, y1 = max(0, int(box["x1"])), max(0, int(box["y1"]))
        x2, y2 = min(w, int(box["x2"])), min(h, int(box["y2"]))
        if x2 <= x1 or y2 <= y1:
            return False, 0.0
        # without shadow detection the raw mask is {0,255} and the reference skips the first threshold;
        # (p > 200) is then the identity, so one code path serves both
        counts, _ = m.roi_counts(np.array([[x1, y1, x2, y2]], np.int32), 200, self.motion_blur_size, 25)
        motion_pixels = int(counts[0])
        ratio = motion_pixels / ((x2 - x1) * (y2 - y1))
        return (motion_pixels >= min_motion_pixels and ratio > self.min_motion_ratio), ratio

    def filter_detections(self, frame, detections: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        if not self.motion_required or len(detections) == 0:
            return detections
        self.total_frames += 1
        kept = []
        for det in detections:
            moving, ratio = self.has_motion_in_bbox(frame, det["bbox"])
            if moving:
                det["has_motion"] = True
                det["motion_ratio"] = ratio
                kept.append(det)
            else:
                self.total_detections_filtered += 1
        if len(kept) < len(detections):
            logger.debug("Motion filter: %d -> %d detections", len(detections), len(kept))
        return kept

    def update_params(self, config: Dict[str, Any]):
        """Hot reload (reference :187-235): model-defining keys rebuild the background model."""
        changed, rebuild = [], False
        for key in ("history", "var_threshold", "detect_shadows"):
            if key in config and config[key] != getattr(self, key):
                setattr(self, key, config[key])
                changed.append("%s: %s" % (key, config[key]))
                rebuild = True
        for key in ("min_motion_area", "min_motion_ratio"):
            if key in config and config[key] != getattr(self, key):
                setattr(self, key, config[key])
                changed.append("%s: %s" % (key, config[key]))
        if "motion_blur_size" in config:
            k = config["motion_blur_size"]
            k = k if k % 2 == 1 else k + 1
            if k != self.motion_blur_size:
                self.motion_blur_size = k
                changed.append("motion_blur_size: %s" % k)
        if rebuild:
            self._stale = True
            logger.info("Motion filter background subtractor recreated with new parameters")
        if changed:
            logger.info("MotionFilter params updated: %s", ", ".join(changed))

    def get_stats(self) -> Dict[str, Any]:
        return {"total_frames": self.total_frames, "total_detections_filtered": self.total_detections_filtered,
                "motion_required": self.motion_required}


class AdaptiveMotionFilter(MotionFilter):
    """Day/night variance threshold (reference :246-304)."""

    def __init__(self, day_var_threshold: int = 16, night_var_threshold: int = 32, day_start_hour: int = 6,
                 day_end_hour: int = 20, **kwargs):
        super().__init__(var_threshold=day_var_threshold, **kwargs)
        self.day_var_threshold = day_var_threshold
        self.night_var_threshold = night_var_threshold
        self.day_start_hour = day_start_hour
        self.day_end_hour = day_end_hour
        self._lock = threading.Lock()

    def _is_daytime(self) -> bool:
        return self.day_start_hour <= datetime.now().hour < self.day_end_hour

    def filter_detections(self, frame, detections):
        wanted = self.day_var_threshold if self._is_daytime() else self.night_var_threshold
        with self._lock:
            if wanted != self.var_threshold:
                self.var_threshold = wanted
                self._stale = True
        return super().filter_detections(frame, detections)
