"""StreamBank -- the per-camera EmptyFrameFilter and MotionFilter of every stream of one GPU, batched.

The reference keeps one EmptyFrameFilter per InferenceEngine (src/inference_engine_yolox.py:589-593) and one
MotionFilter per DetectionProcessor (src/detection_processor.py:274-275), each a chain of synchronous OpenCV
calls per frame (and, for MOG2, per detection: src/motion_filter.py:88,161-162).  With 8 streams per GPU
(BASELINE configs[2]) this class gives the same verdicts -- bit-exact masks, same counters, same
per-detection model updates -- from ONE upload per frame and ONE kernel launch per step for the gate, plus one
background-model pass and one clean-up launch per round of up to 16 detections per stream.
The frames live on the device after `upload`; `device_frames()` hands them to YOLOXDetector.submit_batch /
detect_batch as CUDA tensors, so preprocess reads the same copy (no second upload).
"""
import logging
from typing import Any, Dict, List, Optional

import numpy as np

from .bbox_utils import ensure_valid_bbox
from .engine import StreamBankEngine
from .stream_shard import DeviceView

logger = logging.getLogger(__name__)


class StreamBank:
    def __init__(self, n_streams: int, frame_hw, device=0, history: int = 500, var_threshold: int = 16,
                 detect_shadows: bool = True, motion_blur_size: int = 21, min_motion_ratio: float = 0.05,
                 gate_min_motion_area: int = 200, gate_threshold: int = 25, gate_blur_size: int = 21):
        if motion_blur_size != gate_blur_size:
            raise ValueError("the batched kernels use one blur size for the gate and the mask clean-up")
        self.n = int(n_streams)
        self.h, self.w = int(frame_hw[0]), int(frame_hw[1])
        self.min_motion_ratio = min_motion_ratio
        self.gate_min_motion_area = gate_min_motion_area
        self.gate_threshold = gate_threshold
        self.blur_size = motion_blur_size if motion_blur_size % 2 == 1 else motion_blur_size + 1
        self.engine = StreamBankEngine(self.n, self.h, self.w, history, var_threshold, detect_shadows, self.blur_size,
                                       device)
        # per-stream counters with the reference's names (EmptyFrameFilter.get_stats / MotionFilter.get_stats)
        self.gate_stats = [{"total_frames": 0, "skipped_frames": 0, "motion_frames": 0} for _ in range(self.n)]
        self.motion_stats = [{"total_frames": 0, "total_detections_filtered": 0} for _ in range(self.n)]
        self._tensors = None

    def close(self):
        self.engine.close()

    # ------------------------------------------------------------------ frames
    def upload(self, frames: List[Optional[np.ndarray]]) -> None:
        """Start the (asynchronous) upload of the NEXT step's frames.  has_motion() adopts them; upload() for step
        i+1 may be called as soon as has_motion() of step i has returned, i.e. under the detector and the motion
        filter of step i."""
        self.engine.upload(frames)

    def device_frames(self):
        """One CUDA uint8 tensor (h, w, 3) per stream viewing the bank's copy of the CURRENT step's frame (the one the
        last has_motion() adopted; the bank alternates between two buffers so that the next upload can overlap)."""
        import torch
        if self._tensors is None:
            self._tensors = {}
        key = self.engine.frame_ptr(0)
        if key not in self._tensors:
            self._tensors[key] = [torch.as_tensor(DeviceView(self.engine.frame_ptr(s), (self.h, self.w, 3), "|u1"),
                                                  device="cuda") for s in range(self.n)]
        return self._tensors[key]

    # ------------------------------------------------------------------ EmptyFrameFilter.has_motion, all streams
    def has_motion(self) -> List[Optional[bool]]:
        """Verdict per stream for the frames of the last upload (None: the stream had no frame), reference
        src/empty_frame_filter.py:51-98: first frame -> True, else changed pixels >= min_motion_area."""
        counts = self.engine.gate(self.gate_threshold)
        out: List[Optional[bool]] = []
        for s, c in enumerate(counts.tolist()):
            if c == -2:
                out.append(None)
                continue
            st = self.gate_stats[s]
            st["total_frames"] += 1
            moving = c < 0 or c >= self.gate_min_motion_area
            st["motion_frames" if moving else "skipped_frames"] += 1
            out.append(moving)
        return out

    # ------------------------------------------------------------------ MotionFilter.filter_detections, all streams
    def filter_detections(self, detections: List[Optional[List[Dict[str, Any]]]],
                          min_motion_pixels: int = 10) -> List[List[Dict[str, Any]]]:
        """detections[s]: the detection dicts of stream s for the frame of the last upload (None / [] = none).
        Same result as calling the reference's MotionFilter.filter_detections(frame_s, detections[s]) on every
        stream: one model update per detection, detection i judged on the mask after its own update."""
        rois, offsets, geo = [], [0], []
        for s in range(self.n):
            dets = detections[s] or []
            for det in dets:
                box = ensure_valid_bbox(det["bbox"])
                x1, y1 = max(0, int(box["x1"])), max(0, int(box["y1"]))
                x2, y2 = min(self.w, int(box["x2"])), min(self.h, int(box["y2"]))
                if x2 <= x1 or y2 <= y1:
                    rois.append((0, 0, 0, 0))      # no pixels to count, but the model is still updated once
                    geo.append(None)
                else:
                    rois.append((x1, y1, x2, y2))
                    geo.append((x2 - x1) * (y2 - y1))
            offsets.append(len(rois))
        counts = self.engine.motion(np.array(rois, np.int32).reshape(-1, 4), np.array(offsets, np.int32),
                                    200, self.blur_size, 25) if rois else np.zeros((0,), np.int32)
        out: List[List[Dict[str, Any]]] = []
        for s in range(self.n):
            dets = detections[s] or []
            if not dets:
                out.append(list(dets))
                continue
            st = self.motion_stats[s]
            st["total_frames"] += 1
            kept = []
            for i, det in enumerate(dets):
                k = offsets[s] + i
                area = geo[k]
                pixels = int(counts[k])
                ratio = pixels / area if area else 0.0
                if area and pixels >= min_motion_pixels and ratio > self.min_motion_ratio:
                    det["has_motion"] = True
                    det["motion_ratio"] = ratio
                    kept.append(det)
                else:
                    st["total_detections_filtered"] += 1
            out.append(kept)
        return out

    def get_stats(self) -> Dict[str, Any]:
        gate = []
        for st in self.gate_stats:
            rate = st["skipped_frames"] / max(st["total_frames"], 1)
            gate.append(dict(st, skip_rate=rate, skip_rate_percent=rate * 100))
        return {"empty_frame_filter": gate, "motion_filter": [dict(m, motion_required=True) for m in self.motion_stats],
                "kernels": self.engine.last_ms()}
