"""InferenceEngine -- drop-in for the reference's src/inference_engine_yolox.py (:32-805).

Per-camera thread between an input queue of frame dicts and an output queue of result dicts,
either driving its own detector (standalone: sparse keyframes, empty-frame gate, memory
pressure handling, OOM retry) or handing frames to a SharedInferenceCoordinator (batched).
Constructor keywords, attributes, result-dict and `get_stats()` keys follow the reference.
The detector, the empty-frame gate and Stage 2 underneath are the libtcd-backed classes of this
package.  The reference's GPU->CPU load fallback (:236-272) does not exist here: north_star rules
out a CPU path, so a failed GPU load makes `load_model()` return False.
"""
import logging
import time
from queue import Empty, Queue
from threading import Event, Lock, Thread
from typing import Any, Dict, List, Optional

from .bbox_utils import ensure_valid_bbox
from .constants import (ERROR_SLEEP_SECONDS, FPS_CALCULATION_INTERVAL_SECONDS, LOG_DROPPED_EVERY_N,
                        MIN_TIME_DELTA, QUEUE_GET_TIMEOUT_SECONDS, THREAD_JOIN_TIMEOUT_SECONDS)
from .empty_frame_filter import EmptyFrameFilter
from .memory_manager import MemoryManager, MemoryPressure

logger = logging.getLogger(__name__)


def _is_cuda_oom(exc) -> bool:
    name = type(exc).__name__
    return name == "OutOfMemoryError" or "out of memory" in str(exc).lower()


class InferenceEngine:
    def __init__(self, model_name: str = "yolox-s", model_path: str = "models/yolox/yolox_s.pth",
                 device: str = "cuda:0", conf_threshold: float = 0.25, nms_threshold: float = 0.45,
                 input_size: tuple = (640, 640), input_queue: Optional[Queue] = None,
                 output_queue: Optional[Queue] = None, min_box_area: int = 0, max_det: int = 300,
                 use_two_stage: bool = False, two_stage_pipeline: Optional[Any] = None,
                 class_confidence_overrides: Optional[Dict[str, float]] = None,
                 class_size_constraints: Optional[Dict[str, Dict[str, int]]] = None, wildlife_only: bool = True,
                 shared_coordinator: Optional[Any] = None, camera_id: Optional[str] = None,
                 camera_name: Optional[str] = None, empty_frame_filter_config: Optional[Dict[str, Any]] = None,
                 sparse_detection_config: Optional[Dict[str, Any]] = None, detector_type: str = "yolox",
                 rtdetr_config_path: Optional[str] = None, **kwargs):
        self.model_name, self.model_path, self.device = model_name, model_path, device
        self.conf_threshold, self.nms_threshold, self.input_size = conf_threshold, nms_threshold, input_size
        self.input_queue, self.output_queue = input_queue, output_queue
        self.min_box_area, self.max_det = min_box_area, max_det
        self.use_two_stage, self.two_stage_pipeline = use_two_stage, two_stage_pipeline
        self.class_confidence_overrides = class_confidence_overrides or {}
        self.class_size_constraints = class_size_constraints or {}
        self.wildlife_only = wildlife_only
        self.shared_coordinator = shared_coordinator
        self.camera_id = camera_id or "default"
        self.camera_name = camera_name or "Default Camera"
        self.detector_type = detector_type.lower()
        if self.detector_type not in ("yolox", "rtdetr"):
            raise ValueError("Invalid detector_type '%s'. Must be 'yolox' or 'rtdetr'." % detector_type)
        self.rtdetr_config_path = rtdetr_config_path
        self._extra = kwargs  # e.g. state_dict= for in-memory weights; unknown keys are tolerated like upstream

        self.detector: Optional[Any] = None
        self.is_loaded = False
        self.stop_event = Event()
        self.inference_thread: Optional[Thread] = None
        self.settings_lock = Lock()

        self.inference_count = 0
        self.total_inference_count = 0
        self.total_inference_time = 0.0
        self.avg_inference_time = 0.0
        self.fps = 0.0
        self.last_fps_check = time.time()
        self.dropped_results = 0
        self.last_drop_warning_time = 0
        self.drop_count_since_warning = 0

        self._filter_id: Optional[int] = None      # device-side copy of the per-class tables (K7), see _device_filter
        self._filter_owner: Optional[Any] = None
        self.memory_manager = MemoryManager(device=self.device)
        self.degradation_active = False
        self.original_input_size = self.input_size

        self.empty_frame_filter: Optional[EmptyFrameFilter] = None
        efc = empty_frame_filter_config
        if efc and efc.get("enabled", False):
            self.empty_frame_filter = EmptyFrameFilter(min_motion_area=efc.get("min_motion_area", 200),
                                                       threshold=efc.get("threshold", 25),
                                                       blur_size=efc.get("blur_size", 21),
                                                       device=self.device if str(self.device).startswith("cuda") else 0)
        self.sparse_detection_enabled = False
        self.keyframe_interval = 1
        self.sparse_mode = "simple"
        sdc = sparse_detection_config
        if sdc and sdc.get("enabled", False):
            self.sparse_detection_enabled = True
            k = sdc.get("keyframe_interval", 3)
            if k < 1:
                logger.warning("Invalid keyframe_interval=%s, must be >= 1. Using default of 3.", k)
                k = 3
            self.keyframe_interval = k
            self.sparse_mode = sdc.get("mode", "simple")
        self.frame_counter = 0
        self.last_detections: List[Dict[str, Any]] = []

    # ------------------------------------------------------------------ lifecycle
    def load_model(self) -> bool:
        try:
            if self.shared_coordinator:  # the coordinator owns the detector
                self.is_loaded = True
                return True
            if self.detector_type == "rtdetr":
                if not self.rtdetr_config_path:
                    raise ValueError("rtdetr_config_path is required for RT-DETR detector")
                from .rtdetr_detector import RTDETRDetector
                self.detector = RTDETRDetector(config_path=self.rtdetr_config_path, model_path=self.model_path,
                                               device=self.device, conf_threshold=self.conf_threshold,
                                               input_size=self.input_size, wildlife_only=self.wildlife_only)
            else:
                from .yolox_detector import YOLOXDetector
                self.detector = YOLOXDetector(model_name=self.model_name, model_path=self.model_path,
                                              device=self.device, conf_threshold=self.conf_threshold,
                                              nms_threshold=self.nms_threshold, input_size=self.input_size,
                                              wildlife_only=self.wildlife_only,
                                              state_dict=self._extra.get("state_dict"), max_batch=1)
            if not self.detector.load_model():
                logger.error("Failed to load %s model on %s (no CPU fallback in the B200 build)",
                             self.detector_type.upper(), self.device)
                return False
            self.is_loaded = True
            return True
        except Exception as e:
            logger.error("Failed to load inference engine: %s", e, exc_info=True)
            return False

    def start(self) -> bool:
        if not self.is_loaded and not self.load_model():
            return False
        if not self.input_queue or not self.output_queue:
            logger.error("Input and output queues must be provided")
            return False
        self.stop_event.clear()
        self.inference_thread = Thread(target=self._inference_loop, daemon=True)
        self.inference_thread.start()
        return True

    def stop(self):
        self.stop_event.set()
        if self.inference_thread:
            self.inference_thread.join(timeout=THREAD_JOIN_TIMEOUT_SECONDS)
            if self.inference_thread.is_alive():
                logger.error("Inference thread did not stop after %ss (blocked?)", THREAD_JOIN_TIMEOUT_SECONDS)

    # ------------------------------------------------------------------ thread body
    def _account(self, inference_time):
        self.inference_count += 1
        self.total_inference_count += 1
        self.total_inference_time += inference_time
        self.avg_inference_time = self.total_inference_time / self.total_inference_count
        now = time.time()
        if now - self.last_fps_check >= FPS_CALCULATION_INTERVAL_SECONDS:
            self.fps = self.inference_count / (now - self.last_fps_check)
            self.last_fps_check = now
            self.inference_count = 0

    @staticmethod
    def _result(frame, frame_id, ts, inference_time, detections, camera_id, camera_name):
        return {"frame_id": frame_id, "timestamp": ts, "inference_time": inference_time, "detections": detections,
                "frame_shape": frame.shape, "camera_id": camera_id, "camera_name": camera_name}

    def _inference_loop(self):
        while not self.stop_event.is_set():
            try:
                try:
                    item = self.input_queue.get(timeout=QUEUE_GET_TIMEOUT_SECONDS)
                except Empty:
                    continue
                frame, ts, fid = item["frame"], item["timestamp"], item["frame_id"]
                cid = item.get("camera_id", self.camera_id)
                cname = item.get("camera_name", self.camera_name)
                t0 = time.time()
                if self.shared_coordinator:
                    cb = (lambda raw, f=frame, i=fid, t=ts, c=cid, n=cname, s=t0:
                          self._handle_inference_callback(raw, f, i, t, c, n, s))
                    tid = self._device_filter()
                    if tid is not None:
                        self.shared_coordinator.infer_async(frame=frame, camera_id=cid, callback=cb, filter_id=tid)
                    else:
                        self.shared_coordinator.infer_async(frame=frame, camera_id=cid, callback=cb)
                else:
                    dets = self._run_inference(frame)
                    dt = time.time() - t0
                    self._account(dt)
                    self._queue_result(self._result(frame, fid, ts, dt, dets, cid, cname))
            except (RuntimeError, ValueError, TypeError, AttributeError) as e:
                logger.error("Error in inference loop for camera %s: %s", self.camera_id, e, exc_info=True)
                if self.stop_event.is_set():
                    break
                time.sleep(ERROR_SLEEP_SECONDS)

    def _handle_inference_callback(self, raw_detections, frame, frame_id, frame_timestamp, camera_id, camera_name,
                                   inference_start):
        """Runs on the coordinator thread (reference :390-447)."""
        try:
            dets = self._post_process_detections(raw_detections, frame)
            dt = time.time() - inference_start
            self._account(dt)
            self._queue_result(self._result(frame, frame_id, frame_timestamp, dt, dets, camera_id, camera_name))
        except Exception as e:
            logger.error("Error in inference callback for camera %s: %s", camera_id, e, exc_info=True)

    def _queue_result(self, result: Dict[str, Any]):
        try:
            self.output_queue.put_nowait(result)
        except Exception:
            self.dropped_results += 1
            self.drop_count_since_warning += 1
            now = time.time()
            since = now - self.last_drop_warning_time
            if self.dropped_results % LOG_DROPPED_EVERY_N == 0 or since >= 10.0:
                logger.warning("Output queue full: dropped %d results (%.2f/s, %.1f%% overall)", self.dropped_results,
                               self.drop_count_since_warning / max(since, MIN_TIME_DELTA),
                               100.0 * self.dropped_results / max(self.total_inference_count, 1))
                self.last_drop_warning_time = now
                self.drop_count_since_warning = 0

    # ------------------------------------------------------------------ A6: per-class filtering
    def _device_filter(self, refresh: bool = False) -> Optional[int]:
        """This camera's conf / area / max_det tables as a device-side filter table of the detector that will
        run its frames (its own, or the coordinator's), so that the rows crossing PCIe are already the ones
        _post_process_detections keeps.  None when the detector has no such facility (any other detector)."""
        det = self.detector
        if det is None and self.shared_coordinator is not None and getattr(self.shared_coordinator,
                                                                            "supports_filter_ids", False):
            det = getattr(self.shared_coordinator, "detector", None)
        if det is None or not hasattr(det, "register_class_filter"):
            return None
        if self._filter_id is not None and self._filter_owner is det and not refresh:
            return self._filter_id
        with self.settings_lock:
            args = (self.conf_threshold, dict(self.class_confidence_overrides), self.min_box_area,
                    dict(self.class_size_constraints), self.max_det)
        try:
            self._filter_id = det.register_class_filter(*args, table=self._filter_id if self._filter_owner is det
                                                        else None)
            self._filter_owner = det
        except Exception as e:
            logger.warning("device-side class filter unavailable (%s); filtering on the host only", e)
            self._filter_id, self._filter_owner = None, None
        return self._filter_id

    def _post_process_detections(self, detections: List[Dict[str, Any]], frame) -> List[Dict[str, Any]]:
        """ensure_valid_bbox (area becomes float), per-class confidence override else the global
        threshold, global minimum area, per-class min/max area, stop at max_det, then Stage 2
        (reference :479-540).  Input order is preserved."""
        with self.settings_lock:
            conf, min_area, max_det = self.conf_threshold, self.min_box_area, self.max_det
            overrides = self.class_confidence_overrides.copy()
            limits = self.class_size_constraints.copy()
        kept = []
        for det in detections:
            det["bbox"] = ensure_valid_bbox(det["bbox"])
            name, area = det["class_name"], det["bbox"]["area"]
            if det["confidence"] < overrides.get(name, conf):
                continue
            if min_area > 0 and area < min_area:
                continue
            lim = limits.get(name)
            if lim is not None and (("min" in lim and area < lim["min"]) or ("max" in lim and area > lim["max"])):
                continue
            kept.append(det)
            if len(kept) >= max_det:
                break
        if self.use_two_stage and self.two_stage_pipeline:
            kept = self.two_stage_pipeline.process_detections(frame, kept)
        return kept

    # ------------------------------------------------------------------ standalone path
    def _process_and_store_detections(self, frame):
        tid = self._device_filter()
        raw = self.detector.detect(frame) if tid is None else self.detector.detect(frame, filter_id=tid)
        dets = self._post_process_detections(raw, frame)
        self.last_detections = dets
        return dets

    def _run_inference(self, frame) -> List[Dict[str, Any]]:
        if self.detector is None:
            logger.error("Detector not loaded (coordinator mode should not call _run_inference)")
            return []
        self.frame_counter += 1
        if self.sparse_detection_enabled and self.frame_counter % self.keyframe_interval != 0:
            return self.last_detections.copy() if self.last_detections else []  # shallow, like the reference (:586)
        if self.empty_frame_filter and not self.empty_frame_filter.has_motion(frame):
            self.last_detections = []
            return self.last_detections
        pressure = self.memory_manager.check_memory_pressure()
        if pressure in (MemoryPressure.CRITICAL, MemoryPressure.EXTREME):
            self._apply_degradation(self.memory_manager.reduce_memory_usage(pressure))
        try:
            return self._process_and_store_detections(frame)
        except Exception as e:
            if not _is_cuda_oom(e):
                logger.error("Unexpected error during inference: %s", e, exc_info=True)
                return []
            logger.error("GPU OOM during inference for camera %s: %s", self.camera_id, e)
            rec = self.memory_manager.handle_oom_error()
            self._apply_degradation(rec)
            if "wait_time" in rec:
                time.sleep(rec["wait_time"])
            try:
                self.memory_manager.record_recovery()
                return self._process_and_store_detections(frame)
            except Exception as retry_e:
                logger.error("Inference failed after OOM recovery: %s", retry_e)
                return []

    def infer_single(self, frame) -> List[Dict[str, Any]]:
        if not self.is_loaded:
            logger.error("Model not loaded")
            return []
        if self.shared_coordinator:
            logger.error("infer_single() not supported in coordinator mode - use async inference")
            return []
        return self._run_inference(frame)

    # ------------------------------------------------------------------ hot reload / degradation
    def update_settings(self, conf_threshold: Optional[float] = None, nms_threshold: Optional[float] = None,
                        min_box_area: Optional[int] = None, max_det: Optional[int] = None,
                        class_confidence_overrides: Optional[Dict[str, float]] = None,
                        class_size_constraints: Optional[Dict[str, Dict[str, int]]] = None):
        with self.settings_lock:
            changed = []
            for key, val in (("conf_threshold", conf_threshold), ("nms_threshold", nms_threshold)):
                if val is not None and val != getattr(self, key):
                    setattr(self, key, val)
                    if self.detector:
                        setattr(self.detector, key, val)  # the detector pushes it to the device before its next launch
                    changed.append("%s: %s" % (key, val))
            for key, val in (("min_box_area", min_box_area), ("max_det", max_det)):
                if val is not None and val != getattr(self, key):
                    setattr(self, key, val)
                    changed.append("%s: %s" % (key, val))
            if class_confidence_overrides is not None:
                self.class_confidence_overrides = class_confidence_overrides
                changed.append("class_confidence_overrides: %s" % class_confidence_overrides)
            if class_size_constraints is not None:
                self.class_size_constraints = class_size_constraints
                changed.append("class_size_constraints: %s" % class_size_constraints)
            if changed:
                logger.info("Updated inference settings for %s: %s", self.camera_id, ", ".join(changed))
        if changed and self._filter_id is not None:
            self._device_filter(refresh=True)   # outside the lock: it takes its own snapshot

    def _apply_degradation(self, recommendations: Dict[str, Any]):
        if not recommendations:
            return
        if recommendations.get("clear_cache", False):
            self.memory_manager.clear_cache()
        if recommendations.get("reduce_input_size", False):
            size = recommendations.get("suggested_input_size")
            if size and size != self.input_size:
                logger.warning("Reducing input size: %s -> %s", self.input_size, size)
                self.input_size = size
                if self.detector:
                    self.detector.input_size = size
                    if getattr(self.detector, "exp", None) is not None:
                        self.detector.exp.test_size = size
                self.degradation_active = True
        if recommendations.get("cpu_fallback", False):
            logger.error("cpu_fallback recommended but the B200 build has no CPU path; staying on %s", self.device)

    def get_stats(self) -> Dict[str, Any]:
        stats = {"device": self.device, "fps": self.fps, "avg_inference_time": self.avg_inference_time,
                 "avg_inference_time_ms": self.avg_inference_time * 1000,
                 "total_inferences": self.total_inference_count, "dropped_results": self.dropped_results,
                 "coordinator_mode": self.shared_coordinator is not None,
                 "degradation_active": self.degradation_active,
                 "drop_rate": self.dropped_results / self.total_inference_count if self.total_inference_count else 0.0}
        if self.shared_coordinator:
            stats["coordinator"] = self.shared_coordinator.get_stats()
        mem = self.memory_manager.get_memory_stats()
        if "allocated_mb" in mem:
            stats["gpu_memory_allocated_mb"] = mem["allocated_mb"]
            stats["gpu_memory_reserved_mb"] = mem["reserved_mb"]
        stats["memory"] = mem
        if self.empty_frame_filter:
            stats["empty_frame_filter"] = self.empty_frame_filter.get_stats()
        if self.sparse_detection_enabled:
            stats["sparse_detection"] = {"enabled": True, "keyframe_interval": self.keyframe_interval,
                                         "mode": self.sparse_mode, "frame_counter": self.frame_counter,
                                         "keyframes_processed": self.frame_counter // self.keyframe_interval}
        try:
            import psutil
            proc = psutil.Process()
            stats["memory_mb"] = proc.memory_info().rss / 1024 / 1024
            stats["memory_percent"] = proc.memory_percent()
        except ImportError:
            pass
        return stats
