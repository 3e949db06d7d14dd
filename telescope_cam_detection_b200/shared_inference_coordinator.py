"""SharedInferenceCoordinator -- drop-in for the reference's src/shared_inference_coordinator.py.

Same call surface and semantics (:27-338): `infer_async(frame, callback, camera_id)` queues a
frame; one coordinator thread forms batches of up to `max_batch_size`, waiting at most
`max_batch_wait_ms` for stragglers; at `max_queue_depth` the OLDEST request is dropped and its
callback gets `[]`; a failing batch gives every request in it `callback([])`; callbacks fire
on the coordinator thread in request order; `get_stats()` has the same keys.

What changes underneath: when the detector offers the split-phase `submit_batch/finish_batch`
pair (YOLOXDetector of this package), batches are submitted to every pipeline slot the library
has (`detector.num_slots`: engines x lanes) before the oldest one is read back, so batch k+1..k+n
are collected, uploaded and computing while Python formats the results of batch k, instead of the
GPU idling behind the callbacks.  Results are delivered in submission order.  No allocator hygiene calls: libtcd works out of a static arena, so the
reference's `empty_cache()` every 50 batches (:273-276) has nothing to do.
"""
import logging
import threading
import time
from collections import deque
from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional

logger = logging.getLogger(__name__)


@dataclass
class PendingInference:
    frame: Any
    callback: Callable[[List[Dict[str, Any]]], None]
    enqueue_time: float
    camera_id: Optional[str] = None
    filter_id: Optional[int] = None   # addition: device-side class-filter table of the requesting camera


class SharedInferenceCoordinator:
    def __init__(self, detector: Any, max_batch_size: int = 4, max_batch_wait_ms: float = 10.0,
                 enable_metrics: bool = True, max_queue_depth: int = 60):
        self.detector = detector
        self.max_batch_size = max_batch_size
        self.max_batch_wait_ms = max_batch_wait_ms / 1000.0  # seconds, like the reference attribute
        self.enable_metrics = enable_metrics
        self.max_queue_depth = max_queue_depth
        self.pending_queue: deque = deque()
        self.queue_lock = threading.Lock()
        self.queue_condition = threading.Condition(self.queue_lock)
        self.coordinator_thread: Optional[threading.Thread] = None
        self.stop_event = threading.Event()
        self.running = False
        self.total_batches = 0
        self.total_frames = 0
        self.total_batch_time_ms = 0.0
        self.dropped_frames = 0
        self.batch_sizes: deque = deque(maxlen=1000)
        self.wait_times_ms: deque = deque(maxlen=1000)
        self._pipelined = hasattr(detector, "submit_batch") and hasattr(detector, "finish_batch")
        logger.info("SharedInferenceCoordinator initialized (max_batch_size=%d, max_wait=%.1fms, max_queue_depth=%d)",
                    max_batch_size, max_batch_wait_ms, max_queue_depth)

    # ------------------------------------------------------------------ lifecycle
    def start(self):
        if self.running:
            logger.warning("Coordinator already running")
            return
        self.running = True
        self.stop_event.clear()
        self.coordinator_thread = threading.Thread(target=self._coordinator_loop, name="InferenceCoordinator",
                                                   daemon=True)
        self.coordinator_thread.start()

    def stop(self):
        if not self.running:
            return
        self.running = False
        self.stop_event.set()
        with self.queue_condition:
            self.queue_condition.notify()
        if self.coordinator_thread:
            self.coordinator_thread.join(timeout=2.0)
            if self.coordinator_thread.is_alive():
                logger.warning("Coordinator thread did not stop cleanly")

    def __enter__(self):
        self.start()
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.stop()

    # ------------------------------------------------------------------ producer side
    supports_filter_ids = True

    def infer_async(self, frame: Any, callback: Callable[[List[Dict[str, Any]]], None],
                    camera_id: Optional[str] = None, filter_id: Optional[int] = None):
        if not self.running:
            raise RuntimeError("Coordinator not running - call start() first")
        req = PendingInference(frame=frame, callback=callback, enqueue_time=time.time(), camera_id=camera_id,
                               filter_id=filter_id)
        with self.queue_condition:
            if len(self.pending_queue) >= self.max_queue_depth:
                victim = self.pending_queue.popleft()  # drop the oldest, tell its owner (reference :151-164)
                self.dropped_frames += 1
                if self.dropped_frames % 10 == 0:
                    logger.warning("Inference queue full (%d) - dropped %d frames total", self.max_queue_depth,
                                   self.dropped_frames)
                try:
                    victim.callback([])
                except Exception:
                    pass
            self.pending_queue.append(req)
            self.queue_condition.notify()

    # ------------------------------------------------------------------ coordinator thread
    def _collect_batch(self, block: bool = True) -> List[PendingInference]:
        batch: List[PendingInference] = []
        with self.queue_condition:
            while block and not self.pending_queue and not self.stop_event.is_set():
                self.queue_condition.wait(timeout=0.1)
            if self.stop_event.is_set() or not self.pending_queue:
                return []
            t0 = time.time()
            while len(batch) < self.max_batch_size and self.pending_queue:
                batch.append(self.pending_queue.popleft())
                if len(batch) < self.max_batch_size and not self.pending_queue:
                    remaining = self.max_batch_wait_ms - (time.time() - t0)
                    if remaining <= 0:
                        break
                    self.queue_condition.wait(timeout=remaining)
        return batch

    def _note_waits(self, batch, now):
        if self.enable_metrics:
            for req in batch:
                self.wait_times_ms.append((now - req.enqueue_time) * 1000)

    def _deliver(self, batch, all_detections, elapsed_ms):
        for req, dets in zip(batch, all_detections):
            try:
                req.callback(dets)
            except Exception as e:
                logger.error("Error in callback for camera %s: %s", req.camera_id, e)
        if self.enable_metrics:
            self.total_batches += 1
            self.total_frames += len(batch)
            self.total_batch_time_ms += elapsed_ms
            self.batch_sizes.append(len(batch))
            if self.total_batches % 100 == 0:
                self._log_metrics()

    def _fail(self, batch, err):
        logger.error("Error processing batch: %s", err, exc_info=True)
        for req in batch:
            try:
                req.callback([])
            except Exception as cb_error:
                logger.error("Error calling callback on error: %s", cb_error)

    def _process_batch(self, batch: List[PendingInference]):
        """Synchronous form (any detector with detect_batch), reference :226-288."""
        if not batch:
            return
        t0 = time.time()
        self._note_waits(batch, t0)
        try:
            dets = self.detector.detect_batch([r.frame for r in batch])
            self._deliver(batch, dets, (time.time() - t0) * 1000)
        except Exception as e:
            self._fail(batch, e)

    def _coordinator_loop(self):
        if not self._pipelined:
            while not self.stop_event.is_set():
                try:
                    batch = self._collect_batch()
                    if batch:
                        self._process_batch(batch)
                except Exception as e:
                    logger.error("Error in coordinator loop: %s", e, exc_info=True)
            return
        n_slots = max(1, int(getattr(self.detector, "num_slots", 2) or 2))
        inflight: deque = deque()        # (slot, batch, t_submit), oldest first
        free = list(range(n_slots))
        while not self.stop_event.is_set():
            try:
                batch = []
                if free:
                    # with batches on the GPU, do not block: read the oldest back as soon as nothing is queued
                    batch = self._collect_batch(block=not inflight)
                if batch:
                    t0 = time.time()
                    self._note_waits(batch, t0)
                    slot = free.pop(0)
                    try:
                        fids = [r.filter_id for r in batch]
                        if any(f is not None for f in fids):
                            self.detector.submit_batch(slot, [r.frame for r in batch], fids)
                        else:
                            self.detector.submit_batch(slot, [r.frame for r in batch])
                        inflight.append((slot, batch, t0))
                    except Exception as e:
                        free.append(slot)
                        self._fail(batch, e)
                    if free:
                        with self.queue_lock:
                            more = bool(self.pending_queue)
                        if more:
                            continue     # keep filling slots while requests are waiting
                if inflight:
                    item = inflight.popleft()
                    self._finish(item)
                    free.append(item[0])
            except Exception as e:
                logger.error("Error in coordinator loop: %s", e, exc_info=True)
        while inflight:
            self._finish(inflight.popleft())

    def _finish(self, inflight):
        s, batch, t0 = inflight
        try:
            dets = self.detector.finish_batch(s)
            self._deliver(batch, dets, (time.time() - t0) * 1000)
        except Exception as e:
            self._fail(batch, e)

    # ------------------------------------------------------------------ metrics
    def _averages(self):
        n = max(len(self.batch_sizes), 1)
        avg_size = sum(self.batch_sizes) / n
        avg_time = self.total_batch_time_ms / max(self.total_batches, 1)
        avg_wait = sum(self.wait_times_ms) / len(self.wait_times_ms) if self.wait_times_ms else 0
        fps = self.total_frames / (self.total_batch_time_ms / 1000) if self.total_batch_time_ms > 0 else 0
        return avg_size, avg_time, avg_wait, fps

    def _log_metrics(self):
        if not self.enable_metrics or self.total_batches == 0:
            return
        s, t, w, fps = self._averages()
        logger.info("Coordinator metrics after %d batches: avg batch %.1f frames, %.1f ms, wait %.1f ms, %d frames, "
                    "%.1f fps", self.total_batches, s, t, w, self.total_frames, fps)

    def get_stats(self) -> Dict[str, Any]:
        if not self.enable_metrics or self.total_batches == 0:
            return {"enabled": False, "total_batches": 0, "total_frames": 0}
        s, t, w, fps = self._averages()
        return {"enabled": True, "total_batches": self.total_batches, "total_frames": self.total_frames,
                "avg_batch_size": round(s, 2), "avg_batch_time_ms": round(t, 2), "avg_wait_time_ms": round(w, 2),
                "throughput_fps": round(fps, 1), "queue_depth": len(self.pending_queue)}
