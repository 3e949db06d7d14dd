"""RTSPStreamCaptureGPU: the ingest end of the per-frame path.

Same class surface as /root/reference/src/stream_capture_gpu_ffmpeg.py:19-387 -- constructor arguments, `connect`,
`start`, `stop`, `get_stats`, `get_latest_frame_as_numpy`, and the queue entries
`{'frame', 'timestamp', 'frame_id', 'camera_id', 'camera_name', 'is_gpu_tensor'}` -- over a frame ring in libtcd
(include/tcd.h, tcd_frame_ring_*) instead of the reference's per-frame copy chain (:253-283):

    reference, per frame                          here, per frame
    ---------------------------------------       -----------------------------------------------
    stdout.read(frame_size) -> bytes object       read(fd) straight into a PINNED ring slot (C, no GIL)
    np.frombuffer + torch.from_numpy(..).to()     one cudaMemcpyAsync pinned -> device slot, copy stream
      (pageable: staged through driver memory)
    img.clone() for latest_frame                  latest_frame IS the ring slot
    img.clone() for the queue entry               the queue entry IS the ring slot

Slot ownership replaces the two clones: a queued frame's slot is recycled only after `maxsize + 2` NEWER frames have
been queued (the queue can hold `maxsize` of them and a consumer two more), a frame the full queue refused stays
readable as `latest_frame` until the next frame has arrived, and the slot being read into is never one of those.  The decoder is the reference's own command line (ffmpeg -hwaccel cuda -c:v h264_cuvid
... -f rawvideo -pix_fmt bgr24 pipe:1); `command=` substitutes any program that writes raw bgr24 frames to stdout (the
tests and bench.py use a synthetic writer: there is no ffmpeg binary on this image).  With `keep_frames_on_gpu=False`
the entries are NumPy views of the pinned slots and no GPU is touched when `device` is None.
"""
import ctypes as C
import logging
import subprocess
import time
from queue import Full, Queue
from threading import Event, Lock, Thread
from typing import Optional

import numpy as np

from . import _lib
from .engine import device_index

logger = logging.getLogger(__name__)


class _DeviceSlot:
    """Zero-copy handle of one device ring slot for torch.as_tensor / cupy (CUDA array interface v2)."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class FrameRing:
    """n pinned host slots + n device slots of one frame each (tcd_frame_ring)."""

    def __init__(self, n_slots, frame_shape, device=0):
        self.lib = _lib.load()
        self.shape = tuple(int(x) for x in frame_shape)
        self.frame_bytes = int(np.prod(self.shape))
        self.n = int(n_slots)
        self.device = device_index(device)
        h = C.c_void_p()
        _lib.check(self.lib.tcd_frame_ring_create(self.n, self.frame_bytes, self.device, C.byref(h)), "tcd_frame_ring_create")
        self.h = h
        self._host = [np.ctypeslib.as_array(C.cast(self.lib.tcd_frame_ring_host(self.h, i), C.POINTER(C.c_uint8)),
                                            shape=(self.frame_bytes,)).reshape(self.shape) for i in range(self.n)]
        self._dev = None

    def host(self, slot):
        return self._host[slot]

    def device_ptr(self, slot):
        return self.lib.tcd_frame_ring_device(self.h, slot)

    def device_tensor(self, slot):
        """torch uint8 view of the device slot (no copy)."""
        import torch
        if self._dev is None:
            self._dev = [torch.as_tensor(_DeviceSlot(self.device_ptr(i), self.shape), device="cuda:%d" % self.device)
                         for i in range(self.n)]
        return self._dev[slot]

    def read_fd(self, slot, fd):
        got = C.c_int64(0)
        _lib.check(self.lib.tcd_frame_ring_read_fd(self.h, slot, int(fd), C.byref(got)), "tcd_frame_ring_read_fd")
        return got.value

    def publish(self, slot):
        _lib.check(self.lib.tcd_frame_ring_publish(self.h, slot), "tcd_frame_ring_publish")

    def wait(self, slot, stream=None):
        _lib.check(self.lib.tcd_frame_ring_wait(self.h, slot, C.c_void_p(int(stream)) if stream else None), "tcd_frame_ring_wait")

    def close(self):
        if getattr(self, "h", None):
            self._dev = None
            self._host = []
            self.lib.tcd_frame_ring_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001 -- interpreter shutdown
            pass


class _HostRing:
    """keep_frames_on_gpu=False without a device: plain host slots, same slot discipline."""

    def __init__(self, n_slots, frame_shape):
        self.n = int(n_slots)
        self.shape = tuple(frame_shape)
        self.frame_bytes = int(np.prod(self.shape))
        self._host = [np.empty(self.shape, np.uint8) for _ in range(self.n)]

    def host(self, slot):
        return self._host[slot]

    def read_fd(self, slot, fd):
        import os
        view = memoryview(self._host[slot]).cast("B")
        have = 0
        while have < self.frame_bytes:
            k = os.readv(fd, [view[have:]])
            if k == 0:
                break
            have += k
        return have

    def close(self):
        self._host = []


class RTSPStreamCaptureGPU:
    """Drop-in for the reference class of the same name (stream_capture_gpu_ffmpeg.py:19)."""

    def __init__(self, rtsp_url, frame_queue: Queue, target_width=1280, target_height=720, camera_id="default",
                 camera_name="Default Camera", use_tcp=False, buffer_size=None, max_failures=30, retry_delay=5.0,
                 keep_frames_on_gpu=True, device="cuda:0", ring_slots=None, command=None):
        self.rtsp_url = rtsp_url
        self.frame_queue = frame_queue
        self.target_width = target_width
        self.target_height = target_height
        self.camera_id = camera_id
        self.camera_name = camera_name
        self.use_tcp = use_tcp
        self.buffer_size = buffer_size
        self.max_failures = max_failures
        self.retry_delay = retry_delay
        self.keep_frames_on_gpu = keep_frames_on_gpu
        self.device = device
        self.command = list(command) if command else None
        if buffer_size is not None:   # same notice as the reference (:72-76)
            logger.warning("[%s] buffer_size parameter is not used by FFmpeg GPU capture (FFmpeg uses internal "
                           "buffering). Ignoring buffer_size=%s", camera_id, buffer_size)
        qmax = getattr(frame_queue, "maxsize", 0) or 0
        self.ring_slots = max(6, int(ring_slots)) if ring_slots else max(8, qmax + 5)
        self._hold = self.ring_slots - 3          # queued slots kept before the oldest is recycled
        if qmax <= 0 or self._hold < qmax + 2:
            logger.warning("[%s] frame queue maxsize %d with a ring of %d slots: a queued frame is valid until %d newer "
                           "frames have been queued", camera_id, qmax, self.ring_slots, self._hold)
        self.ring = None
        self.ffmpeg_process: Optional[subprocess.Popen] = None
        self.stop_event = Event()
        self.capture_thread: Optional[Thread] = None
        self.is_connected = False
        self.latest_frame = None
        self.frame_lock = Lock()
        self.frame_count = 0
        self.frame_id_counter = 0
        self.dropped_frames = 0
        self.last_fps_check = time.time()
        self.fps = 0.0

    # -- same command line as the reference (:94-121)
    def _build_ffmpeg_command(self) -> list:
        if self.command is not None:
            return list(self.command)
        cmd = ["ffmpeg", "-hwaccel", "cuda", "-c:v", "h264_cuvid"]
        if self.use_tcp:
            cmd.extend(["-rtsp_transport", "tcp"])
        cmd.extend(["-i", self.rtsp_url, "-f", "rawvideo", "-pix_fmt", "bgr24",
                    "-s", "%dx%d" % (self.target_width, self.target_height), "pipe:1"])
        return cmd

    def _ensure_ring(self):
        if self.ring is not None:
            return
        shape = (self.target_height, self.target_width, 3)
        if self.keep_frames_on_gpu:
            self.ring = FrameRing(self.ring_slots, shape, self.device)
            return
        # host frames: pinned slots when there is a GPU to upload them to later, plain host memory otherwise (this
        # branch moves bytes only; nothing is computed here)
        try:
            self.ring = FrameRing(self.ring_slots, shape, self.device) if self.device is not None else None
        except Exception as e:  # noqa: BLE001
            logger.info("[%s] no pinned ring (%s): plain host slots", self.camera_id, e)
            self.ring = None
        if self.ring is None:
            self.ring = _HostRing(self.ring_slots, shape)

    def connect(self) -> bool:
        logger.info("[%s] Connecting to RTSP stream (GPU decode via FFmpeg): %s", self.camera_id, self.rtsp_url)
        try:
            self._ensure_ring()
            cmd = self._build_ffmpeg_command()
            # bufsize=0: stdout is the raw pipe; the ring reads its fd directly, a Python-side buffer would be a copy
            self.ffmpeg_process = subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, bufsize=0)
            try:   # a 6 MB frame through the default 64 KiB pipe is ~100 wake-ups; 1 MiB is the unprivileged maximum
                import fcntl
                fcntl.fcntl(self.ffmpeg_process.stdout.fileno(), 1031, 1 << 20)   # F_SETPIPE_SZ
            except (ImportError, OSError):
                pass
            time.sleep(0.5 if self.command is None else 0.0)
            if self.command is None and self.ffmpeg_process.poll() is not None:
                stderr = self.ffmpeg_process.stderr.read().decode("utf-8", errors="ignore")
                logger.error("[%s] FFmpeg failed to start: %s", self.camera_id, stderr[:500])
                return False
            self.is_connected = True
            return True
        except Exception as e:  # noqa: BLE001 -- as the reference: any failure is "not connected"
            logger.error("[%s] Connection error: %s", self.camera_id, e)
            return False

    def start(self) -> bool:
        self.stop_event.clear()
        self.capture_thread = Thread(target=self._capture_loop, daemon=True)
        self.capture_thread.start()
        return True

    def _kill_process(self, wait_s=5.0) -> bool:
        p = self.ffmpeg_process
        if p is None:
            return True
        try:
            p.terminate()
            p.wait(timeout=wait_s)
        except subprocess.TimeoutExpired:
            try:
                p.kill()
                p.wait(timeout=2.0)
            except subprocess.TimeoutExpired:
                logger.error("[%s] CRITICAL: FFmpeg process did not die after kill()", self.camera_id)
                return False
        except Exception as e:  # noqa: BLE001
            logger.error("[%s] Error terminating FFmpeg: %s", self.camera_id, e)
            return False
        for f in (p.stdout, p.stderr):
            try:
                f and f.close()
            except Exception:  # noqa: BLE001
                pass
        self.ffmpeg_process = None
        return True

    def stop(self):
        self.stop_event.set()
        # the capture thread may sit in read(): ending the decoder ends the read
        self._kill_process()
        if self.capture_thread:
            self.capture_thread.join(timeout=5.0)
            if self.capture_thread.is_alive():
                logger.error("[%s] CRITICAL: Capture thread did not stop after 5s timeout", self.camera_id)
        self.is_connected = False

    def _capture_loop(self):
        from collections import deque
        consecutive_failures = 0
        frame_size = self.target_height * self.target_width * 3
        free = deque(range(self.ring_slots))
        held = deque()                      # slots of queued frames, oldest first
        latest_slot, latest_queued = -1, True
        slot = free.popleft()
        while not self.stop_event.is_set():
            try:
                if not self.is_connected or self.ffmpeg_process is None:
                    if self.connect():
                        consecutive_failures = 0
                    else:
                        if self.stop_event.wait(self.retry_delay):
                            break
                        continue
                ring = self.ring
                got = ring.read_fd(slot, self.ffmpeg_process.stdout.fileno())
                if got != frame_size:
                    if self.stop_event.is_set():
                        break
                    consecutive_failures += 1
                    logger.warning("[%s] Failed to read frame (%d/%d)", self.camera_id, consecutive_failures, self.max_failures)
                    if consecutive_failures >= self.max_failures:
                        self._reconnect()
                        consecutive_failures = 0
                    elif got == 0:
                        time.sleep(0.01)   # end of stream: the reference spins on empty reads; do not burn a core
                    continue
                if self.keep_frames_on_gpu:
                    ring.publish(slot)
                    ring.wait(slot)            # ~0.25 ms for 1080p: the entry handed out is complete
                    img = ring.device_tensor(slot)
                else:
                    img = ring.host(slot)
                with self.frame_lock:
                    self.latest_frame = img
                timestamp = time.time()
                try:
                    self.frame_queue.put_nowait({
                        "frame": img, "timestamp": timestamp, "frame_id": self.frame_id_counter,
                        "camera_id": self.camera_id, "camera_name": self.camera_name,
                        "is_gpu_tensor": self.keep_frames_on_gpu})
                    self.frame_count += 1
                    self.frame_id_counter += 1
                    queued = True
                except Full:
                    self.dropped_frames += 1
                    queued = False
                # slot ownership (module docstring): the previous latest-only frame is free now; a queued slot is
                # recycled once `_hold` newer ones are queued
                if not latest_queued and latest_slot >= 0:
                    free.append(latest_slot)
                latest_slot, latest_queued = slot, queued
                if queued:
                    held.append(slot)
                    if len(held) > self._hold:
                        free.append(held.popleft())
                slot = free.popleft()
                now = time.time()
                if now - self.last_fps_check >= 1.0:
                    self.fps = self.frame_count / (now - self.last_fps_check)
                    self.last_fps_check = now
                    self.frame_count = 0
                    if self.dropped_frames > 0:
                        logger.debug("[%s] FPS: %.1f, Dropped: %d", self.camera_id, self.fps, self.dropped_frames)
                        self.dropped_frames = 0
                consecutive_failures = 0
            except Exception as e:  # noqa: BLE001 -- as the reference (:313-322)
                if self.stop_event.is_set():
                    break
                consecutive_failures += 1
                logger.warning("[%s] Stream error (%d/%d): %s", self.camera_id, consecutive_failures, self.max_failures, e)
                if consecutive_failures >= self.max_failures:
                    self._reconnect()
                    consecutive_failures = 0
                if self.stop_event.wait(1.0):
                    break

    def _reconnect(self) -> bool:
        if not self._kill_process(wait_s=2.0):
            self.is_connected = False
            return False
        self.is_connected = False
        if self.stop_event.wait(2.0 if self.command is None else 0.05):
            return False
        return self.connect()

    def get_stats(self) -> dict:
        return {"is_connected": self.is_connected, "fps": self.fps, "total_frames": self.frame_count,
                "dropped_frames": self.dropped_frames, "queue_size": self.frame_queue.qsize()}

    def get_latest_frame_as_numpy(self) -> Optional[np.ndarray]:
        with self.frame_lock:
            f = self.latest_frame
            if f is None:
                return None
            if isinstance(f, np.ndarray):
                return f
            return f.cpu().numpy()
