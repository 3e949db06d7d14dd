"""JPEG encode of frames on the result path.

`imencode('.jpg', frame, [cv2.IMWRITE_JPEG_QUALITY, q])` -> `(True, np.ndarray[uint8])` with cv2's signature, for the call
sites /root/reference/src/web_server.py:923 (MJPEG stream: one encode per camera per tick) and
src/snapshot_saver.py:160 (ring-buffer compression); `imwrite(path, frame)` for snapshot_saver.py:358-404.  Frames may
be NumPy arrays or CUDA uint8 tensors (the ingest ring's device frames never have to come back to the host as pixels);
`JpegEncoder.encode_batch` takes a list (all cameras of a tick in one call, five kernel launches for the batch).

The file is byte-identical to what cv2 (libjpeg-turbo) writes for the same frame and quality: baseline sequential JPEG,
YCbCr 4:2:0, Annex K tables, JFIF 1.01 header (csrc/jpeg.cu; oracle/jpeg.py; tests/test_jpeg.py).  Only what the
reference uses is built: 3-channel BGR uint8, the quality parameter.  Anything else raises.
"""
import ctypes as C
import threading
from typing import List, Optional, Sequence

import numpy as np

from . import _lib
from .engine import _is_torch, device_index

IMWRITE_JPEG_QUALITY = 1     # cv2's constant


class JpegEncoder:
    def __init__(self, max_hw=(1080, 1920), max_batch: int = 8, device="cuda:0"):
        self.lib = _lib.load()
        self.max_hw = (int(max_hw[0]), int(max_hw[1]))
        self.max_batch = int(max_batch)
        self.device = device_index(device)
        h = C.c_void_p()
        _lib.check(self.lib.tcd_jpeg_encoder_create(self.max_hw[0], self.max_hw[1], self.max_batch, self.device, C.byref(h)),
                   "tcd_jpeg_encoder_create")
        self.h = h
        self._out = None

    def close(self):
        if getattr(self, "h", None):
            self.lib.tcd_jpeg_encoder_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    def encode_batch(self, frames: Sequence, quality: int = 95) -> List[np.ndarray]:
        """-> one uint8 array (the .jpg file) per frame."""
        n = len(frames)
        if n == 0:
            return []
        if n > self.max_batch:
            out = []
            for i in range(0, n, self.max_batch):
                out.extend(self.encode_batch(frames[i:i + self.max_batch], quality))
            return out
        on_dev = _is_torch(frames[0]) and frames[0].is_cuda
        keep, ptrs, hs, ws, strides = [], [], [], [], []
        for f in frames:
            if _is_torch(f):
                import torch
                if f.dtype != torch.uint8 or f.ndim != 3 or f.shape[2] != 3:
                    raise TypeError("frames must be HxWx3 uint8")
                if f.is_cuda != on_dev:
                    raise TypeError("encode_batch: frames must be all on the host or all on the device")
                if f.stride(2) != 1 or f.stride(1) != 3:
                    f = f.contiguous()
                if not on_dev:
                    f = f.numpy()
                else:
                    keep.append(f)
                    ptrs.append(f.data_ptr()); hs.append(f.shape[0]); ws.append(f.shape[1]); strides.append(f.stride(0))
                    continue
            a = np.asarray(f)
            if on_dev:
                raise TypeError("encode_batch: frames must be all on the host or all on the device")
            if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
                raise TypeError("frames must be HxWx3 uint8 (got %s %s)" % (a.dtype, a.shape))
            if a.strides[2] != 1 or a.strides[1] != 3:
                a = np.ascontiguousarray(a)
            keep.append(a)
            ptrs.append(a.ctypes.data); hs.append(a.shape[0]); ws.append(a.shape[1]); strides.append(a.strides[0])
        if on_dev:
            import torch
            torch.cuda.current_stream(keep[0].device).synchronize()   # the frames were produced on the caller's stream
        caps = [int(self.lib.tcd_jpeg_max_bytes(h, w)) for h, w in zip(hs, ws)]
        bufs = [np.empty(c, np.uint8) for c in caps]
        sizes = (C.c_int64 * n)()
        _lib.check(self.lib.tcd_jpeg_encode(self.h, _lib.ptr_array(ptrs), _lib.i32_array(hs), _lib.i32_array(ws),
                                            _lib.i64_array(strides), n, 1 if on_dev else 0, int(quality),
                                            _lib.ptr_array([b.ctypes.data for b in bufs]), _lib.i64_array(caps), sizes),
                   "tcd_jpeg_encode")
        return [b[:sizes[i]].copy() for i, b in enumerate(bufs)]


_default: Optional[JpegEncoder] = None
_default_lock = threading.Lock()


def _encoder_for(h, w, device):
    global _default
    with _default_lock:
        e = _default
        if e is None or e.max_hw[0] < h or e.max_hw[1] < w or e.device != device_index(device):
            if e is not None:
                e.close()
            e = _default = JpegEncoder((max(h, 1080), max(w, 1920)), 8, device)
        return e


def _quality(params) -> int:
    q = 95
    if params:
        p = list(params)
        if len(p) % 2:
            raise ValueError("params must be (id, value) pairs")
        for k, v in zip(p[0::2], p[1::2]):
            if int(k) == IMWRITE_JPEG_QUALITY:
                q = int(v)
            else:
                raise NotImplementedError("imencode: only IMWRITE_JPEG_QUALITY is built (got parameter id %d)" % int(k))
    return min(max(q, 0), 100)


def imencode(ext: str, img, params=None, device="cuda:0"):
    """cv2.imencode for '.jpg' / '.jpeg' and 3-channel uint8 frames."""
    if ext.lower() not in (".jpg", ".jpeg"):
        raise NotImplementedError("imencode: only JPEG is built (got %r)" % ext)
    if _is_torch(img) and img.is_cuda:
        device = img.device.index or 0
    out = _encoder_for(int(img.shape[0]), int(img.shape[1]), device).encode_batch([img], _quality(params))[0]
    return True, out


def imwrite(path: str, img, params=None, device="cuda:0") -> bool:
    """cv2.imwrite for .jpg paths (snapshot_saver.py:358-404): False on an I/O failure, like cv2."""
    ok, buf = imencode(".jpg" if str(path).lower().endswith((".jpg", ".jpeg")) else str(path)[str(path).rfind("."):], img, params,
                       device)
    try:
        with open(path, "wb") as f:
            f.write(buf.tobytes())
        return True
    except OSError:
        return False


def prepare_detection_message(detection_result: dict) -> dict:
    """The WebSocket message of one result (reference src/web_server.py:849-873): same keys, same defaults."""
    return {
        "type": "detections",
        "camera_id": detection_result.get("camera_id", "default"),
        "camera_name": detection_result.get("camera_name", "Default Camera"),
        "frame_id": detection_result.get("frame_id"),
        "timestamp": detection_result.get("timestamp"),
        "processing_timestamp": detection_result.get("processing_timestamp"),
        "latency_ms": detection_result.get("total_latency_ms", 0),
        "inference_time_ms": detection_result.get("inference_time_ms", 0),
        "detections": detection_result.get("detections", []),
        "detection_counts": detection_result.get("detection_counts", {}),
        "total_detections": detection_result.get("total_detections", 0),
    }


def mjpeg_part(jpeg_bytes) -> bytes:
    """One part of the multipart/x-mixed-replace stream (reference src/web_server.py:926-929)."""
    return b"--frame\r\nContent-Type: image/jpeg\r\n\r\n" + bytes(jpeg_bytes) + b"\r\n"
