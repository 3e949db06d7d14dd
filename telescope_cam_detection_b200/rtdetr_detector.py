"""RTDETRDetector -- drop-in for the reference's src/rtdetr_detector.py (:22-403).

Same constructor, attributes and methods; `detect` / `detect_batch` return the same detection dicts
(score >= conf_threshold, wildlife filter, `area = int(...)`, src/rtdetr_detector.py:267-303,359-388).
Underneath: frames go to libtcd as u8, the PIL-antialias resize (`tcd_pil_resize`), RT-DETRv2-R50 and
the top-300 postprocessor run as hand-written sm_100a kernels inside one CUDA graph per batch size
(rtdetr_engine.py).  No CPU / PyTorch fallback: without libtcd.so or an sm_100 GPU `load_model`
returns False (the reference's convention, :60-204) and `detect*` return empty lists.

Weights: a state dict with the `transformers` RT-DETRv2 key names, either passed in memory
(`state_dict=`) or stored at `model_path` as .npz / torch .pt|.pth.  The reference's YAML
`config_path` is accepted and ignored (depth / widths are read from the weights).
"""
import ctypes as C
import logging
import os
from typing import Any, Dict, List, Optional, Union

import numpy as np

from .coco_constants import COCO_CLASSES, MAMMAL_CLASS_IDS, WILDLIFE_CLASSES

logger = logging.getLogger(__name__)

_BATCH_SIZES = (1, 2, 4, 8, 16, 32)


class _ModelHandle:
    """Stands where the reference keeps an nn.Module (callers test it for None)."""

    def __init__(self, engines):
        self.engines = engines

    def to(self, device):
        if not str(device).startswith("cuda"):
            raise RuntimeError("the B200 build has no CPU path (requested device %r)" % str(device))
        return self

    def eval(self):
        return self


def _frame_desc(frame):
    """-> (ptr, h, w, row_stride, on_device, keepalive) for an HxWx3 uint8 numpy array or torch tensor."""
    if type(frame).__module__.startswith("torch"):
        t = frame
        if t.ndim != 3 or t.shape[2] != 3 or str(t.dtype) != "torch.uint8":
            raise ValueError("frame tensor must be HxWx3 uint8, got %s %s" % (tuple(t.shape), t.dtype))
        if t.stride(2) != 1 or t.stride(1) != 3:
            t = t.contiguous()
        return t.data_ptr(), int(t.shape[0]), int(t.shape[1]), int(t.stride(0)), bool(t.is_cuda), t
    a = np.asarray(frame)
    if a.ndim != 3 or a.shape[2] != 3 or a.dtype != np.uint8:
        raise ValueError("frame must be HxWx3 uint8, got %s %s" % (a.shape, a.dtype))
    if a.strides[2] != 1 or a.strides[1] != 3:
        a = np.ascontiguousarray(a)
    return a.ctypes.data, int(a.shape[0]), int(a.shape[1]), int(a.strides[0]), False, a


class RTDETRDetector:
    def __init__(self, config_path: str = "RT-DETR/rtdetrv2_pytorch/configs/rtdetrv2/rtdetrv2_r18vd_120e_coco.yml",
                 model_path: str = "models/rtdetr/rtdetrv2_r18vd.pth", device: str = "cuda:0",
                 conf_threshold: float = 0.25, input_size: tuple = (640, 640), wildlife_only: bool = True,
                 state_dict: Optional[dict] = None, max_batch: int = 32):
        self.config_path = config_path
        self.model_path = model_path
        self.device = device
        self.conf_threshold = conf_threshold
        self.input_size = input_size
        self.wildlife_only = wildlife_only
        self.model = None
        self.postprocessor = None
        self.transforms = None
        self._state_dict = state_dict
        self._max_batch = int(max_batch)
        self._engines = {}

    # ------------------------------------------------------------------ loading
    def _device_index(self) -> int:
        s = str(self.device)
        if not s.startswith("cuda"):
            raise RuntimeError("RTDETRDetector (B200 build) needs a CUDA device, got %r" % s)
        return int(s.split(":")[1]) if ":" in s else 0

    def _read_weights(self):
        if self._state_dict is not None:
            sd = self._state_dict
        elif str(self.model_path).endswith(".npz"):
            sd = dict(np.load(self.model_path))
        else:
            import torch
            ck = torch.load(self.model_path, map_location="cpu", weights_only=True)
            # the reference takes checkpoint['ema']['module'], else checkpoint['model'] (src/rtdetr_detector.py:139-148)
            if isinstance(ck, dict) and "ema" in ck and isinstance(ck["ema"], dict) and "module" in ck["ema"]:
                sd = ck["ema"]["module"]
            elif isinstance(ck, dict) and "model" in ck and isinstance(ck["model"], dict):
                sd = ck["model"]
            else:
                sd = ck.get("state_dict", ck) if isinstance(ck, dict) else ck
        out = {}
        for k, v in sd.items():
            if hasattr(v, "detach"):
                if not v.dtype.is_floating_point:
                    continue
                v = v.detach().cpu().float().numpy()
            out[k] = np.asarray(v, np.float32)
        from . import rtdetr_keys
        if rtdetr_keys.is_upstream(out):      # upstream rtdetrv2_pytorch names -> the names the builder reads
            out = rtdetr_keys.from_upstream(out)
        return out

    def _engine(self, n):
        from .rtdetr_engine import RTDetrEngine
        size = next((b for b in _BATCH_SIZES if b >= n), None)
        if size is None or size > max(self._max_batch, 1):
            raise ValueError("batch of %d frames exceeds max_batch %d" % (n, self._max_batch))
        if size not in self._engines:
            self._engines[size] = RTDetrEngine(self._weights, batch=size, input_size=tuple(self.input_size),
                                               device=self._device_index())
        return self._engines[size]

    def load_model(self, max_retries: int = 3) -> bool:
        """Load weights and build the batch-1 program.  Returns False on any failure (reference: :60-204)."""
        try:
            self._weights = self._read_weights()
            self._engines = {}
            self._engine(1)
            self.model = _ModelHandle(self._engines)
            logger.info("RT-DETRv2 (libtcd, sm_100a) loaded: input %s, 80 classes, conf %.2f", self.input_size,
                        self.conf_threshold)
            return True
        except Exception as e:  # noqa: BLE001 -- the reference logs and returns False
            logger.error("Failed to load RT-DETR model: %s", e)
            self.model = None
            return False

    # ------------------------------------------------------------------ inference
    def preprocess(self, img) -> tuple:
        """-> (float32 numpy [1,3,H,W] in [0,1] RGB, int64 numpy [1,2] = (orig_w, orig_h)), computed by the
        same kernel the batched path uses (parity tap; the reference returns torch tensors, :206-236)."""
        eng = self._engine(1)
        u8 = self._resize_into(eng, [img], want_u8=True)
        h, w = _frame_desc(img)[1:3]
        x = (u8[0].astype(np.float32) / np.float32(255.0)).transpose(2, 0, 1)[None]
        return x, np.array([[w, h]], np.int64)

    def _resize_into(self, eng, frames, want_u8=False):
        lib = eng.prog.lib
        n = len(frames)
        descs = [_frame_desc(f) for f in frames]
        if len({d[4] for d in descs}) != 1:
            raise ValueError("a batch must be all host frames or all device frames")
        from . import _lib
        ptrs = _lib.ptr_array([d[0] for d in descs])
        hs, ws = _lib.i32_array([d[1] for d in descs]), _lib.i32_array([d[2] for d in descs])
        st = _lib.i64_array([d[3] for d in descs])
        IH, IW = eng.IH, eng.IW
        u8 = np.empty((n, IH, IW, 3), np.uint8) if want_u8 else None
        _lib.check(lib.tcd_pil_resize(ptrs, hs, ws, st, n, 1 if descs[0][4] else 0, IH, IW, C.c_void_p(eng.x_in.ptr),
                                      u8.ctypes.data_as(C.c_void_p) if want_u8 else None,
                                      C.c_void_p(eng.prog.stream())), "tcd_pil_resize")
        return u8

    def _run(self, frames):
        """-> (labels [n,300], boxes [n,300,4], scores [n,300]) numpy, score-descending per frame."""
        n = len(frames)
        eng = self._engine(n)
        padded = list(frames) + [frames[-1]] * (eng.B - n)
        self._resize_into(eng, padded)
        sizes = [[_frame_desc(f)[2], _frame_desc(f)[1]] for f in padded]
        eng.set_orig_sizes(sizes)
        eng.run(graph=True)
        labels, boxes, scores = eng.outputs()
        return labels[:n], boxes[:n], scores[:n]

    def _format(self, labels, boxes, scores) -> List[Dict[str, Any]]:
        """src/rtdetr_detector.py:267-303 / :359-388, one frame."""
        detections = []
        for i in range(len(labels)):
            score = float(scores[i])
            if score < self.conf_threshold:
                continue
            class_id = int(labels[i])
            if self.wildlife_only and class_id not in WILDLIFE_CLASSES:
                continue
            x1, y1, x2, y2 = (float(v) for v in boxes[i])
            class_name = COCO_CLASSES[class_id] if class_id < len(COCO_CLASSES) else f"class_{class_id}"
            detections.append({
                "class_id": class_id,
                "class_name": class_name,
                "confidence": score,
                "bbox": {"x1": x1, "y1": y1, "x2": x2, "y2": y2, "area": int((x2 - x1) * (y2 - y1))},
            })
        return detections

    def detect(self, frame) -> List[Dict[str, Any]]:
        if self.model is None:
            logger.error("Model not loaded")
            return []
        try:
            labels, boxes, scores = self._run([frame])
            return self._format(labels[0], boxes[0], scores[0])
        except Exception as e:  # noqa: BLE001
            logger.error("RT-DETR detection failed: %s", e)
            return []

    def detect_batch(self, frames: List[Union[np.ndarray, Any]]) -> List[List[Dict[str, Any]]]:
        if self.model is None:
            logger.error("Model not loaded")
            return [[] for _ in frames]
        if not frames:
            return []
        labels, boxes, scores = self._run(list(frames))
        return [self._format(labels[b], boxes[b], scores[b]) for b in range(len(frames))]

    def raw_detect_batch(self, frames):
        """The reference model's deploy tuple (labels, boxes, scores) per frame, before thresholding."""
        return self._run(list(frames))

    def is_wildlife_relevant(self, class_id: int) -> bool:
        return class_id in WILDLIFE_CLASSES

    def get_class_category(self, class_id: int) -> str:
        if class_id == 0:
            return "person"
        if class_id == 14:
            return "bird"
        if class_id in MAMMAL_CLASS_IDS:
            return "mammal"
        return "other"
