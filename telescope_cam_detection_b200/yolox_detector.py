"""YOLOXDetector -- drop-in for the reference's src/yolox_detector.py (:42-426).

Same constructor, attributes and methods; `detect` / `detect_batch` return the same detection
dicts.  Underneath, frames go to libtcd as u8 (the reference uploads f32), and preprocess ->
YOLOX-S -> decode/NMS run as hand-written sm_100a kernels inside one CUDA graph per batch size.
There is no CPU or PyTorch fallback: without libtcd.so or an sm_100 GPU `load_model` returns False
(the reference's convention for load failures, :94-123) and `detect*` return empty lists.

Additions the reference does not have (all optional): `state_dict=` to load weights that are
already in memory, `max_batch=`, and the split-phase `submit_batch` / `finish_batch` pair the
pipelined SharedInferenceCoordinator uses.
"""
import logging
import time
from typing import Any, Dict, List, Optional, Union

import numpy as np

from .coco_constants import COCO_CLASSES, MAMMAL_CLASS_IDS, WILDLIFE_CLASSES

logger = logging.getLogger(__name__)

_VARIANTS = {"yolox-s": (0.33, 0.50), "yolox-m": (0.67, 0.75), "yolox-l": (1.0, 1.0), "yolox-x": (1.33, 1.25),
             "yolox-tiny": (0.33, 0.375), "yolox-nano": (0.33, 0.25)}


class _Exp:
    """What callers read from `detector.exp` (src/inference_engine_yolox.py:730-732)."""

    def __init__(self, name, input_size):
        self.depth, self.width = _VARIANTS.get(name, _VARIANTS["yolox-s"])
        self.num_classes = 80
        self.test_size = tuple(input_size)


class _ModelHandle:
    """Stands where the reference keeps an nn.Module: callers only test it for None and call
    .to(device) / .eval() on it (src/inference_engine_yolox.py:743-744)."""

    def __init__(self, engine):
        self.engine = engine

    def to(self, device):
        dev = str(device)
        if not dev.startswith("cuda"):
            raise RuntimeError("the B200 build has no CPU path (requested device %r)" % dev)
        return self

    def eval(self):
        return self


def _device_index(device) -> int:
    s = str(device)
    if not s.startswith("cuda"):
        raise RuntimeError("YOLOXDetector (B200 build) needs a CUDA device, got %r" % s)
    return int(s.split(":")[1]) if ":" in s else 0


class YOLOXDetector:
    def __init__(self, model_name: str = "yolox-s", model_path: str = "models/yolox/yolox_s.pth",
                 device: str = "cuda:0", conf_threshold: float = 0.25, nms_threshold: float = 0.45,
                 input_size: tuple = (640, 640), wildlife_only: bool = True, use_tensorrt: bool = False,
                 state_dict: Optional[dict] = None, max_batch: int = 8, max_src: tuple = (1080, 1920),
                 max_out: int = 1024):
        self.model_name = model_name
        self.model_path = model_path
        self.device = device
        self.conf_threshold = conf_threshold
        self.nms_threshold = nms_threshold
        self.input_size = input_size
        self.wildlife_only = wildlife_only
        self.use_tensorrt = use_tensorrt
        self.model = None
        self.exp = None
        self.is_tensorrt = False
        self._state_dict = state_dict
        self._blob = None
        self._max_batch, self._max_src, self._max_out = int(max_batch), tuple(max_src), int(max_out)
        self._pushed = None       # (conf, nms, input_size) the engine currently holds
        self._sizes = {}          # slot -> [(h, w)] of the frames in flight

    # ------------------------------------------------------------------ loading
    def load_model(self, max_retries: int = 3) -> bool:
        for attempt in range(max_retries):
            try:
                if self.use_tensorrt:
                    raise ValueError("use_tensorrt: TensorRT engines are not part of the B200 build")
                return self._load()
            except (RuntimeError, OSError, IOError) as e:
                if attempt < max_retries - 1:
                    wait = 2 ** attempt
                    logger.warning("Model load failed (attempt %d/%d): %s; retrying in %ds", attempt + 1,
                                   max_retries, e, wait)
                    time.sleep(wait)
                else:
                    logger.error("Failed to load YOLOX model after %d attempts: %s", max_retries, e, exc_info=True)
                    return False
            except Exception as e:  # configuration problems are not retried (reference :119-122)
                logger.error("Failed to load YOLOX model: %s", e, exc_info=True)
                return False
        return False

    def _load(self) -> bool:
        from . import weights
        logger.info("Loading YOLOX model: %s (%s)", self.model_name, self.model_path)
        self.exp = _Exp(self.model_name, self.input_size)
        sd = self._state_dict
        if sd is None:
            import torch
            sd = torch.load(self.model_path, map_location="cpu")["model"]  # same key as the reference (:137)
        self._blob = weights.pack_yolox(sd)
        self._build_engine()
        self.is_tensorrt = False
        logger.info("YOLOX loaded: input %s, %d classes, conf %.2f", self.input_size, self.exp.num_classes,
                    self.conf_threshold)
        return True

    def _build_engine(self):
        from .engine import DetectorEngine
        if self.model is not None:
            self.model.engine.close()
            self.model = None
        eng = DetectorEngine(self._blob, input_size=tuple(self.input_size), max_batch=self._max_batch,
                             num_classes=80, max_out=self._max_out, max_src=self._max_src,
                             device=_device_index(self.device))
        eng.set_thresholds(self.conf_threshold, self.nms_threshold)
        self.model = _ModelHandle(eng)
        self._pushed = (self.conf_threshold, self.nms_threshold, tuple(self.input_size))

    def _sync_settings(self):
        """conf/nms/input_size are plain attributes that callers overwrite (hot reload, OOM degradation:
        src/inference_engine_yolox.py:677-684,730-732); bring the engine up to date before a launch."""
        want = (self.conf_threshold, self.nms_threshold, tuple(self.input_size))
        if want == self._pushed:
            return
        if want[2] != self._pushed[2]:
            self._build_engine()  # new activation arena + graphs for the new network size
            if self.exp is not None:
                self.exp.test_size = want[2]
        else:
            self.model.engine.set_thresholds(want[0], want[1])
            self._pushed = want

    # ------------------------------------------------------------------ data path
    def preprocess(self, img):
        """(H,W,3) u8 BGR, NumPy or CUDA tensor -> (1,3,h,w) float32 CUDA tensor, 0-255, BGR,
        bilinear stretch (reference :186-220).  Parity/debug entry; detect* use the fused layout."""
        import torch
        from . import engine
        dev = _device_index(self.device)
        t = img if isinstance(img, torch.Tensor) else torch.from_numpy(np.ascontiguousarray(img))
        t = t.to("cuda:%d" % dev).contiguous()
        h, w = int(t.shape[0]), int(t.shape[1])
        out = torch.empty((1, 3, self.input_size[0], self.input_size[1]), dtype=torch.float32, device=t.device)
        engine.preprocess([t.data_ptr()], [h], [w], [t.stride(0)], out.data_ptr(), tuple(self.input_size), layout=0)
        return out

    @staticmethod
    def _frame_hw(frame):
        shape = tuple(frame.shape)
        if len(shape) != 3 or shape[2] != 3:
            raise ValueError("frames must be HxWx3 uint8 (BGR), got shape %s" % (shape,))
        return int(shape[0]), int(shape[1])

    def _format_model_output_to_detections(self, output: np.ndarray, orig_h: int, orig_w: int) -> List[Dict[str, Any]]:
        """(n,7) rows [x1,y1,x2,y2,obj,cls,class_id] in network pixels -> detection dicts (reference :222-282)."""
        if output is None or len(output) == 0:
            return []
        out = np.asarray(output, np.float32)
        ratio_h = orig_h / self.input_size[0]
        ratio_w = orig_w / self.input_size[1]
        # float32 products like the reference's np.float32 * python-float scalars, then widened
        xs1 = (out[:, 0] * np.float32(ratio_w)).astype(np.float64)
        ys1 = (out[:, 1] * np.float32(ratio_h)).astype(np.float64)
        xs2 = (out[:, 2] * np.float32(ratio_w)).astype(np.float64)
        ys2 = (out[:, 3] * np.float32(ratio_h)).astype(np.float64)
        conf = (out[:, 4] * out[:, 5]).astype(np.float64)
        cls = out[:, 6].astype(np.int64)
        dets = []
        for i in range(out.shape[0]):
            cid = int(cls[i])
            if self.wildlife_only and cid not in WILDLIFE_CLASSES:
                continue
            x1, y1, x2, y2 = float(xs1[i]), float(ys1[i]), float(xs2[i]), float(ys2[i])
            dets.append({
                "class_id": cid,
                "class_name": COCO_CLASSES[cid] if cid < len(COCO_CLASSES) else "class_%d" % cid,
                "confidence": float(conf[i]),
                "bbox": {"x1": x1, "y1": y1, "x2": x2, "y2": y2, "area": int((x2 - x1) * (y2 - y1))},
            })
        return dets

    def detect(self, frame) -> List[Dict[str, Any]]:
        if self.model is None:
            logger.error("Model not loaded")
            return []
        return self.detect_batch([frame])[0]

    def detect_batch(self, frames: List[Union[np.ndarray, Any]]) -> List[List[Dict[str, Any]]]:
        if self.model is None:
            logger.error("Model not loaded")
            return [[] for _ in frames]
        if not frames:
            return []
        out: List[List[Dict[str, Any]]] = []
        mb = self._max_batch
        for i in range(0, len(frames), mb):  # batches larger than the arena run as consecutive launches
            chunk = frames[i:i + mb]
            self.submit_batch(0, chunk)
            out.extend(self.finish_batch(0))
        return out

    # split-phase pair (pipelined coordinator): slot 0/1 own independent streams and arenas
    def submit_batch(self, slot: int, frames) -> None:
        sizes = [self._frame_hw(f) for f in frames]
        self._sync_settings()
        self.model.engine.submit(slot, frames)
        self._sizes[slot] = sizes

    def finish_batch(self, slot: int) -> List[List[Dict[str, Any]]]:
        sizes = self._sizes.pop(slot)
        rows = self.model.engine.wait(slot)
        return [self._format_model_output_to_detections(r, h, w) for r, (h, w) in zip(rows, sizes)]

    # ------------------------------------------------------------------ helpers of the reference API
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
