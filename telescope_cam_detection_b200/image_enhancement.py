"""ImageEnhancer: crop enhancement before Stage-2 classification.

Same class surface as /root/reference/src/image_enhancement.py:28-216 (constructor arguments, `enhance`,
`enhance_clahe_bilateral`, `enhance_realesrgan`) over libtcd's batched CLAHE + bilateral kernels (csrc/enhance.cu,
include/tcd.h "crop enhancement").  Added: `enhance_batch(crops)` -- every crop of a frame in ONE call (three kernel
launches for the whole batch), which is how TwoStageDetectionPipeline uses it.

Inputs are what the reference takes (HxWx3 BGR or HxW uint8 NumPy arrays); CUDA uint8 torch tensors of the same layout
are accepted too and then stay on the device (the result is a tensor).  Real-ESRGAN is a third-party network
(`realesrgan` / `basicsr`, not on this image) and stays out of scope: method="realesrgan" raises the same ValueError /
load error path the reference takes when the model file or the package is missing.
"""
import ctypes as C
import logging
from typing import List, Optional, Tuple

import numpy as np

from . import _lib
from .engine import _is_torch, device_index

logger = logging.getLogger(__name__)


class ImageEnhancer:
    def __init__(self, method: str = "realesrgan", device: Optional[str] = None,
                 realesrgan_model_path: Optional[str] = None, realesrgan_scale: int = 4, realesrgan_tile: int = 512,
                 realesrgan_tile_pad: int = 10, realesrgan_half: bool = False, clahe_clip_limit: float = 2.0,
                 clahe_tile_grid_size: Tuple[int, int] = (8, 8), bilateral_d: int = 9, bilateral_sigma_color: int = 75,
                 bilateral_sigma_space: int = 75):
        self.method = method
        self.device = device if device is not None else "cuda:0"
        self.realesrgan_model_path = realesrgan_model_path
        self.realesrgan_scale = realesrgan_scale
        self.realesrgan_tile = realesrgan_tile
        self.realesrgan_tile_pad = realesrgan_tile_pad
        self.realesrgan_half = realesrgan_half
        self.realesrgan_model = None
        self.clahe_clip_limit = clahe_clip_limit
        self.clahe_tile_grid_size = tuple(clahe_tile_grid_size)
        self.bilateral_d = bilateral_d
        self.bilateral_sigma_color = bilateral_sigma_color
        self.bilateral_sigma_space = bilateral_sigma_space
        self._h = None
        logger.info("ImageEnhancer initialized: method=%s", method)
        if method == "realesrgan":
            if realesrgan_model_path is None:
                raise ValueError("realesrgan_model_path required when method='realesrgan'")
            self._load_realesrgan()

    def _load_realesrgan(self):
        # the reference imports basicsr / realesrgan here and re-raises on failure (:106-143); neither is on this image
        raise ImportError("Real-ESRGAN (basicsr / realesrgan) is not part of the B200 build: use method='clahe' or 'none'")

    # ------------------------------------------------------------------ libtcd handle
    def _handle(self):
        if self._h is None:
            lib = _lib.load()
            cfg = _lib.EnhanceCfg()
            cfg.clip_limit = float(self.clahe_clip_limit)
            # cv2 tileGridSize is (width, height) = (tiles in x, tiles in y)
            cfg.grid_x, cfg.grid_y = int(self.clahe_tile_grid_size[0]), int(self.clahe_tile_grid_size[1])
            cfg.d = int(self.bilateral_d)
            cfg.sigma_color, cfg.sigma_space = float(self.bilateral_sigma_color), float(self.bilateral_sigma_space)
            h = C.c_void_p()
            _lib.check(lib.tcd_enhancer_create(C.byref(cfg), device_index(self.device), C.byref(h)), "tcd_enhancer_create")
            self._lib, self._h = lib, h
        return self._h

    def close(self):
        if self._h is not None:
            self._lib.tcd_enhancer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ------------------------------------------------------------------ the path
    def enhance_batch(self, images: List) -> List:
        """enhance_clahe_bilateral for a list of crops (all NumPy or all CUDA tensors; mixed channel counts are split
        into one call per channel count).  Returns new arrays / tensors of the inputs' shapes."""
        if not images:
            return []
        on_dev = _is_torch(images[0])
        out = [None] * len(images)
        for cn in (3, 1):
            idx = [i for i, im in enumerate(images) if (im.ndim == 3 and im.shape[2] == 3) == (cn == 3)]
            if not idx:
                continue
            batch = [images[i] for i in idx]
            for im in batch:
                if _is_torch(im) != on_dev:
                    raise TypeError("enhance_batch: crops must be all NumPy arrays or all CUDA tensors")
                if cn == 1 and im.ndim != 2:
                    raise ValueError("crop must be HxWx3 or HxW, got %s" % (tuple(im.shape),))
            res = self._run(batch, cn, on_dev)
            for i, r in zip(idx, res):
                out[i] = r
        return out

    def _run(self, batch, cn, on_dev):
        n = len(batch)
        keep, ptrs, hs, ws, strides, outs, optrs = [], [], [], [], [], [], []
        if on_dev:
            import torch
            for t in batch:
                if t.dtype != torch.uint8 or not t.is_cuda:
                    raise TypeError("device crops must be CUDA uint8 tensors")
                if t.stride(-1) != 1 or (cn == 3 and t.stride(1) != 3):
                    t = t.contiguous()
                keep.append(t)
                ptrs.append(t.data_ptr()); hs.append(t.shape[0]); ws.append(t.shape[1]); strides.append(t.stride(0))
                o = torch.empty(tuple(t.shape), dtype=torch.uint8, device=t.device)
                outs.append(o); optrs.append(o.data_ptr())
            torch.cuda.current_stream(batch[0].device).synchronize()   # the crops were produced on the caller's stream
        else:
            for a in batch:
                a = np.asarray(a)
                if a.dtype != np.uint8:
                    raise TypeError("crops must be uint8, got %s" % a.dtype)
                if a.strides[-1] != 1 or (cn == 3 and a.strides[1] != 3):
                    a = np.ascontiguousarray(a)
                keep.append(a)
                ptrs.append(a.ctypes.data); hs.append(a.shape[0]); ws.append(a.shape[1]); strides.append(a.strides[0])
                o = np.empty(a.shape, np.uint8)
                outs.append(o); optrs.append(o.ctypes.data)
        h = self._handle()
        _lib.check(self._lib.tcd_enhance_clahe_bilateral(
            h, _lib.ptr_array(ptrs), _lib.i32_array(hs), _lib.i32_array(ws), _lib.i64_array(strides), n, cn, 1 if on_dev else 0,
            _lib.ptr_array(optrs), None, 1 if on_dev else 0), "tcd_enhance_clahe_bilateral")
        return outs

    def enhance_clahe_bilateral(self, image):
        return self.enhance_batch([image])[0]

    def enhance_realesrgan(self, image):
        if self.realesrgan_model is None:
            raise RuntimeError("Real-ESRGAN model not loaded")
        raise RuntimeError("Real-ESRGAN is not part of the B200 build")

    def enhance(self, image):
        if self.method == "none":
            return image
        if self.method == "clahe":
            return self.enhance_clahe_bilateral(image)
        if self.method == "realesrgan":
            return self.enhance_realesrgan(image)
        logger.warning("Unknown enhancement method: %s, returning original", self.method)
        return image
