"""Shared helpers for the parity tests (synthetic frames, oracle model cache)."""
import functools
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def noise_frame(seed, h=1080, w=1920):
    """Same convention as the reference's smoke tests (tests/test_system_startup.py:54)."""
    return np.random.default_rng(seed).integers(0, 255, (h, w, 3), dtype=np.uint8)


def structured_frame(seed, h=1080, w=1920):
    """Low-pass 8x8 blocks + noise (like oracle.synth_weights.calib_frame): ~1-3 % of anchors fire."""
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 255, (h // 8 + 1, w // 8 + 1, 3), dtype=np.uint8)
    img = np.repeat(np.repeat(base, 8, 0), 8, 1)[:h, :w].astype(np.int16)
    img += rng.integers(-20, 20, (h, w, 3), dtype=np.int16)
    return np.clip(img, 0, 255).astype(np.uint8)


@functools.lru_cache(maxsize=2)
def oracle_model(seed=0):
    from oracle import synth_weights
    return synth_weights.make_model(seed=seed)


@functools.lru_cache(maxsize=2)
def oracle_blob(seed=0):
    from telescope_cam_detection_b200 import weights
    return weights.pack_yolox(oracle_model(seed).state_dict())


def motion_sequence(seed=7, h=480, w=640, n=40):
    """Synthetic camera stream for the motion tests: static low-pass background + sensor noise,
    moving bright rectangles from frame 8 on, a darkening ("shadow") episode, a few fully static
    frames.  -> list of (frame u8 HxWx3, [bbox dicts]) ; boxes include moving, static, inverted,
    out-of-frame and degenerate ones (the cases src/motion_filter.py:100-113 guards)."""
    rng = np.random.default_rng(seed)
    coarse = rng.integers(40, 200, (h // 16 + 1, w // 16 + 1, 3))
    bg = np.repeat(np.repeat(coarse, 16, 0), 16, 1)[:h, :w].astype(np.int16)
    out = []
    for t in range(n):
        f = bg + rng.integers(-5, 6, (h, w, 3))
        boxes = []
        if t >= 8 and t % 9 != 0:  # every 9th frame stays static
            x = 20 + (t * 11) % (w - 140)
            f[100:180, x:x + 90] = rng.integers(190, 256, (80, 90, 3))
            boxes.append({"x1": x - 10.5, "y1": 92.25, "x2": x + 101.75, "y2": 188.5})
            y = 250 + (t * 5) % 120
            f[y:y + 40, 400:470] = rng.integers(0, 40, (40, 70, 3))
            boxes.append({"x1": 395.0, "y1": float(y - 6), "x2": 476.0, "y2": float(y + 47)})
        if 20 <= t < 30:
            f[300:420, 50:250] = (f[300:420, 50:250] * 0.62).astype(np.int16)
            boxes.append({"x1": 50.0, "y1": 300.0, "x2": 250.0, "y2": 420.0})
        boxes.append({"x1": 500.0, "y1": 20.0, "x2": 620.0, "y2": 90.0})           # static region
        if t % 4 == 0:
            boxes.append({"x1": 300.0, "y1": 200.0, "x2": 180.0, "y2": 120.0})     # inverted corners
        if t % 5 == 0:
            boxes.append({"x1": float(w - 30), "y1": float(h - 25), "x2": float(w + 80), "y2": float(h + 60)})
        if t % 7 == 0:
            boxes.append({"x1": float(w + 5), "y1": 10.0, "x2": float(w + 50), "y2": 60.0})  # off-frame
        out.append((np.clip(f, 0, 255).astype(np.uint8), boxes))
    return out


def enhance_crops():
    """Seeded Stage-2 crops for the enhancement tests: structured colour crops (tile-divisible, not divisible in one /
    both axes, dark, bright), one of pure noise, and a grey crop."""
    rng = np.random.default_rng(2024)
    out = []
    for k, (h, w) in enumerate([(128, 128), (96, 203), (131, 77), (112, 168), (57, 64)]):
        img = structured_frame(300 + k, h, w).astype(np.int16)
        if k == 2:
            img = img // 4            # a dark crop: the linear segments of both colour conversions
        if k == 3:
            img = 128 + img // 2
        out.append(np.clip(img, 0, 255).astype(np.uint8))
    out.append(rng.integers(0, 256, (80, 120, 3), dtype=np.uint8))
    out.append(np.ascontiguousarray(structured_frame(310, 150, 170)[..., 1]))
    return out
