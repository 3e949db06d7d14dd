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
