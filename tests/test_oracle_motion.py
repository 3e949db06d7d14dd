"""CPU: pins oracle/motion.py (NumPy restatement of MOG2 / u8 GaussianBlur / BGR2GRAY and of the
reference's MotionFilter / EmptyFrameFilter logic) against (1) the golden vectors produced by the
reference's own classes (tests/golden/make_golden.py) and (2) cv2, the reference's dependency,
when it is importable.  All integer / index work: bit-exact."""
import os

import numpy as np
import pytest

from util import motion_sequence
from oracle import motion

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_gaussian_taps_are_the_recovered_opencv_table():
    assert motion.gaussian_kernel_q8(21).tolist() == [0, 2, 2, 4, 6, 11, 15, 20, 25, 28, 30, 28, 25, 20, 15, 11, 6,
                                                      4, 2, 2, 0]
    for n in (1, 3, 5, 7, 9, 13, 21, 31):
        assert motion.gaussian_kernel_q8(n).sum() == 256


def test_motion_filter_matches_reference_golden():
    g = np.load(os.path.join(GOLD, "motion_ref.npz"))
    seq = motion_sequence()
    assert [len(b) for _, b in seq] == g["n_per_frame"].tolist()
    mog = motion.Mog2(500, 16, True)
    verdict, ratio = [], []
    for frame, boxes in seq:
        for b in boxes:  # one model update per detection, like src/motion_filter.py:88,161-162
            mask = motion.clean_mask(mog.apply(frame), 21, True)
            ok, r = motion.bbox_verdict(mask, b)
            verdict.append(ok)
            ratio.append(r)
    assert np.array_equal(np.array(verdict), g["verdict"])
    assert np.array_equal(np.array(ratio), g["ratio"])  # integer counts / integer areas: exact
    assert 20 < sum(verdict) < len(verdict) - 20, "golden sequence must exercise both verdicts"


def test_empty_frame_filter_matches_reference_golden():
    g = np.load(os.path.join(GOLD, "empty_ref.npz"))
    for key, (area, thr, blur) in {"default": (200, 25, 21), "tuned": (5000, 10, 9)}.items():
        ef = motion.EmptyFrame(blur)
        got = []
        for frame, _ in motion_sequence():
            n, _ = ef.step(frame, thr)
            got.append(True if n < 0 else n >= area)
        assert got == g[key].tolist(), key
    assert not all(g["default"]) and any(g["default"])


def _small_sequence(h=48, w=64, n=60, seed=1):
    """sensor noise, a moving bright object, a 0.6x darkening episode, a late flickering region"""
    rng = np.random.default_rng(seed)
    bg = rng.integers(30, 220, (h, w, 3)).astype(np.int16)
    for t in range(n):
        f = bg + rng.integers(-6, 7, (h, w, 3))
        if t >= 10:
            x = (t * 3) % (w - 12)
            f[10:26, x:x + 12] = rng.integers(200, 256, (16, 12, 3))
        if 25 <= t < 40:
            f[30:44, 5:30] = (f[30:44, 5:30] * 0.6).astype(np.int16)
        if t >= 45:
            f[2:8, 40:60] = 255 if t % 2 else 250
        yield np.clip(f, 0, 255).astype(np.uint8)


def test_restatement_is_bit_exact_against_cv2():
    cv2 = pytest.importorskip("cv2")
    rng = np.random.default_rng(0)
    for shape in [(37, 53), (200, 260)]:
        img = rng.integers(0, 256, shape, dtype=np.uint8)
        for k in (3, 5, 9, 21, 31):
            assert np.array_equal(cv2.GaussianBlur(img, (k, k), 0), motion.gaussian_blur_u8(img, k)), (shape, k)
        bgr = rng.integers(0, 256, shape + (3,), dtype=np.uint8)
        assert np.array_equal(cv2.cvtColor(bgr, cv2.COLOR_BGR2GRAY), motion.bgr_to_gray(bgr))
    for shadows, hist, vt in [(True, 500, 16), (False, 500, 16), (True, 20, 25)]:
        ref = cv2.createBackgroundSubtractorMOG2(history=hist, varThreshold=vt, detectShadows=shadows)
        mine = motion.Mog2(hist, vt, shadows)
        seen = set()
        for frame in _small_sequence():
            a, b = ref.apply(frame), mine.apply(frame)
            assert np.array_equal(a, b)
            seen |= set(np.unique(a).tolist())
        assert seen == ({0, 127, 255} if shadows else {0, 255})
