"""GPU parity of the motion path through the C ABI: MOG2 masks, mask clean-up, ROI verdicts and
the empty-frame gate are integer / index work -> bit-exact against the golden vectors made by the
reference's own classes and against the CPU oracle (oracle/motion.py) at full 1080p."""
import os

import numpy as np
import pytest

from util import motion_sequence

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_motion_filter_equals_reference_golden():
    from telescope_cam_detection_b200.motion_filter import MotionFilter
    g = np.load(os.path.join(GOLD, "motion_ref.npz"))
    mf = MotionFilter()
    verdict, ratio = [], []
    for frame, boxes in motion_sequence():
        for b in boxes:
            ok, r = mf.has_motion_in_bbox(frame, b)
            verdict.append(ok)
            ratio.append(r)
    assert np.array_equal(np.array(verdict), g["verdict"])
    assert np.array_equal(np.array(ratio), g["ratio"])
    mf.cleanup()


def test_filter_detections_annotates_and_counts_like_the_reference():
    from telescope_cam_detection_b200.motion_filter import MotionFilter
    g = np.load(os.path.join(GOLD, "motion_ref.npz"))
    mf = MotionFilter()
    k = 0
    for frame, boxes in motion_sequence():
        dets = [{"bbox": dict(b), "i": i} for i, b in enumerate(boxes)]
        kept = mf.filter_detections(frame, dets)
        want = g["verdict"][k:k + len(boxes)]
        assert [d["i"] for d in kept] == [i for i in range(len(boxes)) if want[i]]
        for d in kept:
            assert d["has_motion"] is True and d["motion_ratio"] == g["ratio"][k + d["i"]]
        k += len(boxes)
    assert mf.get_stats() == {"total_frames": 40, "total_detections_filtered": int((~g["verdict"]).sum()),
                              "motion_required": True}
    assert mf.filter_detections(motion_sequence(n=1)[0][0], []) == []          # no detections: model untouched
    off = MotionFilter(motion_required=False)
    assert off.filter_detections(None, [1, 2]) == [1, 2]


@pytest.mark.parametrize("shadows,history,vt", [(True, 500, 16), (False, 500, 16), (True, 12, 30)])
def test_mog2_masks_bit_exact_vs_oracle_1080p(shadows, history, vt):
    """full-size frames, device-resident input, repeats > 1 (several updates in registers)"""
    import torch
    from oracle import motion
    from telescope_cam_detection_b200.engine import Mog2Engine
    h, w = 1080, 1920
    rng = np.random.default_rng(5)
    bg = rng.integers(30, 220, (h // 8, w // 8, 3))
    bg = np.repeat(np.repeat(bg, 8, 0), 8, 1).astype(np.int16)
    eng = Mog2Engine(h, w, history, vt, shadows)
    ora = motion.Mog2(history, vt, shadows)
    seen = set()
    for t in range(9):
        f = bg + rng.integers(-6, 7, (h, w, 3))
        if t >= 6:
            f[200:500, 300 + 40 * t:700 + 40 * t] = rng.integers(180, 256, (300, 400, 3))
        if t == 8:
            f[600:900, 100:900] = (f[600:900, 100:900] * 0.6).astype(np.int16)
        f = np.clip(f, 0, 255).astype(np.uint8)
        reps = 3 if t == 4 else 1
        src = torch.from_numpy(f).cuda() if t % 2 else f
        got = eng.apply(src, repeats=reps, want_mask=True)
        for _ in range(reps):
            ref = ora.apply(f)
        assert np.array_equal(got, ref), "frame %d: %d mask pixels differ" % (t, int((got != ref).sum()))
        seen |= set(np.unique(ref).tolist())
        rois = np.array([[0, 0, w, h], [300, 200, 1100, 500], [5, 5, 6, 6], [1900, 1000, 1920, 1080]], np.int32)
        counts, clean = eng.roi_counts(rois, 200, 21, 25, want_mask=True)
        cref = motion.clean_mask(ref, 21, True)
        assert np.array_equal(clean, cref)
        assert counts.tolist() == [int(np.count_nonzero(cref[y1:y2, x1:x2])) for x1, y1, x2, y2 in rois]
    assert 255 in seen and 0 in seen
    eng.close()


def test_empty_frame_filter_equals_reference_golden():
    from telescope_cam_detection_b200.empty_frame_filter import EmptyFrameFilter
    g = np.load(os.path.join(GOLD, "empty_ref.npz"))
    seq = motion_sequence()
    for key, kw in {"default": {}, "tuned": {"min_motion_area": 5000, "threshold": 10, "blur_size": 9}}.items():
        ef = EmptyFrameFilter(**kw)
        assert [ef.has_motion(f) for f, _ in seq] == g[key].tolist(), key
        st = ef.get_stats()
        assert st["total_frames"] == 40 and st["motion_frames"] == int(g[key].sum())
        assert st["skipped_frames"] == 40 - int(g[key].sum())
        ef.reset()
        assert ef.has_motion(seq[0][0]) is True                                   # first frame after reset


def test_empty_gate_blur_plane_and_counts_bit_exact_1080p():
    import torch
    from oracle import motion
    from telescope_cam_detection_b200.engine import EmptyGateEngine
    rng = np.random.default_rng(9)
    eng, ora = EmptyGateEngine(1080, 1920, 21), motion.EmptyFrame(21)
    base = rng.integers(0, 256, (1080, 1920, 3), dtype=np.uint8)
    for t in range(4):
        f = base.copy()
        if t:
            f[100 * t:100 * t + 300, 500:1200] = rng.integers(0, 256, (300, 700, 3), dtype=np.uint8)
        n, blur = eng.step(torch.from_numpy(f).cuda() if t == 2 else f, 25, want_blur=True)
        rn, rblur = ora.step(f, 25)
        assert np.array_equal(blur, rblur) and n == rn
    g_eng, g_ora = EmptyGateEngine(270, 480, 9), motion.EmptyFrame(9)            # grayscale input path
    for t in range(3):
        gray = rng.integers(0, 256, (270, 480), dtype=np.uint8)
        assert g_eng.step(gray, 12)[0] == g_ora.step(gray, 12)[0]
