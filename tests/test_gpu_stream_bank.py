"""GPU: the batched stream bank (BASELINE configs[2]: several camera streams per GPU) against the golden vectors the
REFERENCE's own MotionFilter / EmptyFrameFilter produced -- every stream must give the single-stream verdict
sequence, bit-exact ratios included -- and against the single-stream handles at 1080p."""
import os

import numpy as np
import pytest

from util import motion_sequence

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_every_stream_reproduces_the_reference_goldens():
    """Three streams run the same golden sequence, stream k starting k steps late (idle before and after): gate
    verdicts, per-detection motion verdicts and ratios must equal the reference's for every stream, including the
    one-model-update-per-detection quirk and frames with more detections than others."""
    from telescope_cam_detection_b200.stream_bank import StreamBank
    gm = np.load(os.path.join(GOLD, "motion_ref.npz"))
    ge = np.load(os.path.join(GOLD, "empty_ref.npz"))["default"].tolist()
    seq = motion_sequence()
    n = 3
    first = np.concatenate([[0], np.cumsum(gm["n_per_frame"])])
    bank = StreamBank(n, (480, 640))
    try:
        gate = [[] for _ in range(n)]
        for t in range(len(seq) + n - 1):
            idx = [t - s if 0 <= t - s < len(seq) else None for s in range(n)]
            bank.upload([None if i is None else seq[i][0] for i in idx])
            verdicts = bank.has_motion()
            dets = [None if i is None else [{"bbox": dict(b), "i": j} for j, b in enumerate(seq[i][1])] for i in idx]
            kept = bank.filter_detections(dets)
            for s, i in enumerate(idx):
                if i is None:
                    assert verdicts[s] is None and kept[s] == []
                    continue
                gate[s].append(verdicts[s])
                k0 = int(first[i])
                want = gm["verdict"][k0:k0 + len(seq[i][1])]
                assert [d["i"] for d in kept[s]] == [j for j in range(len(want)) if want[j]], (t, s)
                for d in kept[s]:
                    assert d["has_motion"] is True and d["motion_ratio"] == gm["ratio"][k0 + d["i"]], (t, s)
        for s in range(n):
            assert gate[s] == ge, "stream %d gate verdicts" % s
        st = bank.get_stats()
        for s in range(n):
            assert st["empty_frame_filter"][s]["total_frames"] == 40
            assert st["empty_frame_filter"][s]["skipped_frames"] == 40 - sum(ge)
            assert st["motion_filter"][s]["total_detections_filtered"] == int((~gm["verdict"]).sum())
    finally:
        bank.close()


def test_bank_equals_single_stream_handles_1080p_with_many_boxes():
    """1080p, 4 streams, up to 37 boxes on one stream (three rounds of 16 model updates in registers): counts equal the
    single-stream path (one apply + full-mask clean-up + ROI count per box), which is bit-exact against cv2."""
    from telescope_cam_detection_b200.engine import Mog2Engine, EmptyGateEngine, StreamBankEngine
    h, w, n = 1080, 1920, 4
    rng = np.random.default_rng(11)
    bgs = [np.repeat(np.repeat(rng.integers(30, 220, (h // 8, w // 8, 3)), 8, 0), 8, 1).astype(np.int16) for _ in range(n)]
    bank = StreamBankEngine(n, h, w)
    singles = [Mog2Engine(h, w) for _ in range(n)]
    gates = [EmptyGateEngine(h, w, 21) for _ in range(n)]
    try:
        for t in range(8):
            frames, boxes = [], []
            for s in range(n):
                f = bgs[s] + rng.integers(-6, 7, (h, w, 3))
                if t >= 5:
                    f[100 + 50 * s:400 + 50 * s, 200 + 60 * t:700 + 60 * t] = rng.integers(170, 256, (300, 500, 3))
                frames.append(np.clip(f, 0, 255).astype(np.uint8))
                nb = [0, 1, 5, 37][(s + t) % 4]
                bs = []
                for _ in range(nb):
                    x1, y1 = int(rng.integers(0, w - 40)), int(rng.integers(0, h - 40))
                    bs.append((x1, y1, min(w, x1 + int(rng.integers(8, 700))), min(h, y1 + int(rng.integers(8, 500)))))
                boxes.append(bs)
            if t == 3:
                frames[1] = None                       # a stream without a frame this step
                boxes[1] = []
            bank.upload(frames)
            g = bank.gate(25)
            rois = np.array([b for bs in boxes for b in bs], np.int32).reshape(-1, 4)
            offs = np.cumsum([0] + [len(bs) for bs in boxes]).astype(np.int32)
            counts = bank.motion(rois, offs)
            for s in range(n):
                if frames[s] is None:
                    assert g[s] == -2
                    continue
                assert g[s] == gates[s].step(frames[s], 25)[0], (t, s)
                for j, b in enumerate(boxes[s]):
                    singles[s].apply(frames[s], repeats=1)
                    c, _ = singles[s].roi_counts(np.array([b], np.int32), 200, 21, 25)
                    assert counts[offs[s] + j] == c[0], "step %d stream %d box %d %s" % (t, s, j, b)
    finally:
        bank.close()
        for m in singles:
            m.close()
        for e in gates:
            e.close()
