"""GPU: the configuration bench.py times -- DetectorEngine(max_batch=8, max_out=300, pre_mode=1 letterbox), every
pipeline slot of the library in flight (both lanes of every engine) on DIFFERENT batches -- against the CPU oracle
(letterbox restatement + fp32 YOLOX-S + postprocess).  Tolerances are north_star's: scores 1e-2, boxes 0.5 px,
class ids exact; keep sets must agree wherever the score margin exceeds the bf16 tolerance."""
import numpy as np
import pytest

from util import noise_frame, structured_frame, oracle_model, oracle_blob

pytestmark = pytest.mark.gpu

B, MAXO = 8, 300


@pytest.fixture(scope="module")
def eng():
    from telescope_cam_detection_b200.engine import DetectorEngine
    e = DetectorEngine(oracle_blob(0), input_size=(640, 640), max_batch=B, max_out=MAXO, max_src=(1080, 1920),
                       pre_mode=1)   # exactly bench.py's constructor call
    yield e
    e.close()


def _oracle_rows(frames, mode="letterbox"):
    from oracle.reference_detector import ReferenceYOLOXDetector
    det = ReferenceYOLOXDetector(oracle_model(0), 0.25, 0.45, (640, 640), True, pre_mode=mode)
    out = []
    for i in range(0, len(frames), 8):
        out.extend(det.raw_rows(frames[i:i + 8]))
    return out


def check_rows(got, ref, max_out, what, stats):
    """got: our [n,7] table of one frame; ref: the oracle's rows (all of them, score-descending).
    The network outputs agree within north_star's tolerances (test_forward_b8_letterbox_matches_oracle); what can
    still differ is a greedy-NMS decision whose IoU sits within the bf16 box noise (~0.02) of nms_thre -- with ~500
    overlapping candidates per synthetic frame a few per cent of the rows are such cases, and one flip can release or
    suppress a neighbour.  Hence: >= 94 % of the confident oracle rows must be matched (class exact, score 1e-2, box
    0.75 px) and at most 6 % of ours may be unknown to the oracle."""
    assert got.shape[0] <= max_out
    sc = got[:, 4] * got[:, 5]
    assert np.all(np.diff(sc) <= 1e-7), "%s: rows not score-descending" % what
    assert np.all(sc >= 0.25 - 1e-6)
    if ref.shape[0] == 0:
        assert got.shape[0] <= 2, what
        return
    rsc = ref[:, 4] * ref[:, 5]
    # the table holds the max_out best rows: compare with the oracle rows above the score of our last row (+ margin)
    floor = 0.27 if got.shape[0] < max_out else max(0.27, float(sc[-1]) + 0.02)
    strong = ref[rsc >= floor]

    def matched(row, table):
        same = table[table[:, 6] == row[6]]
        return bool(same.size) and bool(((np.abs(same[:, :4] - row[:4]).max(1) <= 0.75) &
                                         (np.abs(same[:, 4] * same[:, 5] - row[4] * row[5]) <= 1e-2)).any())
    miss = sum(0 if matched(r, got) else 1 for r in strong)
    extra = sum(0 if matched(g, ref) else 1 for g in got[sc >= floor])
    stats.append((what, len(strong), miss, extra))
    # per frame: one flip can cascade through a chain of overlapping boxes (the CPU emulation of the build's bf16
    # rounding points, oracle/yolox_lowp.py, differs from fp32 by up to 6 + 5 rows of 182 on these frames)
    assert miss <= max(3, 0.10 * len(strong)), "%s: %d of %d confident oracle rows unmatched" % (what, miss, len(strong))
    assert extra <= max(3, 0.10 * len(strong)), "%s: %d confident rows the oracle does not have" % (what, extra)


def test_forward_b8_letterbox_matches_oracle(eng):
    import torch
    from oracle.reference_detector import preprocess_cfg
    frames = [structured_frame(300 + i) for i in range(6)] + [noise_frame(306), structured_frame(307, 720, 1280)]
    pred = eng.forward(frames)
    with torch.no_grad():
        ref = oracle_model(0)(torch.cat([preprocess_cfg(f, (640, 640), "letterbox") for f in frames], 0)).numpy()
    assert pred.shape == ref.shape == (8, 8400, 85)
    d_box = np.abs(pred[..., :4] - ref[..., :4]).max()
    d_sc = np.abs(pred[..., 4:] - ref[..., 4:]).max()
    msg = "B=8 letterbox vs fp32 oracle: boxes %.3f px, scores %.4f" % (d_box, d_sc)
    print(msg)
    assert d_sc <= 1e-2 and d_box <= 0.5, msg


def test_every_slot_in_flight_matches_oracle(eng):
    """12 slots (6 engines x 2 lanes), a different batch on each.  (1) every batch alone (synchronous call) against
    the oracle; (2) all slots submitted before the first wait -- then again with the batches rotated -- must return
    BIT-IDENTICAL tables to (1): a table, a staging buffer, a filter table or a graph shared by mistake between
    lanes or engines shows up as a difference."""
    n_slots = eng.num_slots
    assert n_slots >= 2
    batches = []
    for s in range(n_slots):
        fr = [structured_frame(2000 + 8 * s + j) for j in range(B)]
        if s % 3 == 1:
            fr[3] = noise_frame(2500 + s)
        if s % 4 == 2:
            fr[5] = structured_frame(2600 + s, 720, 1280)       # a second camera size in the same batch
        batches.append(fr)
    stats = []
    base = []
    for s, fr in enumerate(batches):
        rows = eng.detect_batch(fr)
        ref = _oracle_rows(fr)
        for b in range(B):
            check_rows(rows[b], ref[b], MAXO, "batch %d frame %d" % (s, b), stats)
        base.append(rows)
    tot = sum(x[1] for x in stats)
    print("confident oracle rows %d, unmatched %d (%.2f %%), ours unknown to the oracle %d" % (
        tot, sum(x[2] for x in stats), 100.0 * sum(x[2] for x in stats) / tot, sum(x[3] for x in stats)))
    assert tot >= 50 * n_slots, "too few confident detections (%d) for the comparison to mean anything" % tot
    assert sum(x[2] for x in stats) <= 0.04 * tot and sum(x[3] for x in stats) <= 0.04 * tot
    for rot in (0, 5):
        order = [(s + rot) % n_slots for s in range(n_slots)]
        for s in range(n_slots):
            eng.submit(s, batches[order[s]])
        for s in range(n_slots):
            rows = eng.wait(s)
            assert len(rows) == B
            for b in range(B):
                assert np.array_equal(rows[b], base[order[s]][b]), "rot %d slot %d frame %d differs from the " \
                    "synchronous run of the same batch" % (rot, s, b)


def test_host_and_device_frames_agree_on_every_slot(eng):
    import torch
    fr = [structured_frame(2100 + j) for j in range(B)]
    dev = [torch.from_numpy(f).cuda() for f in fr]
    eng.submit(0, fr)
    base = eng.wait(0)
    for s in range(eng.num_slots):
        eng.submit(s, dev if s % 2 else fr)
    for s in range(eng.num_slots):
        rows = eng.wait(s)
        for b in range(B):
            assert np.array_equal(rows[b], base[b]), "slot %d frame %d differs from slot 0" % (s, b)


@pytest.mark.parametrize("mode", ["stretch", "letterbox"])
@pytest.mark.parametrize("shape", [(1080, 1920), (720, 1280), (1440, 2560), (333, 517)])
def test_preprocess_modes_match_oracle(mode, shape):
    """tcd_preprocess with letterbox / BGR->RGB / scale / mean / std against the CPU restatement
    (oracle/reference_detector.py::preprocess_cfg; the stretch + identity case is the reference's own ops)."""
    import torch
    from oracle.reference_detector import preprocess_cfg
    from telescope_cam_detection_b200 import engine
    frames = [noise_frame(1, *shape), structured_frame(2, *shape)]
    dev = [torch.from_numpy(f).cuda() for f in frames]
    for swap, scale, mean, std in ((0, 1.0, (0.0, 0.0, 0.0), (1.0, 1.0, 1.0)),
                                   (1, 1.0 / 255.0, (0.485, 0.456, 0.406), (0.229, 0.224, 0.225))):
        out = torch.empty(len(frames), 3, 640, 640, device="cuda", dtype=torch.float32)
        engine.preprocess([t.data_ptr() for t in dev], [shape[0]] * 2, [shape[1]] * 2, [t.stride(0) for t in dev],
                          out.data_ptr(), (640, 640), layout=0, mode=1 if mode == "letterbox" else 0, swap_rb=swap,
                          scale=scale, mean=mean, std=std)
        torch.cuda.synchronize()
        for i, f in enumerate(frames):
            ref = preprocess_cfg(f, (640, 640), mode, bool(swap), scale, mean, std)[0]
            tol = 2e-3 * (1.0 if scale == 1.0 else scale / min(std))
            err = (out[i].cpu() - ref).abs().max().item()
            assert err <= tol, "%s %s swap=%d frame %d: max abs err %.6f > %.6f" % (mode, shape, swap, i, err, tol)
