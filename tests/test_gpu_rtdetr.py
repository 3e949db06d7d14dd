"""End-to-end parity of the libtcd RT-DETRv2-R50 program (bf16 activations, fp32 accumulate, through the
C ABI) against the fp32 CPU oracle (transformers RT-DETRv2, oracle/rtdetr.py) on the same seeded
weights and inputs.  PARITY UNPINNED (the reference's RT-DETR dependency is absent).

Tolerances (north_star): scores within 1e-2 abs, boxes within 0.5 px (network-input pixels, 640),
class ids exact where the score margin exceeds the tolerance.  The query selection (top-300 of 8400
encoder tokens) is a discrete step whose margins (~3e-4 between adjacent ranks) are below bf16
noise, so queries are compared by identity (their anchor index): >= 90 % of the oracle's queries must
be selected, and scores / boxes are compared on the common ones."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

B = 2


@pytest.fixture(scope="module")
def run():
    import torch
    from oracle import rtdetr as oracle
    from telescope_cam_detection_b200.rtdetr_engine import RTDetrEngine

    model = oracle.make_model(0)
    torch.manual_seed(1)
    x = torch.rand(B, 3, 640, 640).to(torch.bfloat16).float()
    ref = oracle.forward(model, x)
    eng = RTDetrEngine(oracle.state_dict_numpy(model), batch=B, keep_taps=True)
    eng.set_input(x.numpy())
    eng.set_orig_sizes([[1920, 1080]] * B)
    eng.run(graph=True)
    yield model, x, ref, eng
    eng.close()


def _rms_rel(got, ref):
    ref = np.asarray(ref, np.float32)
    return float(np.sqrt(((got - ref) ** 2).mean()) / np.sqrt((ref ** 2).mean()))


def test_backbone_and_encoder_stages(run):
    _, _, ref, eng = run
    for i in range(3):
        assert _rms_rel(eng.read_tap("backbone.%d" % (i + 1)), ref["backbone"][i].permute(0, 2, 3, 1).numpy()) <= 0.015
    assert _rms_rel(eng.read_tap("aifi"), ref["aifi"].permute(0, 2, 3, 1).numpy()) <= 0.015
    for i in range(3):
        assert _rms_rel(eng.read_tap("encoder.%d" % i), ref["encoder"][i].permute(0, 2, 3, 1).numpy()) <= 0.02


def test_query_selection_and_decoder(run):
    import torch
    _, _, ref, eng = run
    logits, boxes, topk = eng.raw_outputs()
    rt = ref["topk"].numpy()
    for b in range(B):
        common = np.intersect1d(topk[b], rt[b])
        assert len(common) >= 0.9 * rt.shape[1], "only %d of the oracle's queries selected" % len(common)
        mine = {int(a): i for i, a in enumerate(topk[b])}
        theirs = {int(a): i for i, a in enumerate(rt[b])}
        im = np.array([mine[int(a)] for a in common])
        it = np.array([theirs[int(a)] for a in common])
        # with ~8 % of the query set different, the self-attention context differs too: looser bounds here,
        # the north-star tolerances are asserted on the identical query set in the next test
        for li in range(6):
            h_ref = ref["dec_hidden"][b, li].numpy()[it]
            assert _rms_rel(eng.read_tap("dec%d" % li)[b, 0][im], h_ref) <= 0.08, li
        s_got = 1 / (1 + np.exp(-logits[b][im]))
        s_ref = torch.sigmoid(ref["logits"][b]).numpy()[it]
        assert np.abs(s_got - s_ref).max() <= 3e-2, np.abs(s_got - s_ref).max()
        d_box = np.abs(boxes[b][im] - ref["pred_boxes"][b].numpy()[it]).max() * 640
        assert d_box <= 1.5, d_box


def test_decoder_on_the_oracles_query_set(run):
    """Same weights and input, query selection forced to the oracle's indices: every one of the 300
    queries is compared in order at the north-star tolerances."""
    import torch
    from oracle import rtdetr as oracle
    from telescope_cam_detection_b200.rtdetr_engine import RTDetrEngine
    model, x, ref, _ = run
    eng = RTDetrEngine(oracle.state_dict_numpy(model), batch=B, keep_taps=True, external_topk=True)
    eng.set_input(x.numpy())
    eng.set_orig_sizes([[1920, 1080]] * B)
    eng.set_topk(ref["topk"].numpy())
    eng.run(graph=True)
    logits, boxes, _ = eng.raw_outputs()
    for li in range(6):
        assert _rms_rel(eng.read_tap("dec%d" % li)[:, 0], ref["dec_hidden"][:, li].numpy()) <= 0.03, li
    s_got = 1 / (1 + np.exp(-logits))
    s_ref = torch.sigmoid(ref["logits"]).numpy()
    assert np.abs(s_got - s_ref).max() <= 1e-2, np.abs(s_got - s_ref).max()
    d_box = np.abs(boxes - ref["pred_boxes"].numpy()).max() * 640
    assert d_box <= 0.5, d_box
    # class ids: arg-max class agrees wherever the oracle's margin exceeds the tolerance
    top2 = np.sort(s_ref, 2)[..., -2:]
    clear = (top2[..., 1] - top2[..., 0]) > 2e-2
    assert np.array_equal(s_got.argmax(2)[clear], s_ref.argmax(2)[clear])
    eng.close()


def test_postprocessor_output(run):
    import torch
    from oracle import rtdetr as oracle
    _, _, ref, eng = run
    labels, boxes, scores = eng.outputs()
    rl, rb, rs = oracle.postprocess(ref["logits"], ref["pred_boxes"], torch.tensor([[1920, 1080]] * B))
    assert labels.shape == (B, 300) and boxes.shape == (B, 300, 4)
    assert np.all(scores[:, :-1] >= scores[:, 1:])
    # the engine's own postprocess is exact on the engine's own logits / boxes
    lg, pb, _ = eng.raw_outputs()
    ol, ob, os_ = oracle.postprocess(torch.from_numpy(lg), torch.from_numpy(pb), torch.tensor([[1920, 1080]] * B))
    assert np.array_equal(labels, ol.numpy())
    assert np.allclose(scores, os_.numpy(), atol=1e-6) and np.allclose(boxes, ob.numpy(), atol=2e-3)
    # and close to the oracle's: the k-th score agrees within tolerance for every rank
    assert np.abs(scores - rs.numpy()).max() <= 1e-2


def test_detector_shim_end_to_end(run):
    """RTDETRDetector.detect_batch on 1080p frames: preprocess kernel + program + dict format, against the
    oracle fed by the reference's own preprocess body."""
    import torch
    from oracle import rtdetr as oracle
    from telescope_cam_detection_b200.rtdetr_detector import RTDETRDetector
    from util import structured_frame as synth_frame

    model = run[0]
    det = RTDETRDetector(state_dict=oracle.state_dict_numpy(model), conf_threshold=0.05, wildlife_only=False, max_batch=2)
    assert det.load_model()
    frames = [synth_frame(3), synth_frame(4)]
    x, size = det.preprocess(frames[0])
    xr, sr, _ = oracle.preprocess_ref(frames[0])
    assert np.array_equal(size, sr.numpy()) and np.array_equal(x, xr.numpy())     # resize is bit-exact
    out = det.detect_batch(frames)
    assert len(out) == 2
    xs = torch.cat([oracle.preprocess_ref(f)[0] for f in frames])
    ref = oracle.forward(model, xs)
    rl, rb, rs = oracle.postprocess(ref["logits"], ref["pred_boxes"], torch.tensor([[1920, 1080]] * 2))
    for b in range(2):
        n_ref = int((rs[b] >= 0.05).sum())
        assert abs(len(out[b]) - n_ref) <= max(3, 0.05 * n_ref)
        for d in out[b][:5]:
            assert set(d) == {"class_id", "class_name", "confidence", "bbox"}
            assert set(d["bbox"]) == {"x1", "y1", "x2", "y2", "area"}
            assert d["bbox"]["area"] == int((d["bbox"]["x2"] - d["bbox"]["x1"]) * (d["bbox"]["y2"] - d["bbox"]["y1"]))
        got_top = np.array([d["confidence"] for d in out[b][:20]])
        assert np.abs(got_top - rs[b][:20].numpy()).max() <= 1e-2
    det.wildlife_only = True
    only = det.detect(frames[0])
    assert all(d["class_id"] in (0, 14, 15, 16, 21) for d in only)
    assert det.detect_batch([]) == []


def test_r18vd_variant_loads_and_matches_oracle():
    """The variant the reference's shipped config selects (config/config.yaml:74,91): BasicBlock backbone, CSPRep
    layers with hidden_expansion 0.5 (conv3), three decoder layers.  ADVICE r1: these weights used to fail with a
    KeyError on layer.2."""
    import torch
    from oracle import rtdetr as oracle
    from telescope_cam_detection_b200.rtdetr_engine import RTDetrEngine

    model = oracle.make_model(0, variant="r18vd")
    torch.manual_seed(2)
    x = torch.rand(1, 3, 640, 640).to(torch.bfloat16).float()
    ref = oracle.forward(model, x)
    eng = RTDetrEngine(oracle.state_dict_numpy(model), batch=1, keep_taps=True)
    try:
        eng.set_input(x.numpy())
        eng.set_orig_sizes([[1920, 1080]])
        eng.run(graph=True)
        for i in range(3):
            assert _rms_rel(eng.read_tap("backbone.%d" % (i + 1)), ref["backbone"][i].permute(0, 2, 3, 1).numpy()) <= 0.015
        for i in range(3):
            assert _rms_rel(eng.read_tap("encoder.%d" % i), ref["encoder"][i].permute(0, 2, 3, 1).numpy()) <= 0.02
        logits, boxes, topk = eng.raw_outputs()
        common = np.intersect1d(topk[0], ref["topk"].numpy()[0])
        assert len(common) >= 0.85 * 300, "only %d of the oracle's queries selected" % len(common)
        assert ref["dec_hidden"].shape[1] == 3
    finally:
        eng.close()
