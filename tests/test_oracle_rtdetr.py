"""CPU tests of the RT-DETR oracle (test infrastructure) and of the host-side helpers of the
product's program builder (no GPU, no libtcd compute)."""
import numpy as np
import pytest
import torch

from oracle import rtdetr as oracle
from telescope_cam_detection_b200 import rtdetr_engine as eng


def test_pil_resize_restatement_is_bit_exact_vs_pillow():
    from PIL import Image

    rng = np.random.default_rng(0)
    for (h, w, oh, ow) in ((270, 480, 160, 160), (108, 192, 64, 64), (50, 37, 64, 96)):
        img = rng.integers(0, 255, (h, w, 3), dtype=np.uint8)
        ref = np.asarray(Image.fromarray(img).resize((ow, oh), Image.BILINEAR))
        got = oracle.pil_resize_ref(img, oh, ow)
        assert np.array_equal(ref, got), (h, w, oh, ow, int((ref != got).sum()))


def test_preprocess_ref_follows_the_reference_body():
    rng = np.random.default_rng(1)
    frame = rng.integers(0, 255, (108, 192, 3), dtype=np.uint8)
    t, size, u8 = oracle.preprocess_ref(frame, (64, 64))
    assert t.shape == (1, 3, 64, 64) and t.dtype == torch.float32
    assert size.tolist() == [[192, 108]]           # (orig_w, orig_h), src/rtdetr_detector.py:234
    assert float(t.max()) <= 1.0
    # channel 0 of the tensor is R = channel 2 of the BGR frame
    assert np.array_equal(u8, oracle.pil_resize_ref(frame[:, :, ::-1], 64, 64))


def test_postprocess_matches_transformers_postprocessor_semantics():
    torch.manual_seed(0)
    logits = torch.randn(2, 300, 80)
    boxes = torch.rand(2, 300, 4) * 0.5 + 0.25
    sizes = torch.tensor([[1920, 1080], [640, 480]])
    labels, xyxy, scores = oracle.postprocess(logits, boxes, sizes)
    assert labels.shape == (2, 300) and xyxy.shape == (2, 300, 4)
    assert bool((scores[:, :-1] >= scores[:, 1:]).all())
    flat = torch.sigmoid(logits).flatten(1)
    b, i = 1, 7
    idx = int(torch.topk(flat[b], 300).indices[i])
    assert int(labels[b, i]) == idx % 80
    q = idx // 80
    cx, cy, w, h = boxes[b, q]
    exp = torch.stack([(cx - w / 2) * 640, (cy - h / 2) * 480, (cx + w / 2) * 640, (cy + h / 2) * 480])
    assert torch.allclose(xyxy[b, i], exp, atol=1e-4)


def test_pos_embed_and_anchors_match_transformers():
    from transformers.models.rt_detr_v2 import modeling_rt_detr_v2 as hf

    pe = hf.RTDetrV2SinePositionEmbedding(256, 10000)(width=20, height=20, device="cpu", dtype=torch.float32)[0]
    got = eng.sincos_pos_embed(20, 20, 256)
    assert np.allclose(got, pe.numpy(), atol=1e-6)
    model = oracle.make_model(0)
    anchors, valid = model.model.generate_anchors(((80, 80), (40, 40), (20, 20)))
    a, v = eng.make_anchors([(80, 80), (40, 40), (20, 20)])
    assert np.array_equal(v.astype(bool), valid[0, :, 0].numpy())
    ok = v.astype(bool)
    assert np.allclose(a[ok], anchors[0].numpy()[ok], atol=1e-6)
    assert np.all(a[~ok] == np.finfo(np.float32).max)
    assert int((~ok).sum()) == 316          # the border cells of the stride-8 level


def test_bf16_bits_and_bn_fold():
    x = torch.randn(1000)
    assert np.array_equal(eng.to_bf16_bits(x.numpy()), x.to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16))
    conv = torch.nn.Conv2d(4, 6, 3, bias=False)
    bn = torch.nn.BatchNorm2d(6, eps=1e-5).eval()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5); bn.bias.normal_(); bn.running_mean.normal_(); bn.running_var.uniform_(0.5, 1.5)
    w, b = eng.fold_bn(conv.weight.detach().numpy(), bn.weight.detach().numpy(), bn.bias.detach().numpy(),
                       bn.running_mean.numpy(), bn.running_var.numpy(), 1e-5)
    xin = torch.randn(1, 4, 8, 8)
    with torch.no_grad():
        ref = bn(conv(xin))
        got = torch.nn.functional.conv2d(xin, torch.from_numpy(w), torch.from_numpy(b))
    assert torch.allclose(ref, got, atol=1e-5)


def test_rtdetr_detector_fails_loudly_without_gpu():
    """No CPU fallback: without an sm_100 device load_model returns False (reference convention) and
    detect returns []."""
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from telescope_cam_detection_b200.rtdetr_detector import RTDETRDetector

    det = RTDETRDetector(state_dict={"x": np.zeros(1, np.float32)})
    assert det.load_model() is False
    assert det.detect(np.zeros((8, 8, 3), np.uint8)) == []
    assert det.detect_batch([]) == []


def test_upstream_key_map_round_trip():
    """rtdetr_keys maps an upstream-named (rtdetrv2_pytorch) state dict onto exactly the keys the program builder
    reads, values untouched (q/k/v split of in_proj included).  Completeness check only: the upstream code and
    checkpoints are absent here."""
    from oracle import rtdetr as oracle
    from telescope_cam_detection_b200 import rtdetr_keys

    sd = oracle.state_dict_numpy(oracle.make_model(0))
    want = {k: v for k, v in sd.items() if not (k.startswith("class_embed.") or k.startswith("bbox_embed.")
                                                or k.startswith("model.denoising_class_embed"))}
    up = rtdetr_keys.to_upstream(sd)
    assert rtdetr_keys.is_upstream(up) and not rtdetr_keys.is_upstream(sd)
    assert "encoder.encoder.0.layers.0.self_attn.in_proj_weight" in up
    assert up["decoder.decoder.layers.0.self_attn.in_proj_weight"].shape == (768, 256)
    assert "backbone.res_layers.1.blocks.0.short.conv.conv.weight" in up and "backbone.res_layers.0.blocks.0.short.conv.weight" in up
    back = rtdetr_keys.from_upstream(up)
    assert set(back) == set(want)
    for k in want:
        assert np.array_equal(back[k], want[k]), k
    with pytest.raises(KeyError):
        rtdetr_keys.from_upstream({"backbone.unknown.weight": np.zeros(1, np.float32)})
