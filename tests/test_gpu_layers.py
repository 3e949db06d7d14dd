"""Per-stage activation parity (golden set G2 of SURVEY.md section 8c): every named buffer of
the CUDA layer plan against the matching module output of the CPU oracle on the same frames.
Tolerance: bf16 storage of every activation (rel 2^-8 per layer, compounding) -> the bound is
stated relative to the stage's own dynamic range and loosens with depth."""
import numpy as np
import pytest

from util import structured_frame, noise_frame, oracle_model, oracle_blob

pytestmark = pytest.mark.gpu

# plan buffer name -> (oracle module path, channel slice of the plan buffer or None)
STAGES = [
    ("stem", "backbone.backbone.stem", None),
    ("dark2.0", "backbone.backbone.dark2.0", None),
    ("dark2", "backbone.backbone.dark2", None),
    ("dark3.0", "backbone.backbone.dark3.0", None),
    ("cat_p3", "backbone.backbone.dark3", (128, 256)),
    ("dark4.0", "backbone.backbone.dark4.0", None),
    ("cat_p4", "backbone.backbone.dark4", (256, 512)),
    ("dark5.0", "backbone.backbone.dark5.0", None),
    ("spp_cat", "backbone.backbone.dark5.1.conv1", (0, 256)),
    ("dark5.1", "backbone.backbone.dark5.1", None),
    ("dark5", "backbone.backbone.dark5", None),
    ("cat_n4", "backbone.lateral_conv0", (256, 512)),
    ("f_out0", "backbone.C3_p4", None),
    ("cat_n3", "backbone.reduce_conv1", (128, 256)),
    ("pan_out2", "backbone.C3_p3", None),
    ("cat_n3", "backbone.bu_conv2", (0, 128)),
    ("pan_out1", "backbone.C3_n3", None),
    ("cat_n4", "backbone.bu_conv1", (0, 256)),
    ("pan_out0", "backbone.C3_n4", None),
    ("head.stem.0", "head.stems.0", None),
    ("head.feat.0", "head.cls_convs.0", (0, 128)),
    ("head.feat.0", "head.reg_convs.0", (128, 256)),
    ("head.stem.1", "head.stems.1", None),
    ("head.feat.1", "head.cls_convs.1", (0, 128)),
    ("head.feat.1", "head.reg_convs.1", (128, 256)),
    ("head.stem.2", "head.stems.2", None),
    ("head.feat.2", "head.cls_convs.2", (0, 128)),
    ("head.feat.2", "head.reg_convs.2", (128, 256)),
]


def test_stage_activations_match_oracle():
    import torch
    from oracle.synth_weights import preprocess_ref
    from telescope_cam_detection_b200.engine import DetectorEngine

    frames = [structured_frame(41), noise_frame(42)]
    model = oracle_model(0)
    mods = dict(model.named_modules())
    taps = {}
    hooks = [mods[path].register_forward_hook(lambda m, i, o, p=path: taps.__setitem__(p, o.detach()))
             for _, path, _ in STAGES]
    x = torch.cat([preprocess_ref(f) for f in frames], 0)
    with torch.no_grad():
        model(x)
    for h in hooks:
        h.remove()
    # same taps on the oracle with the build's bf16 rounding points emulated
    from oracle.yolox_lowp import lowp_model
    low = lowp_model(model)
    lmods = dict(low.named_modules())
    ltaps = {}
    for _, path, _ in STAGES:
        lmods[path].register_forward_hook(lambda m, i, o, p=path: ltaps.__setitem__(p, o.detach()))
    with torch.no_grad():
        low(x)

    eng = DetectorEngine(oracle_blob(0), input_size=(640, 640), max_batch=2, max_out=64)
    try:
        eng.forward(frames)
        report, worst, worst_low, first_low = [], 0.0, 0.0, 0.0
        for depth, (buf, path, sl) in enumerate(STAGES):
            got = eng.read_activation(buf)  # [B,H,W,C] f32
            if sl is not None:
                got = got[..., sl[0]:sl[1]]
            ref = taps[path].permute(0, 2, 3, 1).numpy()
            assert got.shape == ref.shape, "%s: shape %s vs oracle %s" % (buf, got.shape, ref.shape)
            scale = float(np.abs(ref).mean()) + 1e-6
            err = float(np.abs(got - ref).mean()) / scale
            lref = ltaps[path].permute(0, 2, 3, 1).numpy()
            lerr = float(np.abs(got - lref).mean()) / scale
            lmax = float(np.abs(got - lref).max()) / scale
            report.append("%-18s <- %-32s vs fp32 %.4f | vs bf16-rounding oracle mean %.5f max %.4f" % (
                buf, path, err, lerr, lmax))
            worst = max(worst, err)
            worst_low = max(worst_low, lerr)
            if depth < 4:
                first_low = max(first_low, lerr)
        print("\n".join(report))
        # bf16 rounding is ~0.2-0.4 % per stored activation; 28 stages deep stays well below 3 %
        assert worst <= 0.03, "\n".join(report)
        # Against the rounding-point emulation the first stages are bit-identical (measured 0.00000
        # through stem/dark2.0).  A first bf16 rounding flip (fp32 summation order) then perturbs its
        # whole receptive field by ~1e-3 relative, which flips a large share of the NEXT roundings:
        # the two bf16 pipelines decorrelate and settle at the bf16 noise level itself (~0.7 %).
        assert first_low <= 1e-3, "\n".join(report)
        assert worst_low <= 0.02, "\n".join(report)
    finally:
        eng.close()
