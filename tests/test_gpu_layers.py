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


def _stage_errors(model, blob, frames, pre_mode=0, max_batch=None):
    """-> list of (buffer, module path, err GPU vs fp32, err bf16-rounding oracle vs fp32, mean / max err GPU vs
    bf16-rounding oracle), every error relative to the stage's mean |activation|."""
    import torch
    from oracle.reference_detector import preprocess_cfg
    from oracle.yolox_lowp import lowp_model
    from telescope_cam_detection_b200.engine import DetectorEngine

    def taps_of(m, x):
        mods, taps = dict(m.named_modules()), {}
        hooks = [mods[path].register_forward_hook(lambda mo, i, o, p=path: taps.__setitem__(p, o.detach()))
                 for _, path, _ in STAGES]
        with torch.no_grad():
            m(x)
        for h in hooks:
            h.remove()
        return taps

    x = torch.cat([preprocess_cfg(f, (640, 640), "letterbox" if pre_mode else "stretch") for f in frames], 0)
    taps = taps_of(model, x)
    ltaps = taps_of(lowp_model(model), x)   # same taps on the oracle with the build's bf16 rounding points emulated
    eng = DetectorEngine(blob, input_size=(640, 640), max_batch=max_batch or len(frames), max_out=64, pre_mode=pre_mode)
    out = []
    try:
        eng.forward(frames)
        for buf, path, sl in STAGES:
            got = eng.read_activation(buf)  # [B,H,W,C] f32
            if sl is not None:
                got = got[..., sl[0]:sl[1]]
            ref = taps[path].permute(0, 2, 3, 1).numpy()
            assert got.shape == ref.shape, "%s: shape %s vs oracle %s" % (buf, got.shape, ref.shape)
            lref = ltaps[path].permute(0, 2, 3, 1).numpy()
            scale = float(np.abs(ref).mean()) + 1e-6
            out.append((buf, path, float(np.abs(got - ref).mean()) / scale, float(np.abs(lref - ref).mean()) / scale,
                        float(np.abs(got - lref).mean()) / scale, float(np.abs(got - lref).max()) / scale))
    finally:
        eng.close()
    return out


def _report(rows):
    return "\n".join("%-18s <- %-32s GPU vs fp32 %.4f | bf16-rounding oracle vs fp32 %.4f | GPU vs bf16-rounding "
                     "oracle mean %.5f max %.4f" % r for r in rows)


@pytest.mark.parametrize("B,pre_mode", [(2, 0), (8, 1)], ids=["B2-stretch", "B8-letterbox"])
def test_stage_activations_match_oracle(B, pre_mode):
    """B = 2 stretch: the reference's own preprocessing.  B = 8 letterbox: the configuration bench.py times -- the
    planner is batch-dependent (whole-SM stem plan, tile grouping, resident weights switch on with the tile count),
    so the plan that produces the headline number is the one compared here."""
    frames = [structured_frame(41), noise_frame(42)] + [structured_frame(43 + i) for i in range(B - 2)]
    rows = _stage_errors(oracle_model(0), oracle_blob(0), frames, pre_mode=pre_mode)
    rep = _report(rows)
    print(rep)
    # bf16 rounding is ~0.2-0.4 % per stored activation; 28 stages deep stays well below 3 %
    assert max(r[2] for r in rows) <= 0.03, rep
    # Against the rounding-point emulation the first stages are bit-identical (measured 0.00000
    # through stem/dark2.0).  A first bf16 rounding flip (fp32 summation order) then perturbs its
    # whole receptive field by ~1e-3 relative, which flips a large share of the NEXT roundings:
    # the two bf16 pipelines decorrelate and settle at the bf16 noise level itself (~0.7 %).
    assert max(r[4] for r in rows[:4]) <= 1e-3, rep
    assert max(r[4] for r in rows) <= 0.02, rep


def test_stage_activations_second_weight_set_nonlinear_silu():
    """SURVEY 8(d)'s BN recipe (gamma ~ U(0.5,1.5), beta ~ N(0,0.1)): pre-activations of order 1-4, so the SiLU
    epilogue works well outside its linear range (the default set keeps gamma at 0.05-0.15, oracle/synth_weights.py).
    A random-weight stack at this gain is chaotic -- the CPU emulation of the build's rounding points is itself
    ~0.5 % from fp32 at the stem and ~30 % at the head (measured) -- so end-to-end tolerances
    do not apply.  What is asserted: bit-nearness to the rounding-point emulation through the first stages, and at
    EVERY depth the GPU is no further from fp32 than the emulation is (x1.5 + 0.2 %): the kernels add no error of
    their own where the activation function is curved."""
    from oracle import synth_weights
    from telescope_cam_detection_b200 import weights
    model = synth_weights.make_model(seed=0, gamma=(0.5, 1.5), beta=(0.0, 0.1))
    frames = [structured_frame(41), noise_frame(42)]
    rows = _stage_errors(model, weights.pack_yolox(model.state_dict()), frames)
    rep = _report(rows)
    print(rep)
    assert max(r[4] for r in rows[:2]) <= 2e-3, rep
    for r in rows:
        assert r[2] <= 1.5 * r[3] + 0.002, "%s: GPU %.4f from fp32, bf16 emulation %.4f\n%s" % (r[0], r[2], r[3], rep)
