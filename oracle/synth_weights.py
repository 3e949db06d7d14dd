"""ORACLE (test infrastructure only) -- seeded synthetic YOLOX-S weights.

The real yolox_s.pth cannot be downloaded here (no network; BASELINE.json asks for
synthetic data / random-init weights).  Upstream's default init gives zero detections
on random weights, so (SURVEY.md section 8d "Weights for parity"):
  * conv weights: torch default init under torch.manual_seed(seed)
  * BN: running stats CALIBRATED on two seeded frames (one structured, one white noise), then
    perturbed a little; gamma~U(0.05,0.15), beta~N(0,0.02).
    WHY SO SMALL A GAMMA: a random-weight BN+SiLU stack is chaotic -- every layer multiplies a
    relative perturbation by ~1.2 (mean-field correlation slope of SiLU at unit pre-activation
    variance), so the 2^-9 rounding of a 16-bit activation format grows ~100x over the 54
    roundings of the longest path and ANY bf16 implementation (including a CPU emulation that
    rounds fp32 results to bf16 at the same points, oracle/yolox_lowp.py) ends 0.5 away from
    the fp32 scores.  Trained networks are not chaotic.  Small gamma keeps SiLU in its
    near-linear range around 0, the slope drops to ~1.0 and the rounding noise only
    accumulates in quadrature (~0.8 % at the head features, the floor for 54 bf16 roundings).
  * prediction convs: rescaled so obj/cls logits have std 0.4 and box logits std 0.05 on the
    calibration frames: 0.8 % feature noise x 0.4 x 4.5 sigma x sigmoid'(<=0.25) keeps bf16
    scores within ~7e-3 of fp32 (north_star tolerance 1e-2) and boxes within ~0.35 px (0.5)
  * head obj/cls biases shifted so 4-7 % of the 8400 anchors pass conf 0.25; a few classes get
    a larger bias so neighbouring anchors share classes; w/h bias 1.3 (boxes 3.7 strides wide)
    puts adjacent-anchor IoU at ~0.57 and diagonal IoU at ~0.36, away from nms_thre 0.45, so
    class-aware NMS removes ~20 % of the candidates
"""
import numpy as np
import torch

from .yolox_s import get_exp, BaseConv


def calib_frame(seed=1234, h=1080, w=1920):
    rng = np.random.default_rng(seed)
    # low-pass noise + blocks: something with structure at several scales
    base = rng.integers(0, 255, (h // 8, w // 8, 3), dtype=np.uint8)
    img = np.repeat(np.repeat(base, 8, 0), 8, 1).astype(np.int16)
    img += rng.integers(-20, 20, (h, w, 3), dtype=np.int16)
    return np.clip(img, 0, 255).astype(np.uint8)


def preprocess_ref(img, size=(640, 640)):
    """Same ops as /root/reference/src/yolox_detector.py:199-218 (numpy branch, CPU)."""
    t = torch.from_numpy(img).permute(2, 0, 1).unsqueeze(0).float()
    if t.shape[2:] != size:
        t = torch.nn.functional.interpolate(t, size=size, mode="bilinear", align_corners=False)
    return t


@torch.no_grad()
def make_model(seed=0, calibrate=True, obj_bias=-1.0, cls_bias=-2.4, wh_bias=1.3,
               obj_std=0.4, cls_std=0.4, reg_std=0.05, cls_boost=3.0,
               gamma=(0.05, 0.15), beta=(0.0, 0.02)):
    torch.manual_seed(seed)
    model = get_exp(None, "yolox-s").get_model()
    model.eval()
    g = torch.Generator().manual_seed(seed + 1)
    bns = [m.bn for m in model.modules() if isinstance(m, BaseConv)]
    for bn in bns:
        c = bn.num_features
        bn.weight.copy_(torch.rand(c, generator=g) * (gamma[1] - gamma[0]) + gamma[0])
        bn.bias.copy_(torch.randn(c, generator=g) * beta[1] + beta[0])
    if calibrate:
        # one structured and one white-noise frame, so both kinds of test input are in-distribution
        noise = np.random.default_rng(4321).integers(0, 255, (1080, 1920, 3), dtype=np.uint8)
        x = torch.cat([preprocess_ref(calib_frame()), preprocess_ref(noise)], 0)
        # sequential calibration: each BN sees inputs produced by already-calibrated layers
        hooks = []

        def hook(mod, inp, out):
            v = inp[0]
            mean = v.mean(dim=(0, 2, 3))
            var = v.var(dim=(0, 2, 3), unbiased=False)
            mod.running_mean.copy_(mean)
            mod.running_var.copy_(var.clamp_min(1e-4))
            return torch.nn.functional.batch_norm(
                v, mod.running_mean, mod.running_var, mod.weight, mod.bias, False, 0.0, mod.eps)

        for bn in bns:
            hooks.append(bn.register_forward_hook(hook))
        model(x)
        for h in hooks:
            h.remove()
        for bn in bns:
            c = bn.num_features
            bn.running_mean.mul_(1 + 0.05 * torch.randn(c, generator=g))
            bn.running_var.mul_(torch.rand(c, generator=g) * 0.4 + 0.8)
    else:
        for bn in bns:
            c = bn.num_features
            bn.running_mean.copy_(torch.randn(c, generator=g) * 0.1)
            bn.running_var.copy_(torch.rand(c, generator=g) + 0.5)
    # prediction convs have no BN: rescale so logits are O(1) on the calibration frame
    if calibrate:
        feats = model.backbone(x)
        for k, f in enumerate(feats):
            f = model.head.stems[k](f)
            cf = model.head.cls_convs[k](f)
            rf = model.head.reg_convs[k](f)
            for conv, feat, std in ((model.head.cls_preds[k], cf, cls_std),
                                    (model.head.obj_preds[k], rf, obj_std),
                                    (model.head.reg_preds[k], rf, reg_std)):
                conv.bias.zero_()
                s = conv(feat).std().clamp_min(1e-6)
                conv.weight.mul_(std / s)
    for conv in model.head.reg_preds:
        conv.bias.zero_()
        conv.bias[2:4] = wh_bias  # boxes ~exp(1.5) strides wide: neighbours overlap, NMS has work to do
    for conv in model.head.obj_preds:
        conv.bias.fill_(obj_bias)
    # a handful of classes (the reference's wildlife set among them) get a head start, so that
    # neighbouring anchors often agree on the class and class-aware NMS has overlaps to resolve
    boost = torch.zeros(model.head.num_classes)
    boost[[0, 14, 15, 16, 21, 2, 56]] = torch.tensor([0.6, 0.5, 0.7, 0.6, 0.5, 0.4, 0.4]) * cls_boost
    for conv in model.head.cls_preds:
        conv.bias.copy_(cls_bias + boost)
    return model


def bn_stats(model):
    out = {}
    for name, m in model.named_modules():
        if isinstance(m, BaseConv):
            out[name + ".mean"] = m.bn.running_mean.numpy().copy()
            out[name + ".var"] = m.bn.running_var.numpy().copy()
    return out


def load_bn_stats(model, stats):
    with torch.no_grad():
        for name, m in model.named_modules():
            if isinstance(m, BaseConv):
                m.bn.running_mean.copy_(torch.from_numpy(stats[name + ".mean"]))
                m.bn.running_var.copy_(torch.from_numpy(stats[name + ".var"]))
    return model
