"""ORACLE (test infrastructure only) -- seeded synthetic YOLOX-S weights.

The real yolox_s.pth cannot be downloaded here (no network; BASELINE.json asks for
synthetic data / random-init weights).  Upstream's default init gives zero detections
on random weights, so (SURVEY.md section 8d "Weights for parity"):
  * conv weights: torch default init under torch.manual_seed(seed)
  * BN: gamma~U(0.5,1.5), beta~N(0,0.1); running stats CALIBRATED on one seeded noise
    frame (trained-network-like activation scale), then perturbed a little
  * head obj/cls biases shifted so a few % of the 8400 anchors pass conf 0.25; w/h bias > 0 so
    neighbouring boxes overlap and NMS suppresses a good share of them
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
def make_model(seed=0, calibrate=True, obj_bias=-5.5, cls_bias=-3.0, wh_bias=1.5,
               obj_std=1.5, cls_std=1.0, reg_std=0.5,
               gamma=(0.5, 1.5), beta=(0.0, 0.1)):
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
        x = preprocess_ref(calib_frame())
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
    for conv in model.head.cls_preds:
        conv.bias.fill_(cls_bias)
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
