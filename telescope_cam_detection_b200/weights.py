"""Weight packer: upstream YOLOX state dict -> libtcd blob (BN folded, fp32, named entries).

Reads the same keys the reference loads with ``ckpt["model"]`` (src/yolox_detector.py:136-137;
key layout in SURVEY.md App. A).  Fails loudly on any missing or unexpected key.

Blob layout (little endian), parsed by csrc/yolox_net.cu::WeightBlob::parse:
    char magic[8] = "TCDW0001"; u32 n_entries; u32 num_classes;
    per entry: char name[96]; i32 cout, cin, k, reserved; f32 w[cout*cin*k*k]; f32 b[cout]
"""
import struct

import numpy as np

BN_EPS = 1e-3  # upstream yolox/exp/yolox_base.py::get_model sets eps=1e-3 on every BatchNorm2d


def _np(t):
    if hasattr(t, "detach"):
        t = t.detach().cpu().numpy()
    return np.asarray(t)


def fold_bn(w, gamma, beta, mean, var, eps=BN_EPS):
    """conv(bias=False) -> BN(eval)  ==  conv with w' = w*g/sqrt(var+eps), b' = beta - mean*g/sqrt(var+eps)."""
    w = _np(w).astype(np.float64)
    s = _np(gamma).astype(np.float64) / np.sqrt(_np(var).astype(np.float64) + eps)
    wf = w * s[:, None, None, None]
    bf = _np(beta).astype(np.float64) - _np(mean).astype(np.float64) * s
    return wf.astype(np.float32), bf.astype(np.float32)


def yolox_conv_names(state_dict):
    """BaseConv prefixes (have .conv.weight + .bn.*) and plain Conv2d prefixes (have .weight + .bias)."""
    base, plain = [], []
    for k in state_dict:
        if k.endswith(".conv.weight"):
            base.append(k[: -len(".conv.weight")])
        elif k.endswith(".weight") and not k.endswith(".bn.weight") and (k[: -len(".weight")] + ".bias") in state_dict:
            p = k[: -len(".weight")]
            if not p.endswith(".bn"):
                plain.append(p)
    return base, plain


def pack_yolox(state_dict, eps=BN_EPS):
    """Returns bytes for tcd_create_detector."""
    sd = dict(state_dict)
    base, plain = yolox_conv_names(sd)
    used = set()
    entries = []
    for p in base:
        keys = [p + ".conv.weight", p + ".bn.weight", p + ".bn.bias", p + ".bn.running_mean", p + ".bn.running_var"]
        for k in keys:
            if k not in sd:
                raise KeyError("pack_yolox: missing key %s" % k)
        used.update(keys)
        used.add(p + ".bn.num_batches_tracked")
        w, b = fold_bn(*[sd[k] for k in keys], eps=eps)
        entries.append((p, w, b))
    num_classes = None
    for p in plain:
        w = _np(sd[p + ".weight"]).astype(np.float32)
        b = _np(sd[p + ".bias"]).astype(np.float32)
        used.update([p + ".weight", p + ".bias"])
        entries.append((p, w, b))
        if ".cls_preds." in p:
            num_classes = w.shape[0]
    extra = [k for k in sd if k not in used]
    if extra:
        raise KeyError("pack_yolox: unexpected keys in state dict: %s" % extra[:8])
    if num_classes is None:
        raise KeyError("pack_yolox: no head.cls_preds.* entries found")
    out = [b"TCDW0001", struct.pack("<II", len(entries), num_classes)]
    for name, w, b in entries:
        cout, cin, kh, kw = w.shape
        if kh != kw or kh not in (1, 3):
            raise ValueError("pack_yolox: %s has unsupported kernel %dx%d" % (name, kh, kw))
        nb = name.encode()
        if len(nb) > 95:
            raise ValueError("pack_yolox: name too long: %s" % name)
        out.append(nb.ljust(96, b"\0"))
        out.append(struct.pack("<iiii", cout, cin, kh, 0))
        out.append(np.ascontiguousarray(w, dtype="<f4").tobytes())
        out.append(np.ascontiguousarray(b, dtype="<f4").tobytes())
    return b"".join(out)
