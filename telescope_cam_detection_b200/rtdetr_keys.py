"""State-dict key map: upstream lyuwenyu/RT-DETR `rtdetrv2_pytorch` (what the reference loads,
src/rtdetr_detector.py:139-172: `checkpoint['ema']['module']` or `checkpoint['model']`) -> the names the
program builder (rtdetr_engine.py) reads, which are the `transformers` RT-DETRv2 names the oracle uses.

The upstream submodule is EMPTY under /root/reference and no checkpoint can be downloaded here, so this map is
written from the upstream module structure (PResNet `conv1.conv1_k` / `res_layers.S.blocks.B.branch2{a,b,c}` /
`short`, HybridEncoder `input_proj` / `encoder.L.layers.0` / `lateral_convs` / `fpn_blocks` / `downsample_convs`
/ `pan_blocks`, RTDETRTransformerv2 `input_proj` / `decoder.layers.L` / `*_head`) and is checked only for
COMPLETENESS: a synthetic upstream-named dict maps onto exactly the key set the builder consumes
(tests/test_oracle_rtdetr.py).  It has not met a real checkpoint -- DESIGN.md section 7a says so.
"""
import re

import numpy as np

_BN = {"weight": "weight", "bias": "bias", "running_mean": "running_mean", "running_var": "running_var"}


def is_upstream(sd):
    return any(k.startswith("backbone.") or k.startswith("decoder.decoder.") for k in sd)


def _split_qkv(out, dst, w_or_b, arr):
    d = arr.shape[0] // 3
    for i, n in enumerate(("q_proj", "k_proj", "v_proj")):
        out["%s.%s.%s" % (dst, n, w_or_b)] = arr[i * d:(i + 1) * d]


def _conv_norm(out, k, v, src_prefix, dst_conv, dst_norm):
    """`<src_prefix>.conv.weight` / `<src_prefix>.norm.<bn>` -> dst names; returns True when k matched."""
    if not k.startswith(src_prefix + "."):
        return False
    rest = k[len(src_prefix) + 1:]
    if rest == "conv.weight":
        out[dst_conv + ".weight"] = v
        return True
    if rest.startswith("norm.") and rest[5:] in _BN:
        out[dst_norm + "." + rest[5:]] = v
        return True
    return rest == "norm.num_batches_tracked"


def from_upstream(sd):
    """-> dict keyed like the transformers model.  Unknown keys raise (a silent drop would be a wrong model)."""
    out = {}
    for k, v in sd.items():
        v = np.asarray(v)
        if k.endswith("num_batches_tracked") or k in ("decoder.anchors", "decoder.valid_mask") or \
                k.startswith("decoder.denoising_class_embed"):
            continue
        # ---- PResNet-vd
        m = re.match(r"backbone\.conv1\.conv1_(\d)\.(conv\.weight|norm\.(\w+))$", k)
        if m:
            i = int(m.group(1)) - 1
            base = "model.backbone.model.embedder.embedder.%d." % i
            out[base + ("convolution.weight" if m.group(2) == "conv.weight" else "normalization." + m.group(3))] = v
            continue
        m = re.match(r"backbone\.res_layers\.(\d)\.blocks\.(\d+)\.(branch2[abc]|short(?:\.conv)?)\.(conv\.weight|norm\.(\w+))$", k)
        if m:
            s, b, part = int(m.group(1)), int(m.group(2)), m.group(3)
            base = "model.backbone.model.encoder.stages.%d.layers.%d." % (s, b)
            if part.startswith("branch2"):
                base += "layer.%d." % "abc".index(part[-1])
            elif part == "short":          # plain 1x1 ConvNormLayer (stage 0)
                base += "shortcut."
            else:                          # variant d: Sequential(pool, conv) -> index 1
                base += "shortcut.1."
            out[base + ("convolution.weight" if m.group(4) == "conv.weight" else "normalization." + m.group(5))] = v
            continue
        # ---- input projections (v2: .conv/.norm, v1: .0/.1)
        m = re.match(r"(encoder|decoder)\.input_proj\.(\d)\.(conv|0)\.weight$", k)
        if m:
            out["model.%s_input_proj.%s.0.weight" % (m.group(1), m.group(2))] = v
            continue
        m = re.match(r"(encoder|decoder)\.input_proj\.(\d)\.(norm|1)\.(\w+)$", k)
        if m:
            out["model.%s_input_proj.%s.1.%s" % (m.group(1), m.group(2), m.group(4))] = v
            continue
        # ---- AIFI / decoder transformer layers
        m = re.match(r"encoder\.encoder\.(\d)\.layers\.(\d)\.(.+)$", k)
        dst = None
        if m:
            dst, rest = "model.encoder.aifi.%s.layers.%s" % (m.group(1), m.group(2)), m.group(3)
            norms = {"norm1": "self_attn_layer_norm", "norm2": "final_layer_norm"}
        else:
            m = re.match(r"decoder\.decoder\.layers\.(\d)\.(.+)$", k)
            if m:
                dst, rest = "model.decoder.layers.%s" % m.group(1), m.group(2)
                norms = {"norm1": "self_attn_layer_norm", "norm2": "encoder_attn_layer_norm", "norm3": "final_layer_norm"}
        if dst is not None:
            head, _, leaf = rest.rpartition(".")
            if rest in ("self_attn.in_proj_weight", "self_attn.in_proj_bias"):
                _split_qkv(out, dst + ".self_attn", "weight" if rest.endswith("weight") else "bias", v)
            elif head == "self_attn.out_proj":
                out["%s.self_attn.o_proj.%s" % (dst, leaf)] = v
            elif head in ("linear1", "linear2"):
                out["%s.mlp.fc%s.%s" % (dst, head[-1], leaf)] = v
            elif head in norms:
                out["%s.%s.%s" % (dst, norms[head], leaf)] = v
            elif rest == "cross_attn.num_points_scale":
                out[dst + ".encoder_attn.n_points_scale"] = v
            elif head.startswith("cross_attn."):
                out["%s.encoder_attn.%s.%s" % (dst, head[len("cross_attn."):], leaf)] = v
            else:
                raise KeyError("rtdetr key map: unexpected transformer-layer key %r" % k)
            continue
        # ---- CCFM (same names under model.encoder.)
        if re.match(r"encoder\.(lateral_convs|fpn_blocks|downsample_convs|pan_blocks)\.", k):
            out["model." + k] = v
            continue
        # ---- decoder heads
        m = re.match(r"decoder\.enc_output\.(proj|norm|0|1)\.(\w+)$", k)
        if m:
            out["model.enc_output.%d.%s" % (0 if m.group(1) in ("proj", "0") else 1, m.group(2))] = v
            continue
        m = re.match(r"decoder\.(enc_score_head|enc_bbox_head\.layers\.\d|query_pos_head\.layers\.\d)\.(\w+)$", k)
        if m:
            pre = "model.decoder." if m.group(1).startswith("query_pos_head") else "model."
            out[pre + m.group(1) + "." + m.group(2)] = v
            continue
        m = re.match(r"decoder\.dec_score_head\.(\d)\.(\w+)$", k)
        if m:
            out["model.decoder.class_embed.%s.%s" % (m.group(1), m.group(2))] = v
            continue
        m = re.match(r"decoder\.dec_bbox_head\.(\d)\.layers\.(\d)\.(\w+)$", k)
        if m:
            out["model.decoder.bbox_embed.%s.layers.%s.%s" % (m.group(1), m.group(2), m.group(3))] = v
            continue
        raise KeyError("rtdetr key map: unexpected upstream key %r" % k)
    return out


def to_upstream(sd):
    """Inverse of from_upstream for the keys the builder reads (test helper: fabricates an upstream-named dict)."""
    out = {}
    qkv = {}
    for k, v in sd.items():
        v = np.asarray(v)
        if k.endswith("num_batches_tracked") or k.startswith("class_embed.") or k.startswith("bbox_embed.") or \
                k.startswith("model.denoising_class_embed"):
            continue
        m = re.match(r"model\.backbone\.model\.embedder\.embedder\.(\d)\.(convolution\.weight|normalization\.(\w+))$", k)
        if m:
            out["backbone.conv1.conv1_%d.%s" % (int(m.group(1)) + 1,
                                                "conv.weight" if m.group(3) is None else "norm." + m.group(3))] = v
            continue
        m = re.match(r"model\.backbone\.model\.encoder\.stages\.(\d)\.layers\.(\d+)\.(layer\.(\d)|shortcut(?:\.1)?)\.(convolution\.weight|normalization\.(\w+))$", k)
        if m:
            part = "branch2" + "abc"[int(m.group(4))] if m.group(4) is not None else ("short" if m.group(3) == "shortcut" else "short.conv")
            out["backbone.res_layers.%s.blocks.%s.%s.%s" % (m.group(1), m.group(2), part,
                                                           "conv.weight" if m.group(6) is None else "norm." + m.group(6))] = v
            continue
        m = re.match(r"model\.(encoder|decoder)_input_proj\.(\d)\.(\d)\.(\w+)$", k)
        if m:
            out["%s.input_proj.%s.%s.%s" % (m.group(1), m.group(2), "conv" if m.group(3) == "0" else "norm", m.group(4))] = v
            continue
        m = re.match(r"model\.encoder\.aifi\.(\d)\.layers\.(\d)\.(.+)$", k)
        src = None
        if m:
            src, rest = "encoder.encoder.%s.layers.%s" % (m.group(1), m.group(2)), m.group(3)
            norms = {"self_attn_layer_norm": "norm1", "final_layer_norm": "norm2"}
        else:
            m = re.match(r"model\.decoder\.layers\.(\d)\.(.+)$", k)
            if m:
                src, rest = "decoder.decoder.layers.%s" % m.group(1), m.group(2)
                norms = {"self_attn_layer_norm": "norm1", "encoder_attn_layer_norm": "norm2", "final_layer_norm": "norm3"}
        if src is not None:
            head, _, leaf = rest.rpartition(".")
            if head in ("self_attn.q_proj", "self_attn.k_proj", "self_attn.v_proj"):
                qkv.setdefault((src, leaf), {})[head[-6]] = v
            elif head == "self_attn.o_proj":
                out["%s.self_attn.out_proj.%s" % (src, leaf)] = v
            elif head in ("mlp.fc1", "mlp.fc2"):
                out["%s.linear%s.%s" % (src, head[-1], leaf)] = v
            elif head in norms:
                out["%s.%s.%s" % (src, norms[head], leaf)] = v
            elif rest == "encoder_attn.n_points_scale":
                out[src + ".cross_attn.num_points_scale"] = v
            elif head.startswith("encoder_attn."):
                out["%s.cross_attn.%s.%s" % (src, head[len("encoder_attn."):], leaf)] = v
            else:
                raise KeyError(k)
            continue
        if re.match(r"model\.encoder\.(lateral_convs|fpn_blocks|downsample_convs|pan_blocks)\.", k):
            out[k[len("model."):]] = v
            continue
        m = re.match(r"model\.enc_output\.(\d)\.(\w+)$", k)
        if m:
            out["decoder.enc_output.%s.%s" % ("proj" if m.group(1) == "0" else "norm", m.group(2))] = v
            continue
        m = re.match(r"model\.(enc_score_head|enc_bbox_head\.layers\.\d)\.(\w+)$", k)
        if m:
            out["decoder.%s.%s" % (m.group(1), m.group(2))] = v
            continue
        m = re.match(r"model\.decoder\.query_pos_head\.layers\.(\d)\.(\w+)$", k)
        if m:
            out["decoder.query_pos_head.layers.%s.%s" % (m.group(1), m.group(2))] = v
            continue
        m = re.match(r"model\.decoder\.class_embed\.(\d)\.(\w+)$", k)
        if m:
            out["decoder.dec_score_head.%s.%s" % (m.group(1), m.group(2))] = v
            continue
        m = re.match(r"model\.decoder\.bbox_embed\.(\d)\.layers\.(\d)\.(\w+)$", k)
        if m:
            out["decoder.dec_bbox_head.%s.layers.%s.%s" % (m.group(1), m.group(2), m.group(3))] = v
            continue
        raise KeyError("to_upstream: unexpected key %r" % k)
    for (src, leaf), parts in qkv.items():
        out["%s.self_attn.in_proj_%s" % (src, leaf)] = np.concatenate([parts["q"], parts["k"], parts["v"]], 0)
    return out
