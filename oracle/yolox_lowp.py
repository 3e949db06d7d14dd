"""ORACLE (test infrastructure only) -- the fp32 YOLOX-S restatement with the CUDA build's
ROUNDING POINTS emulated on the CPU.

The CUDA path stores every activation and every (BN-folded) conv weight as bf16 and
accumulates in fp32 (csrc/conv_umma.cuh).  ``lowp_model`` returns a copy of the fp32 oracle
(oracle/yolox_s.py) that rounds at exactly those points -- network input, folded conv weights,
every BaseConv output (after SiLU), every bottleneck sum -- and is otherwise fp32 PyTorch.

Two uses in tests/: (1) a tight check of the kernels that is independent of how well the
synthetic weights are conditioned (GPU vs this model differ only by fp32 summation order and
the fast exp in SiLU); (2) the statement of how far *any* bf16 implementation sits from the
fp32 reference on a given set of weights (precision section).
PARITY UNPINNED like the model it wraps.
"""
import copy

import torch

from .yolox_s import BaseConv, Bottleneck


def lowp_model(model, dtype=torch.bfloat16):
    m = copy.deepcopy(model).eval()

    def q(t):
        return t.to(dtype).float()

    with torch.no_grad():
        # the residual add happens in the conv epilogue in fp32, BEFORE the one rounding to bf16
        unrounded = {id(mod.conv2) for mod in m.modules() if isinstance(mod, Bottleneck) and mod.use_add}
        for mod in m.modules():
            if isinstance(mod, BaseConv):
                bn = mod.bn
                # same arithmetic as telescope_cam_detection_b200.weights.fold_bn (float64, then f32)
                s = bn.weight.double() / torch.sqrt(bn.running_var.double() + bn.eps)
                w = (mod.conv.weight.double() * s[:, None, None, None]).float()
                b = (bn.bias.double() - bn.running_mean.double() * s).float()
                mod.conv.weight.data = q(w)
                mod.conv.bias = torch.nn.Parameter(b)
                mod.bn = torch.nn.Identity()
                if id(mod) not in unrounded:
                    mod.register_forward_hook(lambda mo, i, o: q(o))
            if isinstance(mod, Bottleneck) and mod.use_add:
                mod.register_forward_hook(lambda mo, i, o: q(o))
        for l in list(m.head.cls_preds) + list(m.head.reg_preds) + list(m.head.obj_preds):
            l.weight.data = q(l.weight.data)
    inner = m.forward

    def forward(x):
        return inner(q(x))

    m.forward = forward
    return m
