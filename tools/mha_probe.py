"""Rotary attention at the Stage-2 classifier's shape (64 crops x 16 heads x 577 tokens x 64): device time per launch,
max error against torch fp32 on the same bf16 operands.  TCD_MHA_UMMA=0 selects the warp-level kernel."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from telescope_cam_detection_b200.eva02_engine import rope_cos_sin  # noqa: E402
from telescope_cam_detection_b200.rtdetr_engine import Act, Program, to_bf16_bits  # noqa: E402


def rot(x):
    return torch.stack([-x[..., 1::2], x[..., ::2]], -1).reshape(x.shape)


def main():
    N, T, H = int(os.environ.get("N", 64)), int(os.environ.get("T", 577)), int(os.environ.get("H", 16))
    D = 64 * H
    torch.manual_seed(0)
    qkv = (torch.randn(N, T, 3 * D) * 1.5).to(torch.bfloat16).float()
    P = Program(0)
    QKV = Act(P.upload(to_bf16_bits(qkv.numpy())), N, 1, T, 3 * D)
    O = P.act(N, 1, T, D)
    G = int(round((T - 1) ** 0.5))
    tab = rope_cos_sin(G, 16)
    P.op("mha", "mha_rope", [QKV.ptr, QKV.ptr, QKV.ptr, O.ptr, P.upload(tab)],
         [3 * D, 0, 3 * D, D, 3 * D, 2 * D, D, N, T, H, 1], [64 ** -0.5])
    P.run(False)
    P.sync()
    ms = P.profile(10)[0]["ms"]
    got = torch.from_numpy(P.read_bf16(O))[:, 0]
    nchk = min(N, 2)
    q, k, v = [qkv[:nchk, :, i * D:(i + 1) * D].view(nchk, T, H, 64).transpose(1, 2) for i in range(3)]
    cos = torch.from_numpy(tab[..., 0]).repeat_interleave(2, -1)
    sin = torch.from_numpy(tab[..., 1]).repeat_interleave(2, -1)
    q = torch.cat([q[:, :, :1], q[:, :, 1:] * cos + rot(q[:, :, 1:]) * sin], 2).to(torch.bfloat16).float()
    k = torch.cat([k[:, :, :1], k[:, :, 1:] * cos + rot(k[:, :, 1:]) * sin], 2).to(torch.bfloat16).float()
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v).transpose(1, 2).reshape(nchk, T, D)
    fl = 4.0 * T * T * D * N
    print(json.dumps({"N": N, "T": T, "heads": H, "ms": ms, "tflops": fl / (ms * 1e-3) / 1e12,
                      "max_err": float((got[:nchk] - ref).abs().max()), "umma": os.environ.get("TCD_MHA_UMMA", "1")}))
    P.close()


if __name__ == "__main__":
    main()
