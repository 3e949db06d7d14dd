"""CPU oracle of the Stage-2 classifier body -- TEST INFRASTRUCTURE ONLY.

The reference builds the body with `timm.create_model(model_name, num_classes=...)`
(src/species_classifier.py:258-262); the shipped configuration selects
`timm/eva02_large_patch14_clip_336.merged2b_ft_inat21` at 336 x 336 (config/config.yaml:279-281).
timm (requirements.txt:4 `timm>=1.0.0`, no exact pin) is NOT installed in this image and there is no
network, so this file restates the published architecture of timm's `Eva` model
(timm/models/eva.py, the `eva02_*_patch14_clip_*` registrations) in plain PyTorch fp32, with timm's
state-dict key names so that a real checkpoint would load unchanged:

  patch_embed.proj (Conv2d 3 -> D, kernel = stride = 14) -> [cls_token | patches] + pos_embed ->
  depth x EvaBlock:  x += attn(norm1(x));  x += mlp(norm2(x))
      attn:  fused qkv Linear (bias = [q_bias, 0, v_bias]), 2-D rotary embedding on q and k of the patch tokens
             (class token skipped; interleaved pairs, timm RotaryEmbeddingCat with in_pixels = False and
             ref_feat_shape = 16 x 16), softmax(q k^T / sqrt(64)) v, LayerNorm over the D attention
             channels (scale_attn_inner = True), proj Linear
      mlp:   SwiGLU: silu(fc1_g(x)) * fc1_x(x) -> LayerNorm(hidden) (scale_mlp = True) -> fc2,
             hidden = int(D * 8 / 3)
  -> norm (LayerNorm) -> class token -> head Linear (num_classes)
LayerNorm eps 1e-6 (timm's default norm layer), head_dim 64 for every registered size.

PARITY UNPINNED: nothing here could be checked against timm itself or against reference vectors (the
reference holds none for this path); the CUDA engine (telescope_cam_detection_b200/
eva02_engine.py) is held to THIS restatement.  `lowp=True` rounds to bf16 at the points where the CUDA
engine stores bf16, to tell implementation error from the precision floor (like oracle/yolox_lowp.py).

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
import math
from dataclasses import dataclass

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F


@dataclass
class Eva02Config:
    img_size: int = 336
    patch_size: int = 14
    embed_dim: int = 1024
    depth: int = 24
    num_heads: int = 16
    num_classes: int = 10000
    ref_grid: int = 16            # ref_feat_shape = (16, 16) = 224 / 14, timm eva02_*_clip_* registrations
    ln_eps: float = 1e-6
    qkv_fused: bool = True

    @property
    def hidden(self):
        return int(self.embed_dim * (4 * 2 / 3))

    @property
    def grid(self):
        return self.img_size // self.patch_size

    @property
    def head_dim(self):
        return self.embed_dim // self.num_heads


# the registered sizes (timm/models/eva.py): name fragment -> (embed_dim, depth, heads)
VARIANTS = {"tiny": (192, 12, 3), "small": (384, 12, 6), "base": (768, 12, 12), "large": (1024, 24, 16),
            "enormous": (1792, 64, 16)}


def config_for(model_name, img_size=None, num_classes=10000):
    name = model_name.lower()
    for frag, (d, depth, heads) in VARIANTS.items():
        if "eva02_" + frag in name:
            size = img_size
            if size is None:
                size = 336 if "336" in name else (448 if "448" in name else 224)
            return Eva02Config(img_size=size, embed_dim=d, depth=depth, num_heads=heads, num_classes=num_classes)
    raise ValueError("not an EVA02 model name: %s" % model_name)


def rope_table(grid, ref_grid, head_dim, temperature=10000.0):
    """(sin, cos), each [grid*grid, head_dim] float32: timm build_rotary_pos_embed(feat_shape=(grid, grid), dim=head_dim,
    in_pixels=False, ref_feat_shape=(ref_grid, ref_grid)): head_dim/4 bands 1 / T^(i / (head_dim/4)), positions
    arange(grid) / grid * ref_grid, row-major (y, x) grid, per position [y bands | x bands], every band twice."""
    nb = head_dim // 4
    bands = 1.0 / (temperature ** (torch.arange(0, nb, dtype=torch.float32) / nb))
    t = torch.arange(grid, dtype=torch.float32) / grid * ref_grid
    gy, gx = torch.meshgrid(t, t, indexing="ij")
    pos = torch.stack([gy, gx], -1).unsqueeze(-1) * bands          # [g, g, 2, nb]
    sin = pos.sin().reshape(grid * grid, -1).repeat_interleave(2, -1)
    cos = pos.cos().reshape(grid * grid, -1).repeat_interleave(2, -1)
    return sin.contiguous(), cos.contiguous()


def _rot(x):
    return torch.stack([-x[..., 1::2], x[..., ::2]], -1).reshape(x.shape)


def _bf(x):
    return x.to(torch.bfloat16).to(torch.float32)


class _Attn(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        D = cfg.embed_dim
        self.cfg = cfg
        if cfg.qkv_fused:
            self.qkv = nn.Linear(D, 3 * D, bias=False)
            self.q_bias = nn.Parameter(torch.zeros(D))
            self.v_bias = nn.Parameter(torch.zeros(D))
        else:
            self.q_proj = nn.Linear(D, D, bias=True)
            self.k_proj = nn.Linear(D, D, bias=False)
            self.v_proj = nn.Linear(D, D, bias=True)
        self.norm = nn.LayerNorm(D, eps=cfg.ln_eps)
        self.proj = nn.Linear(D, D)

    def forward(self, x, sin, cos, r):
        B, T, D = x.shape
        H, hd = self.cfg.num_heads, self.cfg.head_dim
        if self.cfg.qkv_fused:
            bias = torch.cat([self.q_bias, torch.zeros_like(self.q_bias), self.v_bias])
            qkv = r(F.linear(x, r(self.qkv.weight), bias)).reshape(B, T, 3, H, hd).permute(2, 0, 3, 1, 4)
            q, k, v = qkv[0], qkv[1], qkv[2]
        else:
            q = r(F.linear(x, r(self.q_proj.weight), self.q_proj.bias)).reshape(B, T, H, hd).transpose(1, 2)
            k = r(F.linear(x, r(self.k_proj.weight))).reshape(B, T, H, hd).transpose(1, 2)
            v = r(F.linear(x, r(self.v_proj.weight), self.v_proj.bias)).reshape(B, T, H, hd).transpose(1, 2)
        q = torch.cat([q[:, :, :1], q[:, :, 1:] * cos + _rot(q[:, :, 1:]) * sin], 2)
        k = torch.cat([k[:, :, :1], k[:, :, 1:] * cos + _rot(k[:, :, 1:]) * sin], 2)
        q, k = r(q), r(k)
        a = torch.softmax((q @ k.transpose(-1, -2)) * (hd ** -0.5), -1)
        o = r((a @ v).transpose(1, 2).reshape(B, T, D))
        o = r(self.norm(o))
        return F.linear(o, r(self.proj.weight), self.proj.bias)


class _SwiGLU(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        D, Hh = cfg.embed_dim, cfg.hidden
        self.fc1_g = nn.Linear(D, Hh)
        self.fc1_x = nn.Linear(D, Hh)
        self.norm = nn.LayerNorm(Hh, eps=cfg.ln_eps)
        self.fc2 = nn.Linear(Hh, D)

    def forward(self, x, r):
        g = r(F.linear(x, r(self.fc1_g.weight), self.fc1_g.bias))
        u = r(F.linear(x, r(self.fc1_x.weight), self.fc1_x.bias))
        h = r(self.norm(F.silu(g) * u))
        return F.linear(h, r(self.fc2.weight), self.fc2.bias)


class _Block(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.norm1 = nn.LayerNorm(cfg.embed_dim, eps=cfg.ln_eps)
        self.attn = _Attn(cfg)
        self.norm2 = nn.LayerNorm(cfg.embed_dim, eps=cfg.ln_eps)
        self.mlp = _SwiGLU(cfg)

    def forward(self, x, sin, cos, r):
        x = r(x + self.attn(r(self.norm1(x)), sin, cos, r))
        return r(x + self.mlp(r(self.norm2(x)), r))


class _PatchEmbed(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.proj = nn.Conv2d(3, cfg.embed_dim, cfg.patch_size, cfg.patch_size)


class Eva02(nn.Module):
    """forward(x [N,3,S,S] float32, normalised as SpeciesClassifier.preprocess leaves it) -> logits [N,num_classes]."""

    def __init__(self, cfg: Eva02Config, lowp=False):
        super().__init__()
        self.cfg, self.lowp = cfg, lowp
        D, P = cfg.embed_dim, cfg.grid * cfg.grid
        self.patch_embed = _PatchEmbed(cfg)
        self.cls_token = nn.Parameter(torch.zeros(1, 1, D))
        self.pos_embed = nn.Parameter(torch.zeros(1, P + 1, D))
        self.blocks = nn.ModuleList([_Block(cfg) for _ in range(cfg.depth)])
        self.norm = nn.LayerNorm(D, eps=cfg.ln_eps)
        self.head = nn.Linear(D, cfg.num_classes)
        sin, cos = rope_table(cfg.grid, cfg.ref_grid, cfg.head_dim)
        self.register_buffer("rope_sin", sin, persistent=False)
        self.register_buffer("rope_cos", cos, persistent=False)

    def forward(self, x, taps=None):
        r = _bf if self.lowp else (lambda t: t)
        w = self.patch_embed.proj
        t = F.conv2d(r(x), r(w.weight), w.bias, stride=self.cfg.patch_size).flatten(2).transpose(1, 2)
        t = r(t)
        t = r(torch.cat([self.cls_token.expand(t.shape[0], -1, -1) + self.pos_embed[:, :1], t + self.pos_embed[:, 1:]], 1))
        if taps is not None:
            taps["tokens"] = t
        for i, b in enumerate(self.blocks):
            t = b(t, self.rope_sin, self.rope_cos, r)
            if taps is not None:
                taps["block%d" % i] = t
        c = r(self.norm(t[:, 0]))
        return F.linear(c, r(self.head.weight), self.head.bias)


def make_model(cfg: Eva02Config, seed=0, lowp=False, logit_std=2.5, branch_gain=0.5):
    """Seeded, well-conditioned random weights: unit-variance linears scaled so that every residual branch adds
    ~branch_gain / sqrt(depth) of the stream's scale (a trained ViT's regime rather than a random walk), LayerNorm
    gains around 1, and a head scaled to logits of std ~logit_std so that the softmax has a clear top-k."""
    g = torch.Generator().manual_seed(seed)
    m = Eva02(cfg, lowp=lowp)
    bs = branch_gain / math.sqrt(cfg.depth)

    def rn(*shape, std=1.0):
        return torch.randn(*shape, generator=g) * std

    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("norm.weight") or name.endswith("norm1.weight") or name.endswith("norm2.weight"):
                p.copy_(1.0 + 0.2 * (torch.rand(p.shape, generator=g) - 0.5))
            elif ".norm" in name and name.endswith(".bias") or name == "norm.bias":
                p.copy_(rn(*p.shape, std=0.05))
            elif name == "cls_token":
                p.copy_(rn(*p.shape, std=0.5))
            elif name == "pos_embed":
                p.copy_(rn(*p.shape, std=0.3))
            elif name == "patch_embed.proj.weight":
                p.copy_(rn(*p.shape, std=1.0 / math.sqrt(p[0].numel())))
            elif name == "head.weight":
                p.copy_(rn(*p.shape, std=logit_std / math.sqrt(p.shape[1])))
            elif name.endswith("proj.weight") and ".attn.proj" in name or name.endswith("fc2.weight"):
                p.copy_(rn(*p.shape, std=bs / math.sqrt(p.shape[1])))
            elif name.endswith(".weight"):
                p.copy_(rn(*p.shape, std=1.0 / math.sqrt(p.shape[1])))
            else:  # biases
                p.copy_(rn(*p.shape, std=0.02))
    return m.eval()


def state_dict_numpy(model):
    return {k: v.detach().cpu().numpy().astype(np.float32) for k, v in model.state_dict().items()}


def synthetic_input(cfg, n, seed=1):
    """[n,3,S,S] float32 in the value range SpeciesClassifier.preprocess produces (ImageNet-normalised pixels)."""
    g = torch.Generator().manual_seed(seed)
    u8 = torch.randint(0, 256, (n, 3, cfg.img_size // 7, cfg.img_size // 7), generator=g).float()
    img = F.interpolate(u8, size=(cfg.img_size, cfg.img_size), mode="bilinear", align_corners=False) / 255.0
    mean = torch.tensor([0.485, 0.456, 0.406]).view(1, 3, 1, 1)
    std = torch.tensor([0.229, 0.224, 0.225]).view(1, 3, 1, 1)
    return ((img - mean) / std).contiguous()
