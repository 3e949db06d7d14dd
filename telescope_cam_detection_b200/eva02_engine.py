"""The Stage-2 classifier body laid out as a libtcd layer program.

Replaces what `timm.create_model(model_name, num_classes=...)` builds and `self.model(input_tensor)` /
`torch.softmax` / `torch.topk` run in the reference (src/species_classifier.py:258-262, :386-391, :454-461) for the
model family the shipped configuration selects -- `timm/eva02_large_patch14_clip_336.merged2b_ft_inat21`
(config/config.yaml:279-281): the EVA02 vision transformer (patch 14, rotary position embedding on q / k, sub-LayerNorm
inside attention and inside the SwiGLU MLP, class-token read-out; architecture as published in timm/models/eva.py and
restated in oracle/eva02.py, which is what the GPU tests hold this engine to).

This module only *describes* the network from a state dict with timm's key names.  Every op becomes a hand-written
CUDA kernel launch inside libtcd, captured once as a CUDA graph for a fixed maximum batch of crops:
  * every Linear (patch embedding, qkv, proj, fc1_g || fc1_x fused, fc2, head) is a 1x1 conv of the tcgen05 / TMEM /
    TMA implicit-GEMM engine (csrc/conv_umma.cuh) over ALL token rows of the batch at once, bias / residual fused;
  * csrc/vit_kernels.cu: crop -> patch-row gather (the Stage-2 crop-and-resize kernel writing the patch-embedding
    GEMM's operand directly: no [N,3,S,S] tensor between Stage 1 and Stage 2), token assembly, rotary attention
    (head_dim 64), SwiGLU + LayerNorm, softmax, top-k read-out; LayerNorm / top-k / row gather are the RT-DETR kernels.
There is no torch / CPU fallback: without libtcd and an sm_100 device construction fails.
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import check
from .rtdetr_engine import ACT_NONE, Act, Program, to_bf16_bits

ACT_GLU = 4             # gated epilogue of the CTA-pair GEMM (include/tcd.h tcd_conv_desc.act)

# registered EVA02 sizes (timm/models/eva.py): embed_dim -> heads (head_dim is 64 in all of them)
_HEAD_DIM = 64
_MAX_COUT = 2048        # conv engine: a layer's bias is staged in shared memory (kConvMaxBias)
_HEAD_CHUNK = 256       # f32-row output mode of the conv engine: one N tile of <= 256 columns per launch


def is_eva02(model_name):
    return "eva02" in str(model_name).lower()


def rope_cos_sin(grid, ref_grid, head_dim=_HEAD_DIM, temperature=10000.0):
    """[grid*grid, head_dim/2, 2] float32 (cos, sin) per channel PAIR: timm's RotaryEmbeddingCat with in_pixels=False
    (head_dim/4 bands 1/T^(i/(head_dim/4)); positions arange(grid)/grid*ref_grid; pairs [0, hd/4) turn with the row
    position, pairs [hd/4, hd/2) with the column position)."""
    nb = head_dim // 4
    bands = (1.0 / (np.float32(temperature) ** (np.arange(nb, dtype=np.float32) / np.float32(nb)))).astype(np.float32)
    t = (np.arange(grid, dtype=np.float32) / np.float32(grid) * np.float32(ref_grid)).astype(np.float32)
    gy, gx = np.meshgrid(t, t, indexing="ij")
    ang = np.concatenate([gy.reshape(-1, 1) * bands[None], gx.reshape(-1, 1) * bands[None]], 1).astype(np.float32)
    return np.stack([np.cos(ang), np.sin(ang)], -1).astype(np.float32)


def _colsum_bf16(w):
    """Sum over the input channels of every weight row AS THE TENSOR CORE SEES IT (rounded to bf16), float64 accumulate."""
    bits = to_bf16_bits(np.ascontiguousarray(w, np.float32).reshape(w.shape[0], -1))
    return (bits.astype(np.uint32) << 16).view(np.float32).astype(np.float64).sum(1).astype(np.float32)


def _chunks(cout, step, cap=_MAX_COUT):
    """Split cout columns into pieces of <= cap that are multiples of `step` (the last one takes the remainder)."""
    n = max(1, math.ceil(cout / cap))
    per = math.ceil(cout / n / step) * step
    out, c = [], 0
    while c < cout:
        w = min(per, cout - c)
        out.append((c, w))
        c += w
    return out


class Eva02Engine:
    """EVA02 classifier on one GPU for up to `batch` crops per replay.

    sd: dict name -> float32 numpy array with timm's `Eva` key names (fused `attn.qkv.weight` + `q_bias` / `v_bias`,
    or separate q_proj / k_proj / v_proj).  `forward(x)` takes the preprocessed [n,3,S,S] batch; `crop_patches` +
    `run_topk` is the product path from a device frame and crop boxes.
    """

    def __init__(self, sd, batch, img_size=336, ref_grid=16, ln_eps=1e-6, top_k=5, device=0, keep_taps=False, fold_ln=True,
                 ln_in_gemm=True, ln_from_producers=True):
        self.sd = sd
        self.B, self.S = int(batch), int(img_size)
        pw = self._g("patch_embed.proj.weight")
        self.D, self.ps = int(pw.shape[0]), int(pw.shape[2])
        if self.S % self.ps:
            raise ValueError("input size %d is not a multiple of the patch size %d" % (self.S, self.ps))
        if self.D % _HEAD_DIM:
            raise ValueError("embed_dim %d is not a multiple of head_dim 64" % self.D)
        self.G = self.S // self.ps
        self.P, self.T = self.G * self.G, self.G * self.G + 1
        self.heads = self.D // _HEAD_DIM
        self.depth = 1 + max(int(k.split(".")[1]) for k in sd if k.startswith("blocks."))
        self.NC = int(self._g("head.weight").shape[0])
        pos = self._g("pos_embed")
        if pos.shape[1] != self.T:
            raise ValueError("pos_embed holds %d tokens, input size %d needs %d (resampling the table is what timm does "
                             "for other sizes; not supported)" % (pos.shape[1], self.S, self.T))
        self.ln_eps, self.ref_grid = float(ln_eps), int(ref_grid)
        self.K = 3 * self.ps * self.ps
        self.Kp = (self.K + 63) // 64 * 64
        self.top_k = int(min(top_k, self.NC))
        self.lib = _lib.load()
        self.prog = Program(device)
        self.taps = {} if keep_taps else None
        # Every LayerNorm of the network feeds a Linear, so its affine part folds into that Linear at pack time
        # (W' = W diag(gamma), b' = b + W beta, in float64): the LayerNorm kernels then load no parameters -- with them
        # the parameter loads were 4x the data loads per chunk and the kernels L1-bound (ncu: memory pipes 66 % busy at
        # 35 % of DRAM throughput).  fold_ln=False keeps the affine kernels (A/B, parity).
        self.fold_ln = bool(fold_ln)
        # Where every Linear behind a LayerNorm is a CTA-pair GEMM (many crops), the LayerNorm is applied in that GEMM's
        # epilogue from per-row (mean, 1/std): the normalised tensors are never written.  ln_in_gemm=False: A/B, parity.
        self.ln_in_gemm_wanted = bool(ln_in_gemm)
        self.ln_from_producers_wanted = bool(ln_from_producers)
        self._build()
        self.sd = None

    def close(self):
        self.prog.close()

    # ---- weights
    def _g(self, name):
        if name not in self.sd:
            raise KeyError("EVA02 weights: missing key %s" % name)
        return np.asarray(self.sd[name], np.float32)

    def _has(self, name):
        return name in self.sd

    def _tap(self, name, act):
        """Parity aid (keep_taps=True only): the residual stream is updated in place, so a tap is a copy."""
        if self.taps is None:
            return
        P, rows = self.prog, act.W
        if not hasattr(self, "_tap_idx"):
            self._tap_idx = P.upload(np.arange(rows, dtype=np.int32))
        keep = Act(P.alloc(rows * act.C * 2), 1, 1, rows, act.C)
        P.op("tap." + name, "gather", [act.ptr, self._tap_idx, keep.ptr], [1, rows, rows, act.C])
        self.taps[name] = keep

    def _linear(self, name, x, w, b, y, resid=None, step=256, ln_stats=None, ln_parts=0, ln_count=0, stats_out=None, stats_parts=0):
        """y[:, :cout] = x @ w.T + b (+ resid), split along cout so that each launch stays within the engine's bias
        staging limit.  x, y, resid: flat token views [1,1,rows,C].  ln_stats: x is un-normalised and the launch applies
        the LayerNorm whose (mean, 1/std) per row sit at that device address (w, b already hold its affine part)."""
        cout = w.shape[0]
        if cout % step:
            step = 128 if cout % 128 == 0 else 32
        cs = _colsum_bf16(w) if ln_stats is not None else None
        for c0, cw in _chunks(cout, step):
            self.prog.conv(name if cout == cw else "%s[%d:%d]" % (name, c0, c0 + cw), x, w[c0:c0 + cw], b[c0:c0 + cw],
                           act=ACT_NONE, y=y, y_coff=c0, resid=resid, r_coff=c0 if resid is not None else 0,
                           ln_stats=ln_stats, ln_colsum=None if cs is None else cs[c0:c0 + cw], ln_parts=ln_parts, ln_count=ln_count,
                           ln_eps=self.ln_eps, stats_out=stats_out, stats_parts=stats_parts, stats_slot0=c0 // 256)

    def _qkv_weights(self, p, zeros):
        """Fused [q; k; v] weight and bias of block prefix p (fused qkv + q_bias / v_bias, or separate projections)."""
        if self._has(p + "attn.qkv.weight"):
            wqkv = self._g(p + "attn.qkv.weight")
            bq = self._g(p + "attn.q_bias") if self._has(p + "attn.q_bias") else zeros
            bv = self._g(p + "attn.v_bias") if self._has(p + "attn.v_bias") else zeros
            return wqkv, np.concatenate([bq, zeros, bv])
        wqkv = np.concatenate([self._g(p + "attn.q_proj.weight"), self._g(p + "attn.k_proj.weight"),
                               self._g(p + "attn.v_proj.weight")], 0)
        bk = self._g(p + "attn.k_proj.bias") if self._has(p + "attn.k_proj.bias") else zeros
        return wqkv, np.concatenate([self._g(p + "attn.q_proj.bias"), bk, self._g(p + "attn.v_proj.bias")])

    @staticmethod
    def _pad_rows(a, n):
        out = np.zeros((n,) + a.shape[1:], np.float32)
        out[:a.shape[0]] = a
        return out

    def _norm(self, name, a, prefix, y, rows, width=None):
        """LayerNorm `prefix` of `a` for a following Linear -> (tensor the Linear reads, ln_stats for it).  With
        consumer-side normalisation only the row statistics are computed here (8 bytes per row) and the Linear reads `a`
        itself; otherwise the normalised rows are written to `y`."""
        if self.ln_in_gemm:
            stats = self._stats[self._stats_i % len(self._stats)]
            self._stats_i += 1
            self.prog.op(name + ".stats", "row_stats", [a.ptr, stats], [rows, width or a.C, a.C], [self.ln_eps])
            return a, stats
        self._ln(name, a, prefix, y, rows)
        return y, None

    def _ln(self, name, a, prefix, y, rows):
        P = self.prog
        if self.fold_ln:
            P.op(name, "layernorm", [a.ptr, 0, 0, 0, y.ptr], [rows, a.C], [self.ln_eps])
            return
        g = P.upload(self._g(prefix + ".weight"))
        b = P.upload(self._g(prefix + ".bias"))
        P.op(name, "layernorm", [a.ptr, 0, g, b, y.ptr], [rows, a.C], [self.ln_eps])

    def _fold(self, w, b, prefix):
        """(w, b) of a Linear that reads LayerNorm `prefix`'s output -> the pair that reads the un-scaled normalisation."""
        if not self.fold_ln:
            return w, b
        g, be = self._g(prefix + ".weight").astype(np.float64), self._g(prefix + ".bias").astype(np.float64)
        w64 = w.astype(np.float64)
        return (w64 * g[None, :]).astype(np.float32), (b.astype(np.float64) + w64 @ be).astype(np.float32)

    # ---- the program
    def _build(self):
        P, B, T, D, NP = self.prog, self.B, self.T, self.D, self.P
        rows = B * T
        H = int(self._g("blocks.0.mlp.fc1_g.weight").shape[0])
        Hp = (H + 127) // 128 * 128        # 2 * Hp is a multiple of 256: the fused fc1 splits into 128 x 256 GEMM tiles
        self.hidden, self.hidden_pad = H, Hp
        # inputs: a preprocessed f32 batch (forward) or patch rows written by tcd_crop_patches (product path)
        self.in_f32 = P.f32(B * 3 * self.S * self.S)
        self.rows0 = Act(P.alloc(B * NP * self.Kp * 2), 1, 1, B * NP, self.Kp)
        self.boxes_dev = P.alloc(B * 16)
        # patch embedding: Conv2d(3, D, ps, ps) == Linear over the flattened patch (zero-padded to Kp columns)
        pw = np.zeros((D, self.Kp), np.float32)
        pw[:, :self.K] = self._g("patch_embed.proj.weight").reshape(D, self.K)
        emb = Act(P.alloc(B * NP * D * 2), 1, 1, B * NP, D)
        self._linear("patch_embed", self.rows0, pw, self._g("patch_embed.proj.bias"), emb)
        pos = self._g("pos_embed").reshape(T, D).copy()
        pos[0] += self._g("cls_token").reshape(D)
        x = Act(P.alloc(rows * D * 2), 1, 1, rows, D)
        x2 = Act(P.alloc(rows * D * 2), 1, 1, rows, D)
        P.op("tokens", "vit_assemble", [emb.ptr, P.upload(pos), x.ptr], [B, T, D])
        self._tap("tokens", x)
        yb = Act(P.alloc(rows * D * 2), 1, 1, rows, D)
        qkv = Act(P.alloc(rows * 3 * D * 2), 1, 1, rows, 3 * D)
        att = Act(P.alloc(rows * D * 2), 1, 1, rows, D)
        # The gated epilogue exists in the CTA-pair GEMM only: same eligibility rule as csrc/conv_umma.cu
        # (gemm_pair_eligible): >= 74 work items of 256 rows x 256 columns, K >= 512 in 64s.
        self.fuse_glu = D >= 512 and D % 64 == 0 and all(((rows + 255) // 256) * (cw // 256) >= 74 and cw % 256 == 0
                                                         for _, cw in _chunks(2 * Hp, 256))
        # Consumer-side LayerNorm (csrc/gemm_pair.cuh): every Linear that follows a LayerNorm must be a CTA-pair GEMM
        def pair_ok(cin, cout):
            return cin >= 512 and cin % 64 == 0 and all(cw % 256 == 0 and ((rows + 255) // 256) * (cw // 256) >= 74
                                                        for _, cw in _chunks(cout, 256))
        self.ln_in_gemm = (self.fold_ln and self.fuse_glu and self.ln_in_gemm_wanted and pair_ok(D, 3 * D) and pair_ok(D, D)
                           and pair_ok(Hp, D))
        # ... and where, additionally, every LayerNorm INPUT is written by a CTA-pair GEMM or the tcgen05 attention kernel,
        # those producers write per-tile partial (sum, sum of squares) of their rows and no statistics kernel runs at all
        self.ln_from_producers = self.ln_in_gemm and self.ln_from_producers_wanted and D % 256 == 0 and self.T <= 640
        self._stats = [P.f32(rows * 2), P.f32(rows * 2)] if self.ln_in_gemm and not self.ln_from_producers else []
        self._stats_i = 0
        if self.ln_from_producers:
            # partial slots per row, rounded up to even (the consumers read them two per 16-byte load; a slot no producer
            # writes stays at the zero the buffer was allocated with)
            px, pa, ph = (D // 256 + 1) & ~1, self.heads * 2, (Hp // 128 + 1) & ~1
            sx, sx2, satt, sh = P.f32(rows * px * 2), P.f32(rows * px * 2), P.f32(rows * pa * 2), P.f32(rows * ph * 2)
        j = np.arange(Hp)
        self._glu_g = (j // 128) * 256 + ((j % 128) // 32) * 64 + (j % 32)    # GEMM column of the gate of hidden j
        gu = Act(P.alloc(rows * Hp * 2), 1, 1, rows, Hp) if self.fuse_glu else Act(P.alloc(rows * 2 * Hp * 2), 1, 1, rows, 2 * Hp)
        hb = Act(P.alloc(rows * Hp * 2), 1, 1, rows, Hp)
        rope = P.upload(rope_cos_sin(self.G, self.ref_grid))
        zeros = np.zeros(D, np.float32)
        for i in range(self.depth):
            p = "blocks.%d." % i
            n = "b%d." % i
            if self.ln_from_producers:
                # no normalisation or statistics kernel in the block: producers write partial row statistics, consumers
                # normalise in their epilogues
                if i == 0:
                    P.op("tokens.stats", "row_stats", [x.ptr, sx], [rows, D, D, px], [self.ln_eps])
                wqkv, bqkv = self._qkv_weights(p, zeros)
                wqkv, bqkv = self._fold(wqkv, bqkv, p + "norm1")
                self._linear(n + "qkv", x, wqkv, bqkv, qkv, ln_stats=sx, ln_parts=px, ln_count=D)
                has_an = self._has(p + "attn.norm.weight")
                P.op(n + "mha", "mha_rope", [qkv.ptr, qkv.ptr, qkv.ptr, att.ptr, rope, satt if has_an else 0],
                     [3 * D, 0, 3 * D, D, 3 * D, 2 * D, D, B, T, self.heads, 1], [_HEAD_DIM ** -0.5])
                wp, bp = self._g(p + "attn.proj.weight"), self._g(p + "attn.proj.bias")
                if has_an:
                    wp, bp = self._fold(wp, bp, p + "attn.norm")
                    self._linear(n + "proj", att, wp, bp, x2, resid=x, ln_stats=satt, ln_parts=pa, ln_count=D, stats_out=sx2, stats_parts=px)
                else:
                    self._linear(n + "proj", att, wp, bp, x2, resid=x, stats_out=sx2, stats_parts=px)
                wg, bg = self._fold(self._g(p + "mlp.fc1_g.weight"), self._g(p + "mlp.fc1_g.bias"), p + "norm2")
                wx, bx = self._fold(self._g(p + "mlp.fc1_x.weight"), self._g(p + "mlp.fc1_x.bias"), p + "norm2")
                w1 = np.zeros((2 * Hp, D), np.float32)
                b1 = np.zeros(2 * Hp, np.float32)
                w1[self._glu_g], w1[self._glu_g + 32] = self._pad_rows(wg, Hp), self._pad_rows(wx, Hp)
                b1[self._glu_g], b1[self._glu_g + 32] = self._pad_rows(bg, Hp), self._pad_rows(bx, Hp)
                cs1 = _colsum_bf16(w1)
                for c0, cw in _chunks(2 * Hp, 256):
                    P.conv("%sfc1[%d:%d]" % (n, c0, c0 + cw), x2, w1[c0:c0 + cw], b1[c0:c0 + cw], act=ACT_GLU, y=gu, y_coff=c0 // 2,
                           ln_stats=sx2, ln_colsum=cs1[c0:c0 + cw], ln_parts=px, ln_count=D, ln_eps=self.ln_eps,
                           stats_out=sh, stats_parts=ph, stats_slot0=c0 // 256)
                w2 = np.zeros((D, Hp), np.float32)
                wf2, bf2 = self._fold(self._g(p + "mlp.fc2.weight"), self._g(p + "mlp.fc2.bias"), p + "mlp.norm")
                w2[:, :H] = wf2
                self._linear(n + "fc2", gu, w2, bf2, x, resid=x2, ln_stats=sh, ln_parts=ph, ln_count=H, stats_out=sx, stats_parts=px)
                self._tap("block%d" % i, x)
                continue
            a_in, st = self._norm(n + "norm1", x, p + "norm1", yb, rows)
            wqkv, bqkv = self._qkv_weights(p, zeros)
            wqkv, bqkv = self._fold(wqkv, bqkv, p + "norm1")
            self._linear(n + "qkv", a_in, wqkv, bqkv, qkv, ln_stats=st)
            P.op(n + "mha", "mha_rope", [qkv.ptr, qkv.ptr, qkv.ptr, att.ptr, rope],
                 [3 * D, 0, 3 * D, D, 3 * D, 2 * D, D, B, T, self.heads, 1], [_HEAD_DIM ** -0.5])
            a_in, st = att, None
            wp, bp = self._g(p + "attn.proj.weight"), self._g(p + "attn.proj.bias")
            if self._has(p + "attn.norm.weight"):       # scale_attn_inner (the *_clip_* registrations)
                a_in, st = self._norm(n + "attn_norm", att, p + "attn.norm", yb, rows)
                wp, bp = self._fold(wp, bp, p + "attn.norm")
            self._linear(n + "proj", a_in, wp, bp, x2, resid=x, ln_stats=st)
            a_in, st = self._norm(n + "norm2", x2, p + "norm2", yb, rows)
            w1 = np.zeros((2 * Hp, D), np.float32)
            b1 = np.zeros(2 * Hp, np.float32)
            if not self._has(p + "mlp.norm.weight"):
                raise KeyError("EVA02 weights: %smlp.norm.* missing (scale_mlp = False variants are not supported)" % p)
            wg, bg = self._fold(self._g(p + "mlp.fc1_g.weight"), self._g(p + "mlp.fc1_g.bias"), p + "norm2")
            wx, bx = self._fold(self._g(p + "mlp.fc1_x.weight"), self._g(p + "mlp.fc1_x.bias"), p + "norm2")
            if self.fold_ln:
                ng_ptr = nb_ptr = 0
            else:
                ng, nb = np.zeros(Hp, np.float32), np.zeros(Hp, np.float32)
                ng[:H], nb[:H] = self._g(p + "mlp.norm.weight"), self._g(p + "mlp.norm.bias")
                ng_ptr, nb_ptr = P.upload(ng), P.upload(nb)
            if self.fuse_glu:
                # gated epilogue of the CTA-pair GEMM: within every 256-column tile the 32-column slabs alternate
                # gate / value, and the GEMM writes silu(g) * u straight into hb (hidden order, Hp columns)
                w1[self._glu_g], w1[self._glu_g + 32] = self._pad_rows(wg, Hp), self._pad_rows(wx, Hp)
                b1[self._glu_g], b1[self._glu_g + 32] = self._pad_rows(bg, Hp), self._pad_rows(bx, Hp)
                cs1 = _colsum_bf16(w1) if st is not None else None
                for c0, cw in _chunks(2 * Hp, 256):
                    P.conv("%sfc1[%d:%d]" % (n, c0, c0 + cw), a_in, w1[c0:c0 + cw], b1[c0:c0 + cw], act=ACT_GLU, y=gu, y_coff=c0 // 2,
                           ln_stats=st, ln_colsum=None if cs1 is None else cs1[c0:c0 + cw])
                if self.ln_in_gemm:
                    h_in, st = self._norm(n + "mlp_norm", gu, p + "mlp.norm", hb, rows, width=H)
                else:
                    P.op(n + "swiglu_ln", "swiglu_ln", [gu.ptr, ng_ptr, nb_ptr, hb.ptr], [rows, H, Hp, 1], [self.ln_eps])
                    h_in, st = hb, None
            else:
                w1[:H], w1[Hp:Hp + H] = wg, wx
                b1[:H], b1[Hp:Hp + H] = bg, bx
                self._linear(n + "fc1", a_in, w1, b1, gu)
                P.op(n + "swiglu_ln", "swiglu_ln", [gu.ptr, ng_ptr, nb_ptr, hb.ptr], [rows, H, Hp, 0], [self.ln_eps])
                h_in, st = hb, None
            w2 = np.zeros((D, Hp), np.float32)
            wf2, bf2 = self._fold(self._g(p + "mlp.fc2.weight"), self._g(p + "mlp.fc2.bias"), p + "mlp.norm")
            w2[:, :H] = wf2
            self._linear(n + "fc2", h_in, w2, bf2, x, resid=x2, ln_stats=st)
            self._tap("block%d" % i, x)
        # read-out: final LayerNorm on the class tokens only, head, softmax, top-k
        cls = Act(P.alloc(B * D * 2), 1, 1, B, D)
        clsn = Act(P.alloc(B * D * 2), 1, 1, B, D)
        zidx = P.upload(np.zeros(B, np.int32))
        P.op("cls_gather", "gather", [x.ptr, zidx, cls.ptr], [B, T, 1, D])
        self._ln("norm", cls, "norm", clsn, B)
        self.ld_logits = (self.NC + _HEAD_CHUNK - 1) // _HEAD_CHUNK * _HEAD_CHUNK
        self.logits = P.f32(B * self.ld_logits)
        hw, hbias = self._fold(self._g("head.weight"), self._g("head.bias"), "norm")
        if D <= 1024:
            # few rows x many classes: a weight-streaming kernel (csrc/vit_kernels.cu head_gemv), one launch
            P.op("head", "head_gemv", [clsn.ptr, P.upload(to_bf16_bits(hw)), P.upload(hbias), self.logits],
                 [B, D, self.NC, self.ld_logits])
        else:
            for c0 in range(0, self.NC, _HEAD_CHUNK):     # f32-row mode of the conv engine, 256 columns per launch
                cw = min(_HEAD_CHUNK, self.NC - c0)
                d = _lib.ConvDesc()
                d.name = ("head[%d:%d]" % (c0, c0 + cw)).encode()
                d.x, d.N, d.H, d.W, d.x_ctot, d.x_coff, d.cin, d.x_nstride = clsn.ptr, 1, 1, B, D, 0, D, 0
                w = np.ascontiguousarray(hw[c0:c0 + cw]).reshape(cw, D, 1, 1)
                b = np.ascontiguousarray(hbias[c0:c0 + cw])
                d.w_host, d.bias_host = w.ctypes.data, b.ctypes.data
                d.cout, d.k, d.stride, d.act = cw, 1, 1, ACT_NONE
                d.out_f32, d.out_ld, d.out_col, d.rows_per_img, d.row_off = self.logits, self.ld_logits, c0, B, 0
                check(self.lib.tcd_program_add_conv(P.handle, C.byref(d)), "tcd_program_add_conv(head)")
        self.probs = P.f32(B * self.NC)
        P.op("softmax", "softmax", [self.logits, self.probs], [B, self.NC, self.ld_logits, self.NC])
        self.top_idx = P.alloc(B * self.top_k * 4)
        self.top_p = P.f32(B * self.top_k)
        P.op("topk", "topk", [self.probs, self.top_idx], [B, self.NC, self.top_k])
        P.op("take", "take", [self.probs, self.top_idx, self.top_p], [B, self.top_k, self.NC])

    # ---- inputs
    def set_input(self, x_nchw):
        """x: float32 [n,3,S,S] (numpy), n <= batch, normalised as SpeciesClassifier.preprocess leaves it."""
        x = np.ascontiguousarray(x_nchw, np.float32)
        if x.ndim != 4 or x.shape[1:] != (3, self.S, self.S) or x.shape[0] > self.B:
            raise ValueError("expected [n<=%d,3,%d,%d], got %s" % (self.B, self.S, self.S, x.shape))
        self.prog.h2d(self.in_f32, x)
        self.patchify_device(self.in_f32, x.shape[0])
        return x.shape[0]

    def patchify_device(self, x_ptr, n):
        """Patch rows from a DEVICE float32 [n,3,S,S] batch (e.g. a torch tensor's data_ptr)."""
        check(self.lib.tcd_patchify_f32(C.c_void_p(int(x_ptr)), int(n), self.S, self.ps, self.Kp,
                                        C.c_void_p(self.rows0.ptr), C.c_void_p(self.prog.stream())), "tcd_patchify_f32")

    def crop_patches(self, frame_ptr, h, w, stride, boxes, swap_rb=1, scale=1.0 / 255.0, mean=(0.485, 0.456, 0.406),
                     std=(0.229, 0.224, 0.225)):
        """All crops of one device frame (u8 HWC BGR) -> patch rows, one launch (tcd_crop_patches)."""
        boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
        if boxes.shape[0] > self.B:
            raise ValueError("%d crops exceed the engine's batch %d" % (boxes.shape[0], self.B))
        cfg = _lib.make_preproc_cfg(self.S, self.S, 0, swap_rb, scale, mean, std, 0.0)
        check(self.lib.tcd_crop_patches(C.c_void_p(int(frame_ptr)), int(h), int(w), int(stride),
                                        boxes.ctypes.data_as(_lib._I32P), C.c_void_p(self.boxes_dev), boxes.shape[0],
                                        self.S, self.ps, self.Kp, C.byref(cfg), C.c_void_p(self.rows0.ptr),
                                        C.c_void_p(self.prog.stream())), "tcd_crop_patches")
        return boxes.shape[0]

    # ---- replay
    def run(self, graph=True):
        self.prog.run(graph)

    def sync(self):
        self.prog.sync()

    def read_logits(self, n):
        return self.prog.d2h(self.logits, (self.B, self.ld_logits), np.float32)[:n, :self.NC].copy()

    def read_probs(self, n):
        return self.prog.d2h(self.probs, (self.B, self.NC), np.float32)[:n].copy()

    def read_topk(self, n):
        """(probs [n,k] float32 descending, class ids [n,k] int32) -- torch.softmax + torch.topk of the reference."""
        p = self.prog.d2h(self.top_p, (self.B, self.top_k), np.float32)[:n].copy()
        i = self.prog.d2h(self.top_idx, (self.B, self.top_k), np.int32)[:n].copy()
        return p, i

    def forward(self, x_nchw, graph=True):
        """-> logits [n, num_classes] float32."""
        n = self.set_input(x_nchw)
        self.run(graph)
        self.sync()
        return self.read_logits(n)

    def read_tap(self, name):
        a = self.taps[name]
        bits = self.prog.d2h(a.ptr, (self.B, self.T, self.D), np.uint16)
        return (bits.astype(np.uint32) << 16).view(np.float32)

    def flops_per_crop(self):
        """2*MACs of the Linears + attention of one crop (what the program executes, padded columns excluded)."""
        D, T, H = self.D, self.T, self.hidden
        per_tok = 2.0 * D * (3 * D + D + 3 * H)
        attn = 4.0 * T * T * D
        return self.depth * (per_tok * T + attn) + 2.0 * self.P * self.K * D + 2.0 * D * self.NC
