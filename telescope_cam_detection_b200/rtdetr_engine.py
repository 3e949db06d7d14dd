"""RT-DETRv2-R50 laid out as a libtcd layer program (include/tcd.h "layer programs").

Replaces what RTDETRDetector.load_model builds and self.model(img_tensor, orig_size) runs
(src/rtdetr_detector.py:60-204,257,340): upstream lyuwenyu/RT-DETR rtdetrv2_pytorch (un-vendored
submodule) in deploy mode -- PResNet-50-vd backbone, hybrid encoder (AIFI + CCFM with
RepVGG blocks re-parameterised at pack time, which is what model.deploy() does, :164), query
selection, six decoder layers with MSDeformAttn, and the focal-loss postprocessor (top-300 over
queries x classes, no NMS).

This module only *describes* the network: every op below becomes a hand-written CUDA kernel launch
inside libtcd (tcgen05 implicit-GEMM convs / linears, csrc/conv_umma.cuh; transformer kernels,
csrc/tfm_kernels.cuh), captured once as a CUDA graph.  Weights are taken from a state dict with the
key names of the `transformers` RT-DETRv2 implementation (the independent implementation SURVEY
section 8(c) names as the cross-check; a key map from upstream checkpoints is a TODO noted in
DESIGN.md).  There is no torch / CPU fallback: without libtcd and an sm_100 device construction fails.
"""
import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import check

ACT_NONE, ACT_SILU, ACT_RELU, ACT_GELU = 0, 1, 2, 3


class Act:
    """A bf16 NHWC (or token-major, H == 1) device tensor view."""

    def __init__(self, ptr, N, H, W, C, nstride=0):
        self.ptr, self.N, self.H, self.W, self.C, self.nstride = int(ptr), N, H, W, C, nstride

    def tokens(self):
        """[N, H*W, C] view of a dense NHWC map."""
        return Act(self.ptr, self.N, 1, self.H * self.W, self.C, self.nstride)


def fold_bn(w, bn_w, bn_b, mean, var, eps):
    """conv weight [co,ci,k,k] + eval BatchNorm -> (w', b') in float64, returned as float32."""
    s = bn_w.astype(np.float64) / np.sqrt(var.astype(np.float64) + eps)
    w2 = w.astype(np.float64) * s[:, None, None, None]
    b2 = bn_b.astype(np.float64) - mean.astype(np.float64) * s
    return w2.astype(np.float32), b2.astype(np.float32)


def sincos_pos_embed(w, h, dim=256, temperature=10000.0):
    """2-D sin-cos position embedding of the AIFI layer, [h*w, dim], order [sin h, cos h, sin w, cos w]
    (SURVEY App. D step 3)."""
    gw, gh = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32), indexing="xy")
    pos_dim = dim // 4
    omega = (np.arange(pos_dim, dtype=np.float32) / pos_dim)
    omega = (1.0 / (np.float32(temperature) ** omega)).astype(np.float32)
    out_w = gw.reshape(-1, 1) @ omega[None]
    out_h = gh.reshape(-1, 1) @ omega[None]
    return np.concatenate([np.sin(out_h), np.cos(out_h), np.sin(out_w), np.cos(out_w)], axis=1).astype(np.float32)


def make_anchors(shapes, grid_size=0.05, eps=1e-2):
    """Logit-space anchors [sum(HW), 4] and the validity mask."""
    out = []
    for lvl, (h, w) in enumerate(shapes):
        gy, gx = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
        xy = np.stack([(gx + 0.5) / np.float32(w), (gy + 0.5) / np.float32(h)], -1)
        wh = np.ones_like(xy) * np.float32(grid_size) * np.float32(2.0 ** lvl)
        out.append(np.concatenate([xy, wh], -1).reshape(-1, 4))
    a = np.concatenate(out, 0).astype(np.float32)
    valid = ((a > eps) & (a < 1 - eps)).all(-1)
    with np.errstate(divide="ignore", invalid="ignore"):
        logit = np.log(a / (1 - a)).astype(np.float32)
    logit = np.where(valid[:, None], logit, np.finfo(np.float32).max).astype(np.float32)
    return logit, valid.astype(np.float32)


def to_bf16_bits(a):
    """float32 array -> uint16 bf16 bit patterns, round-to-nearest-even."""
    u = np.ascontiguousarray(a, np.float32).view(np.uint32).astype(np.uint64)
    u = u + 0x7FFF + ((u >> 16) & 1)
    return (u >> 16).astype(np.uint16)


class Program:
    """Thin owner of one tcd_program: buffers, op list, replay."""

    def __init__(self, device=0):
        self.lib = _lib.load()
        h = C.c_void_p()
        check(self.lib.tcd_program_create(int(device), C.byref(h)), "tcd_program_create")
        self.handle = h
        self._keep = []

    def close(self):
        if getattr(self, "handle", None):
            self.lib.tcd_program_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def alloc(self, nbytes):
        p = C.c_void_p()
        check(self.lib.tcd_program_alloc(self.handle, int(nbytes), C.byref(p)), "tcd_program_alloc")
        return int(p.value)

    def upload(self, arr):
        arr = np.ascontiguousarray(arr)
        p = C.c_void_p()
        check(self.lib.tcd_program_upload(self.handle, arr.ctypes.data_as(C.c_void_p), arr.nbytes, C.byref(p)),
              "tcd_program_upload")
        return int(p.value)

    def act(self, N, H, W, Cc):
        return Act(self.alloc(N * H * W * Cc * 2), N, H, W, Cc)

    def f32(self, n):
        return self.alloc(n * 4)

    def conv(self, name, x, w, b, k=1, stride=1, act=ACT_NONE, y=None, y_coff=0, x_coff=0, resid=None,
             r_coff=0, resid_pre=0, out_f32=None, out_ld=0, ln_stats=None, ln_colsum=None, ln_parts=0, ln_count=0,
             ln_eps=0.0, stats_out=None, stats_parts=0, stats_slot0=0):
        """w: [cout, cin, k, k] float32 (BN already folded), b: [cout]."""
        w = np.ascontiguousarray(w, np.float32).reshape(w.shape[0], -1, k, k)
        b = np.ascontiguousarray(b, np.float32)
        d = _lib.ConvDesc()
        d.name = name.encode()[:63]
        d.x, d.N, d.H, d.W, d.x_ctot, d.x_coff, d.cin, d.x_nstride = x.ptr, x.N, x.H, x.W, x.C, x_coff, w.shape[1], x.nstride
        d.w_host, d.bias_host = w.ctypes.data, b.ctypes.data
        d.cout, d.k, d.stride, d.act = w.shape[0], k, stride, act
        if out_f32 is not None:
            d.out_f32, d.out_ld, d.out_col = out_f32, out_ld or w.shape[0], 0
            d.rows_per_img, d.row_off = (x.H // stride) * (x.W // stride), 0
        else:
            d.y, d.y_ctot, d.y_coff, d.y_nstride = y.ptr, y.C, y_coff, y.nstride
        if resid is not None:
            d.resid, d.r_ctot, d.r_coff, d.resid_pre, d.r_nstride = resid.ptr, resid.C, r_coff, resid_pre, resid.nstride
        if ln_stats is not None:       # LayerNorm applied in the consumer's epilogue (include/tcd.h tcd_conv_desc.ln_stats)
            cs = np.ascontiguousarray(ln_colsum, np.float32)
            d.ln_stats, d.ln_colsum_host = int(ln_stats), cs.ctypes.data
            d.ln_parts, d.ln_count, d.ln_eps = int(ln_parts), int(ln_count), float(ln_eps)
        if stats_out is not None:      # this launch's tiles also write the LayerNorm statistics of their rows
            d.stats_out, d.stats_parts, d.stats_slot0 = int(stats_out), int(stats_parts), int(stats_slot0)
        check(self.lib.tcd_program_add_conv(self.handle, C.byref(d)), "tcd_program_add_conv(%s)" % name)

    def op(self, name, kind, ptrs, ints, floats=()):
        d = _lib.OpDesc()
        d.name = name.encode()[:63]
        d.kind = _lib.OP_KINDS[kind]
        for j, p in enumerate(ptrs):
            d.p[j] = int(p) if p else None
        for j, v in enumerate(ints):
            d.i[j] = int(v)
        for j, v in enumerate(floats):
            d.f[j] = float(v)
        check(self.lib.tcd_program_add_op(self.handle, C.byref(d)), "tcd_program_add_op(%s)" % name)

    def run(self, graph=True):
        check(self.lib.tcd_program_run(self.handle, 1 if graph else 0), "tcd_program_run")

    def sync(self):
        check(self.lib.tcd_program_sync(self.handle), "tcd_program_sync")

    def stream(self):
        return int(self.lib.tcd_program_stream(self.handle) or 0)

    def h2d(self, dst, arr):
        arr = np.ascontiguousarray(arr)
        check(self.lib.tcd_memcpy_h2d(C.c_void_p(dst), arr.ctypes.data_as(C.c_void_p), arr.nbytes), "tcd_memcpy_h2d")

    def d2h(self, src, shape, dtype):
        out = np.empty(shape, dtype)
        check(self.lib.tcd_memcpy_d2h(out.ctypes.data_as(C.c_void_p), C.c_void_p(src), out.nbytes), "tcd_memcpy_d2h")
        return out

    def read_bf16(self, a):
        """Act -> float32 numpy [N,H,W,C] (dense views only)."""
        bits = self.d2h(a.ptr, (a.N, a.H, a.W, a.C), np.uint16)
        return (bits.astype(np.uint32) << 16).view(np.float32)

    def profile(self, iters=3):
        n = self.lib.tcd_program_num_ops(self.handle)
        ms = (C.c_float * n)()
        check(self.lib.tcd_program_profile(self.handle, int(iters), ms, n), "tcd_program_profile")
        return [{"name": self.lib.tcd_program_op_name(self.handle, i).decode(), "ms": float(ms[i]),
                 "flops": float(self.lib.tcd_program_op_flops(self.handle, i))} for i in range(n)]


class RTDetrEngine:
    """RT-DETRv2-R50 on one GPU for a fixed batch size: `forward(x)` / `detect_tensors(...)`.

    sd: dict name -> float32 numpy array with the `transformers` RTDetrV2ForObjectDetection key names.
    """

    def __init__(self, sd, batch, input_size=(640, 640), num_queries=300, num_classes=80, heads=8,
                 n_points=4, offset_scale=0.5, bn_eps=1e-5, ln_eps=1e-5, device=0, keep_taps=False,
                 external_topk=False):
        self.B, self.IH, self.IW = int(batch), int(input_size[0]), int(input_size[1])
        if self.IH % 32 or self.IW % 32:
            raise ValueError("input size must be a multiple of 32, got %s" % (input_size,))
        self.Q, self.NC, self.heads, self.n_points = num_queries, num_classes, heads, n_points
        self.offset_scale, self.bn_eps, self.ln_eps = offset_scale, bn_eps, ln_eps
        self.sd = sd
        # parity aid: skip the in-graph query selection and use indices written by set_topk() (lets a test
        # compare the decoder on exactly the oracle's query set; the product path never sets it)
        self.external_topk = bool(external_topk)
        self.prog = Program(device)
        self.taps = {} if keep_taps else None
        self._build()
        self.sd = None

    def close(self):
        self.prog.close()

    # ---- weight helpers
    def _g(self, name):
        if name not in self.sd:
            raise KeyError("RT-DETR weights: missing key %s" % name)
        return np.asarray(self.sd[name], np.float32)

    def _convbn(self, prefix, conv="conv", norm="norm"):
        return fold_bn(self._g(prefix + "." + conv + ".weight"), self._g(prefix + "." + norm + ".weight"),
                       self._g(prefix + "." + norm + ".bias"), self._g(prefix + "." + norm + ".running_mean"),
                       self._g(prefix + "." + norm + ".running_var"), self.bn_eps)

    def _resnet_cb(self, prefix):
        return self._convbn(prefix, "convolution", "normalization")

    def _lin(self, prefix):
        w = self._g(prefix + ".weight")
        return w.reshape(w.shape[0], w.shape[1], 1, 1), self._g(prefix + ".bias")

    def _tap(self, name, act):
        if self.taps is not None:
            self.taps[name] = act

    # ---- sub-graphs
    def _linear(self, name, x, prefix, act=ACT_NONE, y=None, out_f32=None, out_ld=0):
        """x: token Act [N,1,T,Cin] -> y [N,1,T,Cout] (allocated when not given)."""
        w, b = self._lin(prefix)
        if out_f32 is not None:
            self.prog.conv(name, x, w, b, act=act, out_f32=out_f32, out_ld=out_ld)
            return None
        if y is None:
            y = self.prog.act(x.N, 1, x.W, w.shape[0])
        self.prog.conv(name, x, w, b, act=act, y=y)
        return y

    def _layernorm(self, name, a, b, prefix, y=None):
        P = self.prog
        if y is None:
            y = P.act(a.N, a.H, a.W, a.C)
        g = P.upload(self._g(prefix + ".weight"))
        be = P.upload(self._g(prefix + ".bias"))
        P.op(name, "layernorm", [a.ptr, b.ptr if b is not None else 0, g, be, y.ptr], [a.N * a.H * a.W, a.C],
             [self.ln_eps])
        return y

    def _self_attention(self, name, x, pos, prefix, pos_bcast):
        """RTDetrV2SelfAttention: q = k = x + pos, v = x; returns o_proj(attn) (no residual)."""
        P = self.prog
        T = x.W
        qk_in = P.act(x.N, 1, T, x.C)
        P.op(name + ".qk_in", "add_rows", [x.ptr, pos.ptr, qk_in.ptr], [x.N, T, x.C, 1 if pos_bcast else 0])
        wq, bq = self._lin(prefix + ".q_proj")
        wk, bk = self._lin(prefix + ".k_proj")
        qk = P.act(x.N, 1, T, 2 * x.C)
        P.conv(name + ".qk_proj", qk_in, np.concatenate([wq, wk], 0), np.concatenate([bq, bk], 0), y=qk)
        v = self._linear(name + ".v_proj", x, prefix + ".v_proj")
        attn = P.act(x.N, 1, T, x.C)
        hd = x.C // self.heads
        assert hd == 32, "head_dim must be 32"
        P.op(name + ".mha", "mha", [qk.ptr, qk.ptr, v.ptr, attn.ptr],
             [2 * x.C, 0, 2 * x.C, x.C, x.C, 0, x.C, x.N, T, T, self.heads], [hd ** -0.5])
        return self._linear(name + ".o_proj", attn, prefix + ".o_proj")

    def _csp_rep(self, name, x, prefix, out=None, out_coff=0):
        """RTDetrV2CSPRepLayer (hidden_expansion 1.0 -> conv3 is Identity): conv1 || conv2 fused along
        Cout, three RepVGG blocks re-parameterised to one 3x3 each, the branch sum folded into the
        last block's epilogue."""
        P = self.prog
        w1, b1 = self._convbn(prefix + ".conv1")
        w2, b2 = self._convbn(prefix + ".conv2")
        h = w1.shape[0]
        T = P.act(x.N, x.H, x.W, 2 * h)
        P.conv(name + ".conv12", x, np.concatenate([w1, w2], 0), np.concatenate([b1, b2], 0), act=ACT_SILU, y=T)
        cur, cur_off = T, 0
        # hidden_expansion != 1 (0.5 in r18vd / r34vd): a 1x1 conv3 maps the h hidden channels back to the output width
        has_c3 = (prefix + ".conv3.conv.weight") in self.sd
        nb = 0
        while ("%s.bottlenecks.%d.conv1.conv.weight" % (prefix, nb)) in self.sd:
            nb += 1
        if has_c3:
            wc, bc = self._convbn(prefix + ".conv3")
            mid = P.act(x.N, x.H, x.W, h)
            if out is None:
                out = P.act(x.N, x.H, x.W, wc.shape[0])
        else:
            if out is None:
                out = P.act(x.N, x.H, x.W, h)
            mid = out
        for i in range(nb):
            w3, b3 = self._convbn("%s.bottlenecks.%d.conv1" % (prefix, i))
            wp, bp = self._convbn("%s.bottlenecks.%d.conv2" % (prefix, i))
            w3 = w3.copy()
            w3[:, :, 1, 1] += wp[:, :, 0, 0]
            last = i == nb - 1
            dst = mid if last else P.act(x.N, x.H, x.W, h)
            off = out_coff if (last and not has_c3) else 0
            P.conv("%s.rep%d" % (name, i), cur, w3, b3 + bp, k=3, act=ACT_SILU, y=dst, y_coff=off,
                   x_coff=cur_off, resid=T if last else None, r_coff=h)
            cur, cur_off = dst, off
        if has_c3:
            P.conv(name + ".conv3", mid, wc, bc, act=ACT_SILU, y=out, y_coff=out_coff)
        return out

    # ---- the network
    def _build(self):
        P, B = self.prog, self.B
        H, W = self.IH, self.IW
        # input: bf16 NHWC with 16 channels (RGB in 0..2, zeros above): written by tcd_pil_resize or set_input
        self.x_in = P.act(B, H, W, 16)
        self._tap("input", self.x_in)

        # -- PResNet-50-vd stem: three 3x3 convs (s2, s1, s1) + 3x3 s2 max-pool
        emb = "model.backbone.model.embedder.embedder."
        c = self.x_in
        for i, (co, s) in enumerate(((32, 2), (32, 1), (64, 1))):
            w, b = self._resnet_cb(emb + str(i))
            y = P.act(B, c.H // s, c.W // s, co)
            P.conv("stem.%d" % i, c, w, b, k=3, stride=s, act=ACT_RELU, y=y)
            c = y
        y = P.act(B, c.H // 2, c.W // 2, c.C)
        P.op("stem.pool", "maxpool", [c.ptr, y.ptr], [B, c.H, c.W, c.C])
        c = y
        self._tap("stem", c)

        # -- four bottleneck stages; the 3x3 carries the stride, the shortcut of a strided block is
        #    avg-pool 2x2 + 1x1 conv ("vd"); conv3's epilogue adds the shortcut BEFORE the ReLU
        depths = []
        st = 0
        while ("model.backbone.model.encoder.stages.%d.layers.0.layer.0.convolution.weight" % st) in self.sd:
            n = 0
            while ("model.backbone.model.encoder.stages.%d.layers.%d.layer.0.convolution.weight" % (st, n)) in self.sd:
                n += 1
            depths.append(n)
            st += 1
        feats = []
        for si, depth in enumerate(depths):
            for li in range(depth):
                pre = "model.backbone.model.encoder.stages.%d.layers.%d." % (si, li)
                nm = "s%d.b%d" % (si, li)
                stride = 2 if (li == 0 and si > 0) else 1
                w1, b1 = self._resnet_cb(pre + "layer.0")
                w2, b2 = self._resnet_cb(pre + "layer.1")
                basic = (pre + "layer.2.convolution.weight") not in self.sd   # r18vd / r34vd: two 3x3 convs per block
                if basic:
                    cout = w2.shape[0]
                    t2 = P.act(B, c.H // stride, c.W // stride, w1.shape[0])
                    P.conv(nm + ".conv1", c, w1, b1, k=3, stride=stride, act=ACT_RELU, y=t2)
                    w3, b3, k3 = w2, b2, 3
                else:
                    w3, b3 = self._resnet_cb(pre + "layer.2")
                    k3 = 1
                    mid, cout = w1.shape[0], w3.shape[0]
                    t1 = P.act(B, c.H, c.W, mid)
                    P.conv(nm + ".conv1", c, w1, b1, act=ACT_RELU, y=t1)
                    t2 = P.act(B, c.H // stride, c.W // stride, mid)
                    P.conv(nm + ".conv2", t1, w2, b2, k=3, stride=stride, act=ACT_RELU, y=t2)
                sc = c
                has_sc = (pre + "shortcut.1.convolution.weight") in self.sd or (pre + "shortcut.convolution.weight") in self.sd
                if has_sc:
                    src = c
                    if stride == 2:
                        src = P.act(B, c.H // 2, c.W // 2, c.C)
                        P.op(nm + ".sc_pool", "avgpool", [c.ptr, src.ptr], [B, c.H, c.W, c.C])
                        ws, bs = self._resnet_cb(pre + "shortcut.1")
                    else:
                        ws, bs = self._resnet_cb(pre + "shortcut")
                    sc = P.act(B, src.H, src.W, cout)
                    P.conv(nm + ".shortcut", src, ws, bs, y=sc)
                out = P.act(B, t2.H, t2.W, cout)
                P.conv(nm + (".conv2" if basic else ".conv3"), t2, w3, b3, k=k3, act=ACT_RELU, y=out, resid=sc, resid_pre=1)
                c = out
            if si >= 1:
                feats.append(c)
                self._tap("backbone.%d" % si, c)
        assert len(feats) == 3, "expected three backbone feature levels"

        # -- encoder input projections (1x1 + BN) written straight into the CCFM concat buffers
        D = int(self._g("model.encoder_input_proj.0.0.weight").shape[0])
        if D != 256:
            raise ValueError("RT-DETR hidden_dim %d is not supported (r18vd / r34vd / r50vd use 256; the attention "
                             "kernels are built for 8 heads of 32)" % D)
        H8, W8, H16, W16, H32, W32 = H // 8, W // 8, H // 16, W // 16, H // 32, W // 32
        catB = P.act(B, H8, W8, 2 * D)     # [up(L1) | proj0]
        catA = P.act(B, H16, W16, 2 * D)   # [up(L2) | proj1]
        catC = P.act(B, H16, W16, 2 * D)   # [down(F0) | L1]
        catD = P.act(B, H32, W32, 2 * D)   # [down(P1) | L2]
        proj2 = P.act(B, H32, W32, D)
        for lvl, (dst, off) in enumerate(((catB, D), (catA, D), (proj2, 0))):
            w = self._g("model.encoder_input_proj.%d.0.weight" % lvl)
            pb = "model.encoder_input_proj.%d.1." % lvl
            w, b = fold_bn(w, self._g(pb + "weight"), self._g(pb + "bias"), self._g(pb + "running_mean"),
                           self._g(pb + "running_var"), self.bn_eps)
            P.conv("enc_proj.%d" % lvl, feats[lvl], w, b, y=dst, y_coff=off)

        # -- AIFI: one post-norm transformer layer over the stride-32 tokens
        T32 = H32 * W32
        x = proj2.tokens()
        pos = Act(P.upload(to_bf16_bits(sincos_pos_embed(W32, H32, D))), 1, 1, T32, D)
        ap = "model.encoder.aifi.0.layers.0"
        o = self._self_attention("aifi", x, pos, ap + ".self_attn", pos_bcast=True)
        x1 = self._layernorm("aifi.ln1", x, o, ap + ".self_attn_layer_norm")
        hdn = self._linear("aifi.fc1", x1, ap + ".mlp.fc1", act=ACT_GELU)
        f = self._linear("aifi.fc2", hdn, ap + ".mlp.fc2")
        x2 = self._layernorm("aifi.ln2", x1, f, ap + ".final_layer_norm")
        s5 = Act(x2.ptr, B, H32, W32, D)
        self._tap("aifi", s5)

        # -- CCFM top-down (FPN) then bottom-up (PAN)
        w, b = self._convbn("model.encoder.lateral_convs.0")
        P.conv("lateral.0", s5, w, b, act=ACT_SILU, y=catD, y_coff=D)                      # L2
        P.op("up.0", "upsample", [catD.ptr, catA.ptr], [B, H32, W32, D, 2 * D, D, 2 * D, 0])
        F1 = self._csp_rep("fpn.0", catA, "model.encoder.fpn_blocks.0")
        w, b = self._convbn("model.encoder.lateral_convs.1")
        P.conv("lateral.1", F1, w, b, act=ACT_SILU, y=catC, y_coff=D)                      # L1
        P.op("up.1", "upsample", [catC.ptr, catB.ptr], [B, H16, W16, D, 2 * D, D, 2 * D, 0])
        F0 = self._csp_rep("fpn.1", catB, "model.encoder.fpn_blocks.1")
        w, b = self._convbn("model.encoder.downsample_convs.0")
        P.conv("down.0", F0, w, b, k=3, stride=2, act=ACT_SILU, y=catC, y_coff=0)
        P1 = self._csp_rep("pan.0", catC, "model.encoder.pan_blocks.0")
        w, b = self._convbn("model.encoder.downsample_convs.1")
        P.conv("down.1", P1, w, b, k=3, stride=2, act=ACT_SILU, y=catD, y_coff=0)
        P2 = self._csp_rep("pan.1", catD, "model.encoder.pan_blocks.1")
        enc_out = [F0, P1, P2]
        for i, a in enumerate(enc_out):
            self._tap("encoder.%d" % i, a)

        # -- decoder input projections write the [B, S, 256] token memory level by level
        shapes = [(H8, W8), (H16, W16), (H32, W32)]
        S = sum(h * w for h, w in shapes)
        self.S = S
        mem = P.act(B, 1, S, D)
        start = 0
        for lvl, (h, w_) in enumerate(shapes):
            w = self._g("model.decoder_input_proj.%d.0.weight" % lvl)
            pb = "model.decoder_input_proj.%d.1." % lvl
            w, b = fold_bn(w, self._g(pb + "weight"), self._g(pb + "bias"), self._g(pb + "running_mean"),
                           self._g(pb + "running_var"), self.bn_eps)
            view = Act(mem.ptr + start * D * 2, B, h, w_, D, nstride=S * D)
            P.conv("dec_proj.%d" % lvl, enc_out[lvl], w, b, y=view)
            start += h * w_
        self._tap("memory", mem)

        # -- query selection
        Q, NC = self.Q, self.NC
        anchors, valid = make_anchors(shapes)
        d_anchors, d_valid = P.upload(anchors), P.upload(valid)
        masked = P.act(B, 1, S, D)
        P.op("enc.mask", "mask_rows", [mem.ptr, d_valid, masked.ptr], [B, S, D])
        eo = self._linear("enc_output.lin", masked, "model.enc_output.0")
        out_mem = self._layernorm("enc_output.ln", eo, None, "model.enc_output.1")
        enc_cls = P.f32(B * S * NC)
        self._linear("enc_score", out_mem, "model.enc_score_head", out_f32=enc_cls, out_ld=NC)
        enc_max = P.f32(B * S)
        P.op("enc.rowmax", "rowmax", [enc_cls, enc_max], [B * S, NC])
        self.topk_idx = P.alloc(B * Q * 4)
        if not self.external_topk:
            P.op("enc.topk", "topk", [enc_max, self.topk_idx], [B, S, Q])
        target = P.act(B, 1, Q, D)
        P.op("enc.gather", "gather", [out_mem.ptr, self.topk_idx, target.ptr], [B, S, Q, D])
        t = self._linear("enc_bbox.0", target, "model.enc_bbox_head.layers.0", act=ACT_RELU)
        t = self._linear("enc_bbox.1", t, "model.enc_bbox_head.layers.1", act=ACT_RELU)
        delta = P.f32(B * Q * 4)
        self._linear("enc_bbox.2", t, "model.enc_bbox_head.layers.2", out_f32=delta, out_ld=4)
        self.ref_sig = P.f32(B * Q * 4)
        ref_bf = P.act(B, 1, Q, 16)
        P.op("ref.init", "ref_init", [delta, d_anchors, self.topk_idx, self.ref_sig, ref_bf.ptr], [4, B, Q])
        self.enc_cls, self.enc_max = enc_cls, enc_max

        # -- decoder
        hs = target
        n_layers = 0
        while ("model.decoder.layers.%d.self_attn.q_proj.weight" % n_layers) in self.sd:
            n_layers += 1
        offs = P.f32(B * Q * self.heads * 3 * self.n_points * 2)
        aw = P.f32(B * Q * self.heads * 3 * self.n_points)
        self.logits = P.f32(B * Q * NC)
        LP = 3 * self.n_points
        for li in range(n_layers):
            lp = "model.decoder.layers.%d" % li
            nm = "dec%d" % li
            q1 = self._linear(nm + ".qpos.0", ref_bf, "model.decoder.query_pos_head.layers.0", act=ACT_RELU)
            qpos = self._linear(nm + ".qpos.1", q1, "model.decoder.query_pos_head.layers.1")
            o = self._self_attention(nm + ".sa", hs, qpos, lp + ".self_attn", pos_bcast=False)
            hs = self._layernorm(nm + ".ln1", hs, o, lp + ".self_attn_layer_norm")
            ca_in = P.act(B, 1, Q, D)
            P.op(nm + ".ca_in", "add_rows", [hs.ptr, qpos.ptr, ca_in.ptr], [B, Q, D, 0])
            self._linear(nm + ".offsets", ca_in, lp + ".encoder_attn.sampling_offsets", out_f32=offs,
                         out_ld=self.heads * LP * 2)
            self._linear(nm + ".attn_w", ca_in, lp + ".encoder_attn.attention_weights", out_f32=aw,
                         out_ld=self.heads * LP)
            value = self._linear(nm + ".value", mem, lp + ".encoder_attn.value_proj")
            sampled = P.act(B, 1, Q, D)
            P.op(nm + ".msda", "msda", [value.ptr, offs, aw, self.ref_sig, sampled.ptr],
                 [B, Q, S, self.heads, 3, self.n_points, H8, W8, H16, W16, H32, W32], [self.offset_scale])
            o2 = self._linear(nm + ".ca_out", sampled, lp + ".encoder_attn.output_proj")
            hs = self._layernorm(nm + ".ln2", hs, o2, lp + ".encoder_attn_layer_norm")
            f1 = self._linear(nm + ".fc1", hs, lp + ".mlp.fc1", act=ACT_RELU)
            f2 = self._linear(nm + ".fc2", f1, lp + ".mlp.fc2")
            hs = self._layernorm(nm + ".ln3", hs, f2, lp + ".final_layer_norm")
            self._tap(nm, hs)
            bp = "model.decoder.bbox_embed.%d.layers." % li
            t = self._linear(nm + ".bbox.0", hs, bp + "0", act=ACT_RELU)
            t = self._linear(nm + ".bbox.1", t, bp + "1", act=ACT_RELU)
            self._linear(nm + ".bbox.2", t, bp + "2", out_f32=delta, out_ld=4)
            P.op(nm + ".ref", "ref_update", [delta, self.ref_sig, ref_bf.ptr], [4, B * Q])
            if li == n_layers - 1:
                self._linear(nm + ".class", hs, "model.decoder.class_embed.%d" % li, out_f32=self.logits, out_ld=NC)

        # -- postprocessor: top-Q over queries x classes, boxes in original pixels
        self.post_idx = P.alloc(B * Q * 4)
        P.op("post.topk", "topk", [self.logits, self.post_idx], [B, Q * NC, Q])
        self.orig_wh = P.f32(B * 2)
        self.out_labels, self.out_boxes, self.out_scores = P.alloc(B * Q * 4), P.f32(B * Q * 4), P.f32(B * Q)
        P.op("post.decode", "detr_post", [self.logits, self.ref_sig, self.post_idx, self.orig_wh, self.out_labels,
                                           self.out_boxes, self.out_scores], [B, Q, NC, Q])

    # ---- running
    def set_input(self, x_nchw):
        """x_nchw: float32 [B,3,H,W] in [0,1] RGB (what RTDETRDetector.preprocess returns) -> device input."""
        x = np.asarray(x_nchw, np.float32)
        assert x.shape == (self.B, 3, self.IH, self.IW), x.shape
        buf = np.zeros((self.B, self.IH, self.IW, 16), np.uint16)
        buf[..., :3] = to_bf16_bits(x.transpose(0, 2, 3, 1))
        self.prog.h2d(self.x_in.ptr, buf)

    def set_topk(self, idx):
        assert self.external_topk, "engine was built with its own query selection"
        self.prog.h2d(self.topk_idx, np.asarray(idx, np.int32).reshape(self.B, self.Q))

    def set_orig_sizes(self, wh):
        self.prog.h2d(self.orig_wh, np.asarray(wh, np.float32).reshape(self.B, 2))

    def run(self, graph=True):
        self.prog.run(graph)
        self.prog.sync()

    def outputs(self):
        """-> labels int32 [B,Q], boxes f32 [B,Q,4] (xyxy, original pixels), scores f32 [B,Q]; score-descending."""
        P, B, Q = self.prog, self.B, self.Q
        return (P.d2h(self.out_labels, (B, Q), np.int32), P.d2h(self.out_boxes, (B, Q, 4), np.float32),
                P.d2h(self.out_scores, (B, Q), np.float32))

    def raw_outputs(self):
        """-> logits f32 [B,Q,NC], pred_boxes f32 [B,Q,4] (cxcywh in [0,1]), encoder top-k indices [B,Q]."""
        P, B, Q = self.prog, self.B, self.Q
        return (P.d2h(self.logits, (B, Q, self.NC), np.float32), P.d2h(self.ref_sig, (B, Q, 4), np.float32),
                P.d2h(self.topk_idx, (B, Q), np.int32))

    def read_tap(self, name):
        return self.prog.read_bf16(self.taps[name])
