"""Unit parity of the RT-DETR-side kernels (through the tcd_program C ABI) against plain torch fp32
on the same bf16-rounded operands.  Tolerances: one bf16 output rounding (rel 2^-8) plus fp32
summation-order noise; integer / index work is exact."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bf16(t):
    import torch
    return t.to(torch.bfloat16).float()


def _prog():
    from telescope_cam_detection_b200.rtdetr_engine import Program
    return Program(0)


def _up_bf16(P, t, shape4):
    from telescope_cam_detection_b200.rtdetr_engine import Act, to_bf16_bits
    ptr = P.upload(to_bf16_bits(t.numpy()))
    return Act(ptr, *shape4)


def test_layernorm_with_residual():
    import torch
    torch.manual_seed(0)
    P = _prog()
    a, b = _bf16(torch.randn(3, 50, 256)), _bf16(torch.randn(3, 50, 256))
    g, be = torch.rand(256) + 0.5, torch.randn(256) * 0.1
    A, Bb = _up_bf16(P, a, (3, 1, 50, 256)), _up_bf16(P, b, (3, 1, 50, 256))
    Y = P.act(3, 1, 50, 256)
    P.op("ln", "layernorm", [A.ptr, Bb.ptr, P.upload(g.numpy()), P.upload(be.numpy()), Y.ptr], [150, 256], [1e-5])
    P.run(False); P.sync()
    ref = torch.nn.functional.layer_norm(a + b, (256,), g, be, 1e-5)
    got = torch.from_numpy(P.read_bf16(Y))[:, 0]
    assert torch.allclose(got, ref, atol=2e-2, rtol=1e-2)
    P.close()


@pytest.mark.parametrize("T", [400, 300, 77])
def test_mha_matches_torch(T):
    import torch
    torch.manual_seed(1)
    P = _prog()
    N, H, D = 2, 8, 32
    qk = _bf16(torch.randn(N, T, 2 * H * D))
    v = _bf16(torch.randn(N, T, H * D))
    QK, V = _up_bf16(P, qk, (N, 1, T, 512)), _up_bf16(P, v, (N, 1, T, 256))
    O = P.act(N, 1, T, 256)
    P.op("mha", "mha", [QK.ptr, QK.ptr, V.ptr, O.ptr], [512, 0, 512, 256, 256, 0, 256, N, T, T, H], [D ** -0.5])
    P.run(False); P.sync()
    q = qk[..., :256].view(N, T, H, D).transpose(1, 2)
    k = qk[..., 256:].view(N, T, H, D).transpose(1, 2)
    vv = v.view(N, T, H, D).transpose(1, 2)
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, vv).transpose(1, 2).reshape(N, T, 256)
    got = torch.from_numpy(P.read_bf16(O))[:, 0]
    assert torch.allclose(got, ref, atol=1.5e-2, rtol=1e-2), float((got - ref).abs().max())
    P.close()


def test_msda_matches_transformers_function():
    import torch
    from transformers.models.rt_detr_v2.modeling_rt_detr_v2 import multi_scale_deformable_attention_v2 as msda_ref
    torch.manual_seed(2)
    P = _prog()
    N, Q, H, D, L, Pn = 2, 37, 8, 32, 3, 4
    shapes = [(20, 24), (10, 12), (5, 6)]
    S = sum(h * w for h, w in shapes)
    value = _bf16(torch.randn(N, S, H * D))
    offsets = torch.randn(N, Q, H, L * Pn, 2) * 2.0
    logits = torch.randn(N, Q, H, L * Pn)
    ref_pts = torch.rand(N, Q, 4) * torch.tensor([1.2, 1.2, 0.4, 0.4]) - torch.tensor([0.1, 0.1, 0.0, 0.0])
    Vv = _up_bf16(P, value, (N, 1, S, H * D))
    O = P.act(N, 1, Q, H * D)
    P.op("msda", "msda", [Vv.ptr, P.upload(offsets.numpy()), P.upload(logits.numpy()), P.upload(ref_pts.numpy()), O.ptr],
         [N, Q, S, H, L, Pn, 20, 24, 10, 12, 5, 6], [0.5])
    P.run(False); P.sync()
    n_points_scale = torch.tensor([1.0 / Pn] * (L * Pn)).unsqueeze(-1)
    off = offsets * n_points_scale * ref_pts[:, :, None, None, 2:] * 0.5
    loc = ref_pts[:, :, None, None, :2] + off                     # [N,Q,H,L*P,2]
    attn = torch.softmax(logits, -1)
    ref = msda_ref(value.view(N, S, H, D), shapes, loc, attn, [Pn] * L, "default")
    got = torch.from_numpy(P.read_bf16(O))[:, 0]
    assert torch.allclose(got, ref, atol=1.5e-2, rtol=1e-2), float((got - ref).abs().max())
    P.close()


def test_topk_exact_with_ties():
    import torch
    torch.manual_seed(3)
    P = _prog()
    N, n, k = 3, 24000, 300
    keys = torch.randn(N, n)
    keys[1] = torch.round(keys[1] * 4) / 4          # heavy ties
    keys[2, ::7] = 5.0
    idx = P.alloc(N * k * 4)
    P.op("topk", "topk", [P.upload(keys.numpy()), idx], [N, n, k])
    P.run(False); P.sync()
    got = P.d2h(idx, (N, k), np.int32)
    for b in range(N):
        order = np.lexsort((np.arange(n), -keys[b].numpy()))[:k]   # value desc, index asc
        assert np.array_equal(got[b], order), b
    P.close()


def test_pools_and_gather_and_rowmax():
    import torch
    torch.manual_seed(4)
    P = _prog()
    x = _bf16(torch.randn(2, 16, 24, 64))
    X = _up_bf16(P, x, (2, 16, 24, 64))
    Ym, Ya = P.act(2, 8, 12, 64), P.act(2, 8, 12, 64)
    P.op("max", "maxpool", [X.ptr, Ym.ptr], [2, 16, 24, 64])
    P.op("avg", "avgpool", [X.ptr, Ya.ptr], [2, 16, 24, 64])
    f = torch.randn(2 * 50, 80)
    rm = P.f32(100)
    P.op("rowmax", "rowmax", [P.upload(f.numpy()), rm], [100, 80])
    src = _bf16(torch.randn(2, 50, 256))
    Sx = _up_bf16(P, src, (2, 1, 50, 256))
    gi = torch.randint(0, 50, (2, 9), dtype=torch.int32)
    G = P.act(2, 1, 9, 256)
    P.op("gather", "gather", [Sx.ptr, P.upload(gi.numpy()), G.ptr], [2, 50, 9, 256])
    P.run(False); P.sync()
    xn = x.permute(0, 3, 1, 2)
    assert torch.equal(torch.from_numpy(P.read_bf16(Ym)), torch.nn.functional.max_pool2d(xn, 3, 2, 1).permute(0, 2, 3, 1))
    assert torch.allclose(torch.from_numpy(P.read_bf16(Ya)), torch.nn.functional.avg_pool2d(xn, 2, 2).permute(0, 2, 3, 1),
                          atol=1e-2, rtol=1e-2)
    assert np.array_equal(P.d2h(rm, (100,), np.float32), f.max(1).values.numpy())
    got = torch.from_numpy(P.read_bf16(G))[:, 0]
    assert torch.equal(got, torch.stack([src[b][gi[b].long()] for b in range(2)]))
    P.close()


def test_ref_update_and_detr_post():
    import torch
    torch.manual_seed(5)
    P = _prog()
    N, Q, C, k = 2, 300, 80, 300
    ref0 = torch.rand(N * Q, 4) * 0.8 + 0.1
    delta = torch.randn(N * Q, 4) * 0.3
    d_ref = P.upload(ref0.numpy())
    rb = P.act(N, 1, Q, 16)
    P.op("upd", "ref_update", [P.upload(delta.numpy()), d_ref, rb.ptr], [4, N * Q])
    logits = torch.randn(N, Q, C)
    idx = torch.stack([torch.topk(logits[b].flatten(), k).indices for b in range(N)]).int()
    wh = torch.tensor([[1920.0, 1080.0], [640.0, 480.0]])
    lab, box, sc = P.alloc(N * k * 4), P.f32(N * k * 4), P.f32(N * k)
    P.op("post", "detr_post", [P.upload(logits.numpy()), d_ref, P.upload(idx.numpy()), P.upload(wh.numpy()), lab, box, sc],
         [N, Q, C, k])
    P.run(False); P.sync()
    x = ref0.clamp(0, 1)
    new = torch.sigmoid(delta + torch.log(x.clamp(min=1e-5) / (1 - x).clamp(min=1e-5)))
    got = torch.from_numpy(P.d2h(d_ref, (N * Q, 4), np.float32))
    assert torch.allclose(got, new, atol=1e-6)
    from oracle.rtdetr import postprocess
    rl, rbx, rs = postprocess(logits, new.view(N, Q, 4), wh.long())
    assert np.array_equal(P.d2h(lab, (N, k), np.int32), rl.numpy())
    assert np.allclose(P.d2h(sc, (N, k), np.float32), rs.numpy(), atol=1e-6)
    assert np.allclose(P.d2h(box, (N, k, 4), np.float32), rbx.numpy(), atol=2e-3)
    assert np.allclose(P.read_bf16(rb)[:, 0, :, :4].reshape(-1, 4), _bf16(new).numpy())
    P.close()


CONV_CASES = [
    # N, H, W, cin, cout, k, stride, act, resid_pre
    (2, 40, 40, 64, 256, 1, 1, 2, 1),     # ResNet conv3: relu(conv + shortcut)
    (2, 64, 64, 3, 32, 3, 2, 2, 0),       # ResNet stem conv0 (3 real channels of 16)
    (2, 1, 300, 256, 1024, 1, 1, 3, 0),   # AIFI fc1 (GELU) as a linear over tokens
    (2, 20, 20, 256, 256, 3, 1, 1, 0),    # RepVGG 3x3 + SiLU
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv_variants(case):
    import torch
    import torch.nn.functional as F
    from telescope_cam_detection_b200.rtdetr_engine import Act
    N, H, W, cin, cout, k, stride, act, resid_pre = case
    torch.manual_seed(6 + cin + cout)
    P = _prog()
    cpad = (cin + 15) // 16 * 16
    x = torch.zeros(N, H, W, cpad)
    x[..., :cin] = _bf16(torch.randn(N, H, W, cin))
    w = torch.randn(cout, cin, k, k) / (cin * k * k) ** 0.5
    b = torch.randn(cout) * 0.1
    X = _up_bf16(P, x, (N, H, W, cpad))
    Ho, Wo = H // stride, W // stride
    Y = P.act(N, Ho, Wo, cout)
    r = _bf16(torch.randn(N, Ho, Wo, cout)) if resid_pre else None
    R = _up_bf16(P, r, (N, Ho, Wo, cout)) if resid_pre else None
    P.conv("c", X, w.numpy(), b.numpy(), k=k, stride=stride, act=act, y=Y, resid=R, resid_pre=resid_pre)
    P.run(False); P.sync()
    ref = F.conv2d(x[..., :cin].permute(0, 3, 1, 2), _bf16(w), b, stride=stride, padding=(k - 1) // 2)
    if resid_pre:
        ref = ref + r.permute(0, 3, 1, 2)
    ref = {0: lambda t: t, 1: F.silu, 2: F.relu, 3: F.gelu}[act](ref)
    got = torch.from_numpy(P.read_bf16(Y)).permute(0, 3, 1, 2)
    err = (got - ref).abs()
    assert bool((err <= 1e-2 + 8e-3 * ref.abs()).all()), float(err.max())
    P.close()


def test_linear_f32_rows_and_strided_level_view():
    """f32 row output (logits / offsets) and a conv writing one pyramid level inside [B, S, C] tokens."""
    import torch
    from telescope_cam_detection_b200.rtdetr_engine import Act
    torch.manual_seed(9)
    P = _prog()
    N, T = 2, 300
    x = _bf16(torch.randn(N, T, 256))
    X = _up_bf16(P, x, (N, 1, T, 256))
    w = torch.randn(192, 256) / 16
    b = torch.randn(192) * 0.1
    out = P.f32(N * T * 192)
    P.conv("lin", X, w.numpy().reshape(192, 256, 1, 1), b.numpy(), out_f32=out, out_ld=192)
    w4 = torch.randn(4, 256) / 16
    out4 = P.f32(N * T * 4)
    P.conv("lin4", X, w4.numpy().reshape(4, 256, 1, 1), np.zeros(4, np.float32), out_f32=out4, out_ld=4)
    # level view: 10x12 map written at token offset 50 of a [N, 200, 64] memory
    m = _bf16(torch.randn(N, 10, 12, 64))
    M = _up_bf16(P, m, (N, 10, 12, 64))
    mem = P.act(N, 1, 200, 64)
    wv = torch.randn(64, 64) / 8
    view = Act(mem.ptr + 50 * 64 * 2, N, 10, 12, 64, nstride=200 * 64)
    P.conv("lvl", M, wv.numpy().reshape(64, 64, 1, 1), np.zeros(64, np.float32), y=view)
    P.run(False); P.sync()
    ref = x @ _bf16(w).T + b
    got = torch.from_numpy(P.d2h(out, (N, T, 192), np.float32))
    assert torch.allclose(got, ref, atol=2e-3, rtol=1e-3)
    got4 = torch.from_numpy(P.d2h(out4, (N, T, 4), np.float32))
    assert torch.allclose(got4, x @ _bf16(w4).T, atol=2e-3, rtol=1e-3)
    memv = torch.from_numpy(P.read_bf16(mem))[:, 0]
    refm = (m.view(N, 120, 64) @ _bf16(wv).T)
    assert torch.allclose(memv[:, 50:170], refm, atol=2e-2, rtol=1e-2)
    assert float(memv[:, :50].abs().max()) == 0.0 and float(memv[:, 170:].abs().max()) == 0.0
    P.close()


@pytest.mark.parametrize("shape", [(1080, 1920, 640, 640), (270, 480, 160, 192), (100, 64, 128, 96)])
def test_pil_resize_bit_exact_vs_pillow(shape):
    import ctypes as C
    from PIL import Image
    from telescope_cam_detection_b200 import _lib
    h, w, oh, ow = shape
    rng = np.random.default_rng(h + w)
    frames = [rng.integers(0, 255, (h, w, 3), dtype=np.uint8) for _ in range(2)]
    P = _prog()
    out = P.act(2, oh, ow, 16)
    u8 = np.empty((2, oh, ow, 3), np.uint8)
    lib = P.lib
    _lib.check(lib.tcd_pil_resize(_lib.ptr_array([f.ctypes.data for f in frames]), _lib.i32_array([h, h]),
                                  _lib.i32_array([w, w]), _lib.i64_array([w * 3, w * 3]), 2, 0, oh, ow,
                                  C.c_void_p(out.ptr), u8.ctypes.data_as(C.c_void_p), C.c_void_p(P.stream())), "pil")
    for i, f in enumerate(frames):
        ref = np.asarray(Image.fromarray(np.ascontiguousarray(f[:, :, ::-1])).resize((ow, oh), Image.BILINEAR))
        assert np.array_equal(u8[i], ref), int((u8[i] != ref).sum())
    import torch
    got = P.read_bf16(out)
    ref_t = torch.from_numpy(u8.astype(np.float32) / 255.0).to(torch.bfloat16).float().numpy()
    assert np.array_equal(got[..., :3], ref_t) and float(np.abs(got[..., 3:]).max()) == 0.0
    P.close()
