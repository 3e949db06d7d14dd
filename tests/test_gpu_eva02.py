"""Stage-2 classifier body (src/species_classifier.py:258-262,386-391): the EVA02 layer
program and its kernels against the fp32 restatement oracle/eva02.py (parity unpinned: timm is absent) and against
plain torch fp32 for the single kernels.

Tolerances: a kernel is compared on bf16-rounded operands and may differ by one bf16 output rounding (rel 2^-8) plus
fp32 summation-order noise.  The network runs bf16 storage / fp32 accumulate against an fp32 oracle: class
probabilities within 1e-2 absolute (north_star's score tolerance), the top-1 class equal wherever the oracle's margin
exceeds that, logits within the floor the bf16 rounding-point emulation (oracle lowp=True) itself shows x 3.
"""
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
    return Act(P.upload(to_bf16_bits(t.numpy())), *shape4)


def _rot(x):
    import torch
    return torch.stack([-x[..., 1::2], x[..., ::2]], -1).reshape(x.shape)


@pytest.mark.parametrize("N,T,H,rope", [(2, 257, 3, True), (1, 577, 16, True), (3, 65, 5, True), (2, 100, 2, False)])
def test_mha_rope_matches_torch(N, T, H, rope):
    import torch
    from telescope_cam_detection_b200.eva02_engine import rope_cos_sin
    torch.manual_seed(3)
    P = _prog()
    D = 64 * H
    qkv = _bf16(torch.randn(N, T, 3 * D))
    QKV = _up_bf16(P, qkv, (N, 1, T, 3 * D))
    O = P.act(N, 1, T, D)
    G = int(round((T - 1) ** 0.5))
    tab = rope_cos_sin(G, 16) if rope else None
    P.op("mha", "mha_rope", [QKV.ptr, QKV.ptr, QKV.ptr, O.ptr, P.upload(tab) if rope else 0],
         [3 * D, 0, 3 * D, D, 3 * D, 2 * D, D, N, T, H, 1], [64 ** -0.5])
    P.run(False); P.sync()
    q, k, v = [qkv[..., i * D:(i + 1) * D].view(N, T, H, 64).transpose(1, 2) for i in range(3)]
    if rope:
        cos = torch.from_numpy(tab[..., 0]).repeat_interleave(2, -1)
        sin = torch.from_numpy(tab[..., 1]).repeat_interleave(2, -1)
        q = torch.cat([q[:, :, :1], q[:, :, 1:] * cos + _rot(q[:, :, 1:]) * sin], 2)
        k = torch.cat([k[:, :, :1], k[:, :, 1:] * cos + _rot(k[:, :, 1:]) * sin], 2)
    ref = torch.nn.functional.scaled_dot_product_attention(_bf16(q), _bf16(k), v).transpose(1, 2).reshape(N, T, D)
    got = torch.from_numpy(P.read_bf16(O))[:, 0]
    assert torch.allclose(got, ref, atol=1.5e-2, rtol=1e-2), float((got - ref).abs().max())
    P.close()


@pytest.mark.parametrize("H", [2730, 512, 853])
def test_swiglu_layernorm(H):
    import torch
    torch.manual_seed(4)
    P = _prog()
    rows, Hp = 37, (H + 127) // 128 * 128
    gu = torch.zeros(rows, 2 * Hp)
    gu[:, :H] = torch.randn(rows, H) * 2
    gu[:, Hp:Hp + H] = torch.randn(rows, H)
    gu[:, H:Hp] = 7.0          # garbage in the padding columns must not leak into the statistics
    gu = _bf16(gu)
    g, b = torch.rand(H) + 0.5, torch.randn(H) * 0.1
    GU = _up_bf16(P, gu, (1, 1, rows, 2 * Hp))
    Y = P.act(1, 1, rows, Hp)
    gp, bp = np.zeros(Hp, np.float32), np.zeros(Hp, np.float32)
    gp[:H], bp[:H] = g.numpy(), b.numpy()
    P.op("sw", "swiglu_ln", [GU.ptr, P.upload(gp), P.upload(bp), Y.ptr], [rows, H, Hp, 0], [1e-6])
    P.run(False); P.sync()
    h = torch.nn.functional.silu(gu[:, :H]) * gu[:, Hp:Hp + H]
    ref = torch.nn.functional.layer_norm(h, (H,), g, b, 1e-6)
    got = torch.from_numpy(P.read_bf16(Y))[0, 0]
    assert torch.allclose(got[:, :H], ref, atol=2e-2, rtol=1e-2), float((got[:, :H] - ref).abs().max())
    assert float(got[:, H:].abs().max()) == 0.0 if Hp > H else True
    P.close()


def test_token_assembly_softmax_topk():
    import torch
    torch.manual_seed(5)
    P = _prog()
    N, T, D = 3, 17, 192
    emb = _bf16(torch.randn(N, T - 1, D))
    pos = torch.randn(T, D)
    E = _up_bf16(P, emb, (N, 1, T - 1, D))
    X = P.act(N, 1, T, D)
    P.op("tok", "vit_assemble", [E.ptr, P.upload(pos.numpy()), X.ptr], [N, T, D])
    C, ld, k = 1000, 1024, 7
    logits = torch.randn(N, ld) * 3
    lg = P.upload(logits.numpy())
    probs, idx, top = P.f32(N * C), P.alloc(N * k * 4), P.f32(N * k)
    P.op("sm", "softmax", [lg, probs], [N, C, ld, C])
    P.op("tk", "topk", [probs, idx], [N, C, k])
    P.op("take", "take", [probs, idx, top], [N, k, C])
    P.run(False); P.sync()
    ref = torch.cat([pos[:1].expand(N, 1, D), emb + pos[1:]], 1)
    got = torch.from_numpy(P.read_bf16(X))[:, 0]
    assert torch.allclose(got, ref, atol=2e-2, rtol=1e-2)
    pr = torch.softmax(logits[:, :C], 1)
    gp = torch.from_numpy(P.d2h(probs, (N, C), np.float32))
    assert torch.allclose(gp, pr, atol=1e-6, rtol=1e-5)
    tp, ti = torch.topk(gp, k, 1)   # the read-out must be the top-k of the probabilities the device holds
    assert np.array_equal(P.d2h(idx, (N, k), np.int32), ti.numpy().astype(np.int32))
    assert np.array_equal(P.d2h(top, (N, k), np.float32), tp.numpy())
    P.close()


@pytest.mark.parametrize("rows,cin,cout,act,resid", [(20000 + 37, 1024, 512, 0, True), (19200, 512, 768, 1, False),
                                                     (9700, 2816, 1024, 0, True)])
def test_token_linear_on_cta_pairs(rows, cin, cout, act, resid):
    """Long-K token linears with enough tiles take the cta_group::2 GEMM (csrc/gemm_pair.cuh: 256 x 256 per CTA pair):
    odd tile counts, a ragged last tile, bias, SiLU, residual, a column slice of a wider output."""
    import torch
    from telescope_cam_detection_b200.rtdetr_engine import Act
    torch.manual_seed(7)
    P = _prog()
    x = _bf16(torch.randn(rows, cin))
    w = _bf16(torch.randn(cout, cin) / cin ** 0.5)
    b = torch.randn(cout) * 0.1
    r = _bf16(torch.randn(rows, cout + 256)) if resid else None
    X = _up_bf16(P, x, (1, 1, rows, cin))
    Y = Act(P.alloc(rows * (cout + 256) * 2), 1, 1, rows, cout + 256)
    R = _up_bf16(P, r, (1, 1, rows, cout + 256)) if resid else None
    P.conv("lin", X, w.numpy().reshape(cout, cin, 1, 1), b.numpy(), act=act, y=Y, y_coff=256, resid=R, r_coff=256 if resid else 0)
    P.run(False); P.sync()
    got = torch.from_numpy(P.read_bf16(Y))[0, 0]
    ref = x @ w.T + b
    if act == 1:
        ref = torch.nn.functional.silu(ref)
    if resid:
        ref = ref + r[:, 256:]
    assert torch.allclose(got[:, 256:], ref, atol=3e-2, rtol=1.5e-2), float((got[:, 256:] - ref).abs().max())
    assert float(got[:, :256].abs().max()) == 0.0      # the columns left of the slice are untouched
    P.close()


@pytest.mark.parametrize("H,Hp,act", [(1024, 1024, 0), (2730, 2816, 0), (512, 512, 4)])
def test_layernorm_applied_in_the_pair_gemm_epilogue(H, Hp, act):
    """row_stats (8 bytes per row) + a CTA-pair GEMM that reads the UN-normalised rows and normalises in its epilogue:
    y = rstd * (x W'^T - mean * colsum(W')) + b'  ==  Linear(LayerNorm(x)) with the affine part folded into W', b'."""
    import torch
    from telescope_cam_detection_b200.eva02_engine import _colsum_bf16
    from telescope_cam_detection_b200.rtdetr_engine import Act
    torch.manual_seed(9)
    P = _prog()
    rows, cout = 19000 + 3, 512
    x = torch.zeros(rows, Hp)
    x[:, :H] = torch.randn(rows, H) * 2.0 + 0.4           # a row mean the correction term has to remove
    x = _bf16(x)
    g, be = torch.rand(H) + 0.5, torch.randn(H) * 0.1
    n_out = cout * (2 if act == 4 else 1)
    w = torch.randn(n_out, H) / H ** 0.5
    b = torch.randn(n_out) * 0.1
    wf = torch.zeros(n_out, Hp)
    wf[:, :H] = _bf16(w * g[None])                        # what the engine packs: W diag(gamma), rounded like the GEMM operand
    bf = b + w @ be
    X = _up_bf16(P, x, (1, 1, rows, Hp))
    stats = P.f32(rows * 2)
    P.op("stats", "row_stats", [X.ptr, stats], [rows, H, Hp], [1e-6])
    Y = Act(P.alloc(rows * cout * 2), 1, 1, rows, cout)
    P.conv("lin", X, wf.numpy().reshape(n_out, Hp, 1, 1), bf.numpy(), act=act, y=Y, ln_stats=stats, ln_colsum=_colsum_bf16(wf.numpy()))
    P.run(False); P.sync()
    st = torch.from_numpy(P.d2h(stats, (rows, 2), np.float32))
    xv = x[:, :H]
    assert torch.allclose(st[:, 0], xv.mean(1), atol=1e-4) and torch.allclose(st[:, 1], (xv.var(1, unbiased=False) + 1e-6).rsqrt(), rtol=1e-4)
    ln = (xv - xv.mean(1, keepdim=True)) * (xv.var(1, unbiased=False, keepdim=True) + 1e-6).rsqrt()
    ref = ln @ wf[:, :H].T + bf
    if act == 4:          # interleaved 32-column slabs: gate / value
        r4 = ref.view(rows, n_out // 64, 2, 32)
        ref = (torch.nn.functional.silu(r4[:, :, 0]) * r4[:, :, 1]).reshape(rows, cout)
    got = torch.from_numpy(P.read_bf16(Y))[0, 0]
    assert torch.allclose(got, ref, atol=4e-2, rtol=2e-2), float((got - ref).abs().max())
    P.close()


def test_gated_fc1_on_cta_pairs_and_layernorm():
    """fc1_g || fc1_x with the gated epilogue (act 4: interleaved 32-column slabs, silu(g) * u written directly) followed
    by the LayerNorm-only mode of swiglu_ln, against torch fp32 on the same bf16 operands."""
    import torch
    from telescope_cam_detection_b200.rtdetr_engine import Act
    torch.manual_seed(8)
    P = _prog()
    rows, D, H = 19000 + 5, 512, 300
    Hp = 384
    x = _bf16(torch.randn(rows, D))
    wg, wx = _bf16(torch.randn(H, D) / D ** 0.5), _bf16(torch.randn(H, D) / D ** 0.5)
    bg, bx = torch.randn(H) * 0.1, torch.randn(H) * 0.1
    j = np.arange(Hp)
    gcol = (j // 128) * 256 + ((j % 128) // 32) * 64 + (j % 32)
    w1, b1 = np.zeros((2 * Hp, D), np.float32), np.zeros(2 * Hp, np.float32)
    w1[gcol[:H]], w1[gcol[:H] + 32] = wg.numpy(), wx.numpy()
    b1[gcol[:H]], b1[gcol[:H] + 32] = bg.numpy(), bx.numpy()
    X = _up_bf16(P, x, (1, 1, rows, D))
    Hh = Act(P.alloc(rows * Hp * 2), 1, 1, rows, Hp)
    Y = P.act(1, 1, rows, Hp)
    P.conv("fc1", X, w1.reshape(2 * Hp, D, 1, 1), b1, act=4, y=Hh)
    g, b = torch.rand(H) + 0.5, torch.randn(H) * 0.1
    gp, bp = np.zeros(Hp, np.float32), np.zeros(Hp, np.float32)
    gp[:H], bp[:H] = g.numpy(), b.numpy()
    P.op("ln", "swiglu_ln", [Hh.ptr, P.upload(gp), P.upload(bp), Y.ptr], [rows, H, Hp, 1], [1e-6])
    P.run(False); P.sync()
    h = torch.nn.functional.silu(x @ wg.T + bg) * (x @ wx.T + bx)
    got_h = torch.from_numpy(P.read_bf16(Hh))[0, 0]
    assert torch.allclose(got_h[:, :H], h, atol=3e-2, rtol=1.5e-2), float((got_h[:, :H] - h).abs().max())
    assert float(got_h[:, H:].abs().max()) == 0.0          # silu(0) * 0 in the padded hidden columns
    ref = torch.nn.functional.layer_norm(got_h[:, :H], (H,), g, b, 1e-6)
    got = torch.from_numpy(P.read_bf16(Y))[0, 0]
    assert torch.allclose(got[:, :H], ref, atol=2e-2, rtol=1e-2), float((got[:, :H] - ref).abs().max())
    P.close()


@pytest.mark.parametrize("B,K,NC", [(64, 1024, 10000), (5, 192, 1000), (33, 320, 77)])
def test_head_gemv(B, K, NC):
    import torch
    from telescope_cam_detection_b200.rtdetr_engine import to_bf16_bits
    torch.manual_seed(6)
    P = _prog()
    x, w, b = _bf16(torch.randn(B, K)), _bf16(torch.randn(NC, K) * 0.1), torch.randn(NC)
    ld = NC + 24
    out = P.f32(B * ld)
    P.op("head", "head_gemv", [P.upload(to_bf16_bits(x.numpy())), P.upload(to_bf16_bits(w.numpy())), P.upload(b.numpy()), out],
         [B, K, NC, ld])
    P.run(False); P.sync()
    got = torch.from_numpy(P.d2h(out, (B, ld), np.float32))
    ref = x.double() @ w.double().T + b.double()
    assert torch.allclose(got[:, :NC].double(), ref, atol=2e-4, rtol=1e-5), float((got[:, :NC].double() - ref).abs().max())
    assert float(got[:, NC:].abs().max()) == 0.0     # nothing written beyond the classes
    P.close()


def test_crop_patches_equal_crop_resize_rows():
    """tcd_crop_patches == tcd_crop_resize (the tested Stage-2 gather) re-laid out as patch rows and rounded to bf16."""
    import torch
    from telescope_cam_detection_b200 import engine
    from telescope_cam_detection_b200.rtdetr_engine import to_bf16_bits
    from telescope_cam_detection_b200.eva02_engine import Eva02Engine
    from oracle import eva02
    from util import noise_frame
    cfg = eva02.Eva02Config(img_size=112, embed_dim=64, depth=1, num_heads=1, num_classes=40)
    eng = Eva02Engine(eva02.state_dict_numpy(eva02.make_model(cfg, 0)), batch=4, img_size=112, top_k=5)
    frame = torch.from_numpy(noise_frame(2, 480, 640)).cuda()
    boxes = np.array([[10, 20, 200, 300], [0, 0, 640, 480], [300, 100, 412, 212], [5, 5, 45, 60]], np.int32)
    n = eng.crop_patches(frame.data_ptr(), 480, 640, frame.stride(0), boxes)
    eng.sync()
    rows = eng.prog.d2h(eng.rows0.ptr, (4, 64, eng.Kp), np.uint16)
    ref = torch.empty((4, 3, 112, 112), dtype=torch.float32, device="cuda")
    engine.crop_resize(frame.data_ptr(), 480, 640, frame.stride(0), boxes, 112, ref.data_ptr(), layout=0)
    torch.cuda.synchronize()
    r = ref.cpu().view(4, 3, 8, 14, 8, 14).permute(0, 2, 4, 1, 3, 5).reshape(4, 64, 588).numpy()
    want = to_bf16_bits(r)
    # same arithmetic in two kernels: equal up to an fp32 contraction difference that flips a bf16 rounding
    assert n == 4 and (rows[:, :, :588] != want).mean() < 1e-3
    back = (rows[:, :, :588].astype(np.uint32) << 16).view(np.float32)
    assert np.allclose(back, r, atol=1e-2, rtol=8e-3)
    assert not rows[:, :, 588:].any()
    # the f32 patchify entry is a pure re-layout + rounding of the NCHW tensor: exact
    eng.patchify_device(ref.data_ptr(), 4)
    eng.sync()
    assert np.array_equal(eng.prog.d2h(eng.rows0.ptr, (4, 64, eng.Kp), np.uint16)[:, :, :588], want)
    eng.close()


def _compare(cfg, batch, n, seed=0, fused=True, tap_tol=0.06):
    import torch
    from telescope_cam_detection_b200.eva02_engine import Eva02Engine
    from oracle import eva02
    cfg.qkv_fused = fused
    model = eva02.make_model(cfg, seed)
    lowp = eva02.make_model(cfg, seed, lowp=True)
    x = eva02.synthetic_input(cfg, n, seed=seed + 1)
    taps = {}
    with torch.no_grad():
        ref = model(x, taps)
        floor = float((lowp(x) - ref).abs().max())
    eng = Eva02Engine(eva02.state_dict_numpy(model), batch=batch, img_size=cfg.img_size, top_k=5, keep_taps=True)
    got = torch.from_numpy(eng.forward(x.numpy()))
    # residual stream after every block: relative to the stream's own scale
    for name in ["tokens"] + ["block%d" % i for i in range(cfg.depth)]:
        t = torch.from_numpy(eng.read_tap(name))[:n]
        err = float((t - taps[name]).abs().max() / taps[name].abs().max())
        assert err <= tap_tol, (name, err)
    d = float((got - ref).abs().max())
    assert d <= max(3 * floor, 0.05), (d, floor)
    pr, pg = torch.softmax(ref, 1), torch.softmax(got, 1)
    assert float((pr - pg).abs().max()) <= 1e-2
    top2 = torch.topk(pr, 2, 1).values
    sure = (top2[:, 0] - top2[:, 1]) > 2e-2
    assert bool((pr.argmax(1)[sure] == pg.argmax(1)[sure]).all())
    p, i = eng.read_topk(n)
    dev_probs = eng.read_probs(n)
    assert np.allclose(dev_probs, pg.numpy(), atol=1e-6, rtol=1e-4)
    tp, ti = torch.topk(torch.from_numpy(dev_probs), 5, 1)
    assert np.array_equal(i, ti.numpy().astype(np.int32)) and np.array_equal(p, tp.numpy())
    # the captured graph replays to the same bits as the eager pass
    again = eng.forward(x.numpy(), graph=True)
    eager = eng.forward(x.numpy(), graph=False)
    assert np.array_equal(again, eager)
    eng.close()
    return d, floor


def test_eva02_small_config_vs_oracle():
    from oracle import eva02
    _compare(eva02.Eva02Config(img_size=224, embed_dim=192, depth=4, num_heads=3, num_classes=1000), batch=5, n=5)


def test_eva02_padded_hidden_separate_qkv_partial_batch():
    """hidden = int(320 * 8 / 3) = 853 (padded to 896), separate q/k/v projections, 3 crops in an engine built for 4."""
    from oracle import eva02
    _compare(eva02.Eva02Config(img_size=112, embed_dim=320, depth=3, num_heads=5, num_classes=333), batch=4, n=3,
             seed=2, fused=False)


def test_eva02_large_shapes_one_block():
    """The shipped model's layer shapes (D 1024, 16 heads, hidden 2730, 577 tokens, 10 000 classes), one block deep."""
    from oracle import eva02
    _compare(eva02.Eva02Config(img_size=336, embed_dim=1024, depth=1, num_heads=16, num_classes=10000), batch=2, n=2,
             seed=3)


def test_eva02_many_crops_pair_gemm_and_gated_mlp():
    """40 crops at 336 (23 080 token rows): the linears take the CTA-pair GEMM and fc1 its gated epilogue."""
    from telescope_cam_detection_b200.eva02_engine import Eva02Engine
    from oracle import eva02
    cfg = eva02.Eva02Config(img_size=336, embed_dim=512, depth=1, num_heads=8, num_classes=100)
    sd = eva02.state_dict_numpy(eva02.make_model(cfg, 4))
    x = eva02.synthetic_input(cfg, 40, seed=5).numpy()
    eng = Eva02Engine(sd, batch=40, img_size=336, top_k=5)
    # LayerNorms applied inside the consumer GEMMs' epilogues, their statistics written by the producers' epilogues
    assert eng.fuse_glu and eng.ln_in_gemm and eng.ln_from_producers
    a = eng.forward(x)
    eng.close()
    eng = Eva02Engine(sd, batch=40, img_size=336, top_k=5, ln_from_producers=False)   # statistics by row_stats kernels
    assert eng.ln_in_gemm and not eng.ln_from_producers
    c = eng.forward(x)
    eng.close()
    assert float(np.abs(a - c).max()) <= 0.08, float(np.abs(a - c).max())
    eng = Eva02Engine(sd, batch=40, img_size=336, top_k=5, ln_in_gemm=False)   # normalised rows written by LayerNorm kernels
    assert not eng.ln_in_gemm
    b = eng.forward(x)
    eng.close()
    assert float(np.abs(a - b).max()) <= 0.08, float(np.abs(a - b).max())     # two bf16 pipelines, same network
    _compare(cfg, batch=40, n=40, seed=4)


def test_species_classifier_native_body():
    """SpeciesClassifier('timm/eva02_...') builds the native body; classify_crops (gather -> patch rows) agrees with
    classify on the sliced crop (preprocess -> NCHW -> patchify) and with the oracle on the reference's tensor."""
    import torch
    from telescope_cam_detection_b200.species_classifier import SpeciesClassifier
    from oracle import eva02
    from util import structured_frame
    cfg = eva02.Eva02Config(img_size=224, embed_dim=192, depth=2, num_heads=3, num_classes=500)
    # (default logit scale: at logit_std 4 the CPU bf16 rounding-point emulation is itself 1.2e-2 from fp32 on the
    # probabilities -- the 1e-2 tolerance below is meaningful for distributions no more peaked than this one)
    model = eva02.make_model(cfg, 5)
    clf = SpeciesClassifier(model_name="timm/eva02_tiny_patch14_224.mim_in22k", input_size=224, use_hierarchical=False,
                            confidence_threshold=0.0, weights=model.state_dict(), max_batch=3, batch_buckets=(1,))
    assert clf.load_model(num_classes=500) and clf.engine is not None
    frame = structured_frame(4, 480, 640)
    boxes = [(10, 20, 200, 300), (0, 0, 640, 480), (300, 100, 524, 324), (50, 60, 120, 200)]   # 4 crops > max_batch 3
    res = clf.classify_crops(frame, boxes, top_k=5)
    assert len(res) == 4
    for (x1, y1, x2, y2), r in zip(boxes, res):
        single = clf.classify(np.ascontiguousarray(frame[y1:y2, x1:x2]), top_k=5)
        assert single[0]["class_id"] == r[0]["class_id"] and len(single) == len(r)
        assert np.allclose([d["confidence"] for d in single], [d["confidence"] for d in r], atol=2e-3)
        with torch.no_grad():
            pr = torch.softmax(model(clf.preprocess(np.ascontiguousarray(frame[y1:y2, x1:x2])).cpu()), 1)[0]
        for d in r:
            assert abs(d["confidence"] - float(pr[d["class_id"]])) <= 1e-2
            assert d["species"] == "species_%d" % d["class_id"]
        if float(pr.max()) - float(torch.topk(pr, 2).values[1]) > 2e-2 and r:
            assert r[0]["class_id"] == int(pr.argmax())
    batch = clf.classify_batch([np.ascontiguousarray(frame[y1:y2, x1:x2]) for x1, y1, x2, y2 in boxes], top_k=3)
    assert [[d["class_id"] for d in b] for b in batch] == [[d["class_id"] for d in r[:3]] for r in res]
    assert sorted(clf._engines) == [1, 3]           # single crops ran on the one-slot program, built on first use
    assert clf.load_model(num_classes=500)          # idempotent
    clf.close()
    wrong = SpeciesClassifier(model_name="timm/eva02_tiny_patch14_224", input_size=224, weights=model.state_dict())
    assert not wrong.load_model(num_classes=1000)   # head size mismatch fails loudly, like load_state_dict would
