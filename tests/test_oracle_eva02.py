"""CPU checks of the Stage-2 classifier oracle (oracle/eva02.py, parity unpinned: timm absent) and of the host side of
the native body (eva02_engine.py tables / column splits, SpeciesClassifier.load_model's failure behaviour)."""
import numpy as np
import pytest
import torch

from oracle import eva02


def test_rope_table_is_a_rotation_by_position_times_band():
    """Independent statement: channel pair j of a patch at (row y, col x) turns by angle pos * 1/10000^(j' / 16), pos =
    y * 16/G for j < 16 and x * 16/G for j >= 16 (j' = j mod 16) -- checked through complex multiplication."""
    G, hd = 24, 64
    sin, cos = eva02.rope_table(G, 16, hd)
    x = torch.randn(G * G, hd)
    got = x * cos + eva02._rot(x) * sin
    z = torch.view_as_complex(x.reshape(G * G, hd // 2, 2).contiguous())
    j = torch.arange(hd // 2)
    band = 1.0 / (10000.0 ** ((j % 16).float() / 16))
    ys, xs = torch.meshgrid(torch.arange(G), torch.arange(G), indexing="ij")
    pos = torch.where(j[None] < 16, ys.reshape(-1, 1).float(), xs.reshape(-1, 1).float()) * (16.0 / G)
    ref = torch.view_as_real(z * torch.polar(torch.ones_like(pos), pos * band[None])).reshape(G * G, hd)
    assert torch.allclose(got, ref, atol=1e-4)


def test_engine_rope_table_equals_oracle():
    from telescope_cam_detection_b200.eva02_engine import rope_cos_sin
    for G in (8, 16, 24):
        sin, cos = eva02.rope_table(G, 16, 64)
        t = rope_cos_sin(G, 16)
        assert t.shape == (G * G, 32, 2)
        assert np.abs(t[..., 0] - cos.numpy()[:, ::2]).max() < 1e-6 and np.abs(t[..., 1] - sin.numpy()[:, ::2]).max() < 1e-6
        assert np.array_equal(cos.numpy()[:, ::2], cos.numpy()[:, 1::2])


def test_shipped_model_configuration():
    cfg = eva02.config_for("timm/eva02_large_patch14_clip_336.merged2b_ft_inat21")
    assert (cfg.embed_dim, cfg.depth, cfg.num_heads, cfg.img_size, cfg.hidden, cfg.grid) == (1024, 24, 16, 336, 2730, 24)
    with torch.device("meta"):
        n = sum(p.numel() for p in eva02.Eva02(cfg).parameters())
    assert 300e6 < n < 320e6      # ViT-L/14 trunk (~304 M) + 10 000-class head
    with pytest.raises(ValueError):
        eva02.config_for("tf_efficientnet_b4_ns")


def test_state_dict_uses_timm_key_names():
    cfg = eva02.Eva02Config(img_size=112, embed_dim=64, depth=1, num_heads=1, num_classes=10)
    keys = set(eva02.make_model(cfg).state_dict())
    for k in ("patch_embed.proj.weight", "cls_token", "pos_embed", "blocks.0.norm1.weight", "blocks.0.attn.qkv.weight",
              "blocks.0.attn.q_bias", "blocks.0.attn.v_bias", "blocks.0.attn.norm.weight", "blocks.0.attn.proj.bias",
              "blocks.0.mlp.fc1_g.weight", "blocks.0.mlp.fc1_x.bias", "blocks.0.mlp.norm.weight", "blocks.0.mlp.fc2.weight",
              "norm.weight", "head.bias"):
        assert k in keys, k
    assert not any("rope" in k for k in keys)      # the rotary tables are non-persistent buffers


def test_fused_and_separate_qkv_agree():
    cfg = eva02.Eva02Config(img_size=112, embed_dim=128, depth=2, num_heads=2, num_classes=50)
    a = eva02.make_model(cfg, 1)
    cfg2 = eva02.Eva02Config(img_size=112, embed_dim=128, depth=2, num_heads=2, num_classes=50, qkv_fused=False)
    b = eva02.Eva02(cfg2).eval()
    sd = dict(a.state_dict())
    for i in range(cfg.depth):
        p = "blocks.%d.attn." % i
        w = sd.pop(p + "qkv.weight")
        sd[p + "q_proj.weight"], sd[p + "k_proj.weight"], sd[p + "v_proj.weight"] = w[:128], w[128:256], w[256:]
        sd[p + "q_proj.bias"], sd[p + "v_proj.bias"] = sd.pop(p + "q_bias"), sd.pop(p + "v_bias")
    b.load_state_dict(sd)
    x = eva02.synthetic_input(cfg, 2)
    with torch.no_grad():
        assert torch.allclose(a(x), b(x), atol=1e-5)


def test_lowp_emulation_stays_near_fp32():
    cfg = eva02.Eva02Config(img_size=112, embed_dim=128, depth=3, num_heads=2, num_classes=100)
    x = eva02.synthetic_input(cfg, 3)
    with torch.no_grad():
        a, b = eva02.make_model(cfg, 0)(x), eva02.make_model(cfg, 0, lowp=True)(x)
    assert float((a.softmax(1) - b.softmax(1)).abs().max()) < 1e-2


def test_column_splits_respect_the_engine_limits():
    from telescope_cam_detection_b200.eva02_engine import _chunks
    for cout, step in ((5632, 256), (3072, 256), (1024, 256), (576, 32), (1792, 128), (4096, 256)):
        parts = _chunks(cout, step)
        assert parts[0][0] == 0 and sum(w for _, w in parts) == cout
        assert all(w <= 2048 and w % step == 0 for _, w in parts)
        assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(len(parts) - 1))


def test_species_classifier_fails_loudly_without_a_body():
    from telescope_cam_detection_b200.species_classifier import SpeciesClassifier
    # EVA02 name, no checkpoint: the reference would download from the hub; here load_model reports failure
    assert SpeciesClassifier(model_name="timm/eva02_large_patch14_clip_336.merged2b_ft_inat21", input_size=336).load_model(10000) is False
    # any other timm name without timm installed and without a caller-supplied module: failure, not a silent stand-in
    assert SpeciesClassifier(model_name="tf_efficientnet_b4_ns").load_model(1000, max_retries=1) is False
    if not torch.cuda.is_available():
        cfg = eva02.Eva02Config(img_size=112, embed_dim=64, depth=1, num_heads=1, num_classes=10)
        clf = SpeciesClassifier(model_name="eva02_tiny_patch14_224", input_size=112, weights=eva02.make_model(cfg).state_dict())
        assert clf.load_model(10) is False and clf.engine is None     # no sm_100 device: no CPU fallback


def test_checkpoint_readers(tmp_path):
    """The native body takes its weights from a torch pickle (the reference's checkpoint_path form) or a safetensors
    file (what timm's hub models ship); both come back as the same state dict."""
    from telescope_cam_detection_b200.species_classifier import SpeciesClassifier
    cfg = eva02.Eva02Config(img_size=112, embed_dim=64, depth=1, num_heads=1, num_classes=10)
    sd = eva02.make_model(cfg).state_dict()
    torch.save({"model_state_dict": sd}, tmp_path / "a.pth")
    torch.save(sd, tmp_path / "b.pth")
    from safetensors.numpy import save_file
    save_file({k: v.numpy() for k, v in sd.items()}, str(tmp_path / "c.safetensors"))
    for name in ("a.pth", "b.pth", "c.safetensors"):
        got = SpeciesClassifier._read_checkpoint(str(tmp_path / name))
        assert set(got) == set(sd)
        assert np.array_equal(np.asarray(got["head.weight"]), sd["head.weight"].numpy())


def test_layernorm_folded_into_the_consumer_gemm_is_the_same_function():
    """The engine never materialises LayerNorm(x): it folds (gamma, beta) into the consumer's (W, b) (`_fold`) and the
    GEMM epilogue applies  y = rstd * (x W'^T - mean * colsum(W')) + b'  from per-row (sum, sum of squares) partials.
    In fp32 / float64 that must equal Linear(LayerNorm(x)) (timm's order, oracle/eva02.py blocks), partials in any split."""
    from telescope_cam_detection_b200 import eva02_engine as E

    rng = np.random.default_rng(5)
    rows, d, cout, eps = 37, 256, 96, 1e-6
    x = (rng.standard_normal((rows, d)) * 1.7 + 0.4).astype(np.float32)
    g = rng.uniform(0.5, 1.5, d).astype(np.float32)
    be = (rng.standard_normal(d) * 0.2).astype(np.float32)
    w = (rng.standard_normal((cout, d)) / np.sqrt(d)).astype(np.float32)
    b = (rng.standard_normal(cout) * 0.1).astype(np.float32)

    x64 = x.astype(np.float64)
    mu, var = x64.mean(1, keepdims=True), x64.var(1, keepdims=True)
    want = ((x64 - mu) / np.sqrt(var + eps) * g + be) @ w.astype(np.float64).T + b

    class Holder:                       # `_fold` reads the LayerNorm parameters through `_g`
        fold_ln = True

        def _g(self, name):
            return {"n.weight": g, "n.bias": be}[name]

    w2, b2 = E.Eva02Engine._fold(Holder(), w, b, "n")
    # statistics the way the producers write them: (sum, sum of squares) partials over column slices, summed in order
    parts = 4
    sl = np.array_split(np.arange(d), parts)
    su = sum(x64[:, s].sum(1) for s in sl)
    sq = sum((x64[:, s] ** 2).sum(1) for s in sl)
    mean = su / d
    rstd = 1.0 / np.sqrt(np.maximum(sq / d - mean * mean, 0.0) + eps)
    colsum = w2.astype(np.float64).sum(1)
    got = rstd[:, None] * (x64 @ w2.astype(np.float64).T - mean[:, None] * colsum[None]) + b2
    assert np.abs(got - want).max() < 2e-5
    # the column sums the engine uploads are those of the bf16-ROUNDED weights (what the tensor core multiplies by)
    cs = E._colsum_bf16(w2)
    w2_bf16 = (E.to_bf16_bits(w2).astype(np.uint32) << 16).view(np.float32)
    assert np.allclose(cs, w2_bf16.astype(np.float64).sum(1), rtol=0, atol=1e-5)
    assert np.abs(cs - colsum).max() < 0.05          # near the fp32 sums, not equal: rounding of 256 weights per row


def test_program_buckets_and_replay_chunking_with_a_stub_engine(monkeypatch):
    """Host logic of the native Stage 2 without a GPU: the smallest program that holds n crops is chosen (built once),
    more than max_batch crops run as several replays whose rows come back in crop order, close() closes every program."""
    from telescope_cam_detection_b200 import eva02_engine
    from telescope_cam_detection_b200.species_classifier import SpeciesClassifier

    built, closed = [], []

    class Stub:
        def __init__(self, sd, batch, img_size, top_k, device):
            self.batch, self.top_k = batch, top_k
            built.append(batch)

        def close(self):
            closed.append(self.batch)

        def patchify_device(self, ptr, n):
            assert n <= self.batch
            self.n, self.tag = n, ptr

        def run(self):
            pass

        def sync(self):
            pass

        def read_topk(self, n):
            p = np.tile(np.linspace(1.0, 0.1, self.top_k, dtype=np.float32), (n, 1))
            i = np.tile(np.arange(self.top_k, dtype=np.int32), (n, 1)) + 1000 * self.tag
            return p, i

    monkeypatch.setattr(eva02_engine, "Eva02Engine", Stub)
    sc = SpeciesClassifier.__new__(SpeciesClassifier)
    sc.batch_buckets, sc.max_batch, sc.input_size, sc._dev_idx, sc._sd = (4, 16), 64, 336, 0, {}
    sc.engine = Stub({}, 64, 336, 32, 0)
    sc._engines = {64: sc.engine}
    sizes = [sc._engine_for(n).batch for n in (1, 4, 5, 16, 17, 64)]
    assert sizes == [4, 4, 16, 16, 64, 64]
    assert built == [64, 4, 16]                       # every size built once
    assert sc._engine_for(3) is sc._engine_for(2)

    class Part:                                       # what `_topk` needs of a CUDA tensor
        def __init__(self, n, s0=0):
            self.shape, self.s0, self.device = (n, 3, 336, 336), s0, "cpu"

        def contiguous(self):
            return self

        def float(self):
            return self

        def __getitem__(self, sl):
            return Part(min(sl.stop, self.shape[0]) - sl.start, sl.start)

        def data_ptr(self):
            return self.s0

    import torch
    monkeypatch.setattr(torch.cuda, "current_stream", lambda dev=None: type("S", (), {"synchronize": lambda self: None})())
    p, i = sc._topk(Part(64 + 64 + 5), top_k=5)      # 64 + 64 + 5 crops -> programs 64, 64, 16
    assert p.shape == (133, 5) and i.shape == (133, 5)
    assert (i[:64, 0] == 0).all() and (i[64:128, 0] == 64000).all() and (i[128:, 0] == 128000).all()
    sc.close()
    assert sorted(closed) == [4, 16, 64] and sc.engine is None and sc._engines == {}
