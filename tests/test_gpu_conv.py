"""Per-layer parity of the tcgen05/TMA implicit-GEMM conv (through the C ABI) against
torch fp32 conv2d on the same bf16-rounded operands.  Tolerance: bf16 output rounding
(rel 2^-8) + fp32 accumulation-order noise."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

CASES = [
    # N, H, W, cin, cout, k, stride, silu, resid, x_ctot, x_coff, y_ctot, y_coff
    (1, 16, 16, 64, 64, 1, 1, 0, 0, 64, 0, 64, 0),
    (2, 80, 80, 128, 128, 3, 1, 1, 0, 128, 0, 128, 0),
    (2, 80, 80, 128, 256, 3, 1, 1, 0, 128, 0, 256, 0),
    (2, 40, 40, 128, 128, 3, 1, 1, 1, 256, 0, 256, 0),
    (2, 20, 20, 256, 256, 3, 1, 1, 0, 512, 256, 512, 0),
    (2, 160, 160, 32, 32, 3, 1, 1, 1, 64, 0, 64, 0),
    (2, 160, 160, 32, 32, 1, 1, 1, 0, 64, 0, 32, 0),
    (1, 320, 320, 12, 32, 3, 1, 1, 0, 16, 0, 32, 0),
    (2, 320, 320, 32, 64, 3, 2, 1, 0, 32, 0, 64, 0),
    (2, 80, 80, 128, 256, 3, 2, 1, 0, 256, 128, 256, 0),
    (2, 40, 40, 256, 512, 3, 2, 1, 0, 512, 256, 512, 0),
    (3, 20, 20, 1024, 512, 1, 1, 1, 0, 1024, 0, 512, 0),
    (1, 24, 40, 64, 64, 3, 1, 1, 0, 64, 0, 128, 64),
    # CTA-pair kernel (conv_pair.cuh): several rounds per cluster, odd patch count (a no-op half pair), partial patches
    (8, 80, 80, 128, 256, 3, 1, 1, 0, 128, 0, 256, 0),
    (3, 40, 40, 128, 128, 3, 1, 1, 1, 256, 128, 256, 128),
    (1, 48, 40, 128, 128, 3, 1, 0, 0, 128, 0, 384, 256),
]

# N, H, W, cin per group, cout, groups: the fused cls_b | reg_b convs of the head (pair kernel at 80 / 40, generic at 20)
GROUPED = [(8, 80, 80, 128, 256, 2), (2, 40, 40, 128, 256, 2), (2, 20, 20, 128, 256, 2), (2, 40, 40, 64, 128, 2)]


@pytest.mark.parametrize("case", GROUPED)
def test_grouped_conv_matches_torch(case):
    import torch
    import torch.nn.functional as F
    from telescope_cam_detection_b200 import engine

    torch.backends.cudnn.allow_tf32 = False
    N, H, W, cin, cout, groups = case
    g = torch.Generator(device="cpu").manual_seed(77 + H + cin)
    dev = "cuda:0"
    x = torch.randn(N, H, W, groups * cin, generator=g).to(dev).to(torch.bfloat16)
    w = (torch.randn(cout, cin, 3, 3, generator=g) / (cin * 9) ** 0.5).float()
    b = torch.randn(cout, generator=g).float() * 0.1
    y = torch.zeros((N, H, W, cout), device=dev, dtype=torch.bfloat16)
    engine.conv2d_nhwc_bf16(x.data_ptr(), N, H, W, groups * cin, 0, cin, w.numpy(), b.numpy(), 3, 1, 1,
                            y.data_ptr(), cout, 0, groups=groups)
    torch.cuda.synchronize()
    ref = F.silu(F.conv2d(x.float().permute(0, 3, 1, 2), w.to(dev).to(torch.bfloat16).float(), b.to(dev), padding=1,
                          groups=groups))
    got = y.float().permute(0, 3, 1, 2)
    err = (got - ref).abs()
    assert bool((err <= 1e-2 + 8e-3 * ref.abs()).all()), "max err %.4g" % err.max().item()



@pytest.mark.parametrize("case", CASES)
def test_conv_matches_torch(case):
    import torch
    import torch.nn.functional as F
    from telescope_cam_detection_b200 import engine

    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    N, H, W, cin, cout, k, stride, silu, resid, x_ctot, x_coff, y_ctot, y_coff = case
    g = torch.Generator(device="cpu").manual_seed(1234 + H + cin + cout)
    dev = "cuda:0"
    cin_pad = (cin + 15) // 16 * 16
    x = torch.randn(N, H, W, x_ctot, generator=g).to(dev).to(torch.bfloat16)
    if cin_pad != cin:
        x[..., x_coff + cin:x_coff + cin_pad] = 0
    w = (torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5).float()
    b = torch.randn(cout, generator=g).float() * 0.1
    Ho, Wo = H // stride, W // stride
    y = torch.full((N, Ho, Wo, y_ctot), 7.0, device=dev, dtype=torch.bfloat16)
    r = torch.randn(N, Ho, Wo, y_ctot, generator=g).to(dev).to(torch.bfloat16) if resid else None
    if resid:
        y.copy_(r)  # in-place residual, like the CSP bottleneck
    engine.conv2d_nhwc_bf16(x.data_ptr(), N, H, W, x_ctot, x_coff, cin, w.numpy(), b.numpy(), k, stride, silu,
                            y.data_ptr(), y_ctot, y_coff,
                            resid_ptr=y.data_ptr() if resid else 0, r_ctot=y_ctot, r_coff=y_coff)
    torch.cuda.synchronize()
    xin = x[..., x_coff:x_coff + cin].float().permute(0, 3, 1, 2)
    wq = w.to(dev).to(torch.bfloat16).float()
    ref = F.conv2d(xin, wq, b.to(dev), stride=stride, padding=(k - 1) // 2)
    if silu:
        ref = F.silu(ref)
    if resid:
        ref = ref + r[..., y_coff:y_coff + cout].float().permute(0, 3, 1, 2)
    got = y[..., y_coff:y_coff + cout].float().permute(0, 3, 1, 2)
    err = (got - ref).abs()
    tol = 1e-2 + 8e-3 * ref.abs()
    assert bool((err <= tol).all()), "max err %.4g at %s (ref %.4g)" % (
        err.max().item(), np.unravel_index(int(err.argmax()), err.shape), ref.flatten()[err.argmax()].item())
    # untouched channels of the output buffer stay untouched
    if y_ctot != cout:
        other = torch.ones(y_ctot, dtype=torch.bool)
        other[y_coff:y_coff + cout] = False
        expect = r[..., other].float() if resid else torch.full_like(y[..., other].float(), 7.0)
        assert torch.equal(y[..., other].float(), expect)
