"""Kernels of the native 3-D ResNet path vs. plain PyTorch fp32 references of the same ops (``-m gpu``)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _cl(x):
    return x.permute(0, 2, 3, 4, 1).contiguous()


def _cf(x):
    return x.permute(0, 4, 1, 2, 3).contiguous()


def _close(a, b, atol=1e-2, what=""):
    a, b = a.float(), b.float()
    scale = b.abs().max().item() + 1e-12
    err = (a - b).abs().max().item()
    assert err <= atol * scale, f"{what}: max err {err:.4g} vs scale {scale:.4g}"


STRIDED = [
    # N, Cin, D,H,W, Cout, k, stride, pad
    (2, 64, 16, 19, 16, 128, 3, 2, 1),      # layer2.0.conv1 of ResNet3D at a reduced extent
    (2, 64, 16, 19, 16, 128, 1, 2, 0),      # layer2.0.downsample
    (3, 128, 8, 10, 8, 256, 3, 2, 1),       # layer3.0.conv1
    (3, 128, 9, 10, 7, 256, 1, 2, 0),
    (2, 256, 4, 5, 4, 256, 3, 1, 1),        # layer3 body
]


@pytest.mark.parametrize("N,Cin,D,H,W,Cout,k,stride,pad", STRIDED)
def test_strided_and_pointwise_conv_tc(N, Cin, D, H, W, Cout, k, stride, pad):
    """Element-strided TMA boxes (fprop / wgrad) and dgrad through zero-insertion upsampling."""
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(5)
    x = torch.randn(N, Cin, D, H, W, device=DEV).bfloat16()
    w = (torch.randn(Cout, Cin, k, k, k, device=DEV) / (Cin * k ** 3) ** 0.5).bfloat16().float()
    xf = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    ref = F.conv3d(xf, wr, None, stride=stride, padding=pad)
    wf, wd = V.pack_conv_weights(w)
    y, part = V.conv3d_fprop(_cl(x), wf, None, k, stride, pad, stats=True)
    torch.cuda.synchronize()
    assert tuple(y.shape[1:4]) == tuple(ref.shape[2:])
    _close(_cf(y), ref, what="fprop")
    yb = y.float().reshape(-1, Cout)
    torch.testing.assert_close(part[:, 0].sum(0), yb.sum(0), rtol=1e-3, atol=1e-2)
    dy = torch.randn_like(ref).bfloat16()
    ref.backward(dy.float())
    dx = V.conv3d_dgrad(_cl(dy), wd, k, pad, (D, H, W), stride=stride)
    _close(_cf(dx), xf.grad, what="dgrad")
    for splits in (1, 2):
        dw = V.conv3d_wgrad(_cl(dy), _cl(x), k, stride, pad, splits=splits)
        _close(dw, wr.grad, what=f"wgrad splits={splits}")


@pytest.mark.parametrize("shape,k,s,p", [((2, 64, 21, 25, 21), 3, 2, 1), ((1, 128, 9, 8, 7), 3, 2, 1), ((2, 64, 10, 9, 11), 2, 2, 0),
                                         ((2, 64, 9, 9, 9), 3, 3, 0)])
def test_maxpool3d_overlapping_fwd_bwd(shape, k, s, p):
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(6)
    x = torch.randn(*shape, device=DEV).bfloat16()
    xf = x.float().requires_grad_(True)
    ref = F.max_pool3d(xf, k, s, p)
    out, idx = V.maxpool3d_fwd(_cl(x), k, s, p)
    assert torch.equal(_cf(out).float(), ref)
    dout = torch.randn_like(ref).bfloat16()
    ref.backward(dout.float())
    dx = V.maxpool3d_bwd(_cl(dout), idx, shape[2:], k, s, p)
    _close(_cf(dx), xf.grad, atol=8e-3, what="maxpool bwd")      # sums of up to 8 bf16 values, rounded once


def test_elementwise_upsample_avgpool():
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(7)
    a = torch.randn(2, 5, 6, 7, 64, device=DEV).bfloat16()
    b = torch.randn_like(a)
    assert torch.equal(V.ew_bf16(a, b, 0).float(), torch.relu(a.float() + b.float()).bfloat16().float())
    assert torch.equal(V.ew_bf16(a, b, 1).float(), torch.where(b > 0, a, torch.zeros_like(a)).float())
    assert torch.equal(V.ew_bf16(a, b, 2).float(), (a.float() + b.float()).bfloat16().float())
    u = V.upsample_zero(a, 2)
    assert u.shape == (2, 9, 11, 13, 64)
    assert torch.equal(u[:, ::2, ::2, ::2], a) and float(u.float().abs().sum()) == pytest.approx(float(a.float().abs().sum()), rel=1e-6)
    x = torch.randn(3, 256, 8, 10, 8, device=DEV).bfloat16()
    xf = x.float().requires_grad_(True)
    ref = F.avg_pool3d(xf, 3).flatten(1)
    feat = V.avgpool3d_fwd(_cl(x), 3)
    torch.testing.assert_close(feat, ref, rtol=1e-5, atol=1e-5)
    g = torch.randn_like(ref)
    ref.backward(g)
    dx = V.avgpool3d_bwd(g, (3, 8, 10, 8, 256), 3)
    _close(_cf(dx), xf.grad, atol=5e-3, what="avgpool bwd")


@pytest.mark.parametrize("B,Fdim,Hd", [(8, 3072, 512), (5, 768, 512), (16, 3072, 512)])
def test_linear_head_matches_torch(B, Fdim, Hd):
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(8)
    feat = torch.randn(B, Fdim, device=DEV)
    y = (torch.rand(B, device=DEV) > 0.5).float()
    W1 = (torch.randn(Hd, Fdim, device=DEV) / Fdim ** 0.5).requires_grad_(True)
    b1 = (torch.randn(Hd, device=DEV) * 0.1).requires_grad_(True)
    W2 = (torch.randn(1, Hd, device=DEV) / Hd ** 0.5).requires_grad_(True)
    b2 = (torch.randn(1, device=DEV) * 0.1).requires_grad_(True)
    fr = feat.clone().requires_grad_(True)
    logits = F.linear(F.linear(fr, W1, b1), W2, b2).squeeze(1)
    loss = F.binary_cross_entropy_with_logits(logits, y)
    loss.backward()
    out = V.linear_head_fwd_bwd(feat, y, W1.detach(), b1.detach(), W2.detach(), b2.detach(), train=True)
    torch.testing.assert_close(out["logits"], logits.detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(out["loss"], loss.detach(), rtol=1e-4, atol=1e-5)
    for name, ref in (("gW1", W1.grad), ("gb1", b1.grad), ("gW2", W2.grad), ("gb2", b2.grad), ("dfeat", fr.grad)):
        torch.testing.assert_close(out[name], ref, rtol=1e-3, atol=1e-6, msg=name)
    counters = torch.zeros(3, dtype=torch.float64, device=DEV)
    ev = V.linear_head_fwd_bwd(feat, y, W1.detach(), b1.detach(), W2.detach(), b2.detach(), train=False, counters=counters)
    pr = torch.sigmoid(logits.detach())
    assert counters[0].item() == float(((pr >= 0.5).float() == y).sum())
    assert counters[2].item() == B
    ref_loss = F.binary_cross_entropy_with_logits(pr, y, reduction="sum")      # the reference's double sigmoid
    assert counters[1].item() == pytest.approx(ref_loss.item(), rel=1e-4)


HALO_CASES = [
    # N, Cin, D,H,W, Cout, pad, block_n
    (1, 64, 4, 5, 6, 64, 1, 64),
    (2, 64, 6, 14, 32, 64, 1, 64),          # ResNet layer1 row pitch (34), several row blocks
    (2, 64, 9, 13, 11, 128, 0, 128),        # AlexNet conv2 style (no padding, 128 output channels)
    (2, 64, 9, 13, 11, 128, 0, 64),         # two N tiles
    (2, 128, 7, 9, 9, 64, 2, 64),           # dgrad geometry: pad' = 2, two Cin blocks
    (3, 192, 4, 5, 4, 128, 1, 64),          # three Cin blocks
]


def test_conv3d_halo_reports_unsupported_shapes():
    from neuroimagedisttraining_b200.ops import volumetric as V
    lib = V._lib()
    assert lib.nidt_conv3d_halo_supported(3, 4, 5, 4, 256, 256, 1, 128) == 0      # four Cin blocks of planes do not fit in shared memory
    assert lib.nidt_conv3d_halo_supported(2, 8, 8, 200, 64, 64, 1, 64) == 0        # row pitch above 128
    assert lib.nidt_conv3d_halo_supported(2, 8, 8, 8, 32, 64, 1, 64) == 0          # Cin not a multiple of 64


@pytest.mark.parametrize("N,Cin,D,H,W,Cout,pad,block_n", HALO_CASES)
def test_conv3d_halo_resident_kernel(N, Cin, D, H, W, Cout, pad, block_n):
    """Halo-resident 3x3x3 conv (shifted SWIZZLE_128B descriptors into planes kept in shared memory)."""
    from neuroimagedisttraining_b200.ops import volumetric as V
    Do, Ho, Wo = D + 2 * pad - 2, H + 2 * pad - 2, W + 2 * pad - 2
    assert V._lib().nidt_conv3d_halo_supported(N, Do, Ho, Wo, Cin, Cout, pad, block_n) == 1
    torch.manual_seed(9)
    x = torch.randn(N, Cin, D, H, W, device=DEV).bfloat16()
    w = (torch.randn(Cout, Cin, 3, 3, 3, device=DEV) / (Cin * 27) ** 0.5).bfloat16().float()
    bias = torch.randn(Cout, device=DEV) * 0.1
    ref = F.conv3d(x.float(), w, bias, stride=1, padding=pad)
    wf, wd = V.pack_conv_weights(w)
    y, part = V.conv3d_halo_fprop(_cl(x), wf, bias, pad, block_n=block_n, stats=True)
    torch.cuda.synchronize()
    _close(_cf(y), ref, what="halo fprop")
    yb = y.float().reshape(-1, Cout)
    torch.testing.assert_close(part[:, 0].sum(0), yb.sum(0), rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(part[:, 1].sum(0), (yb * yb).sum(0), rtol=1e-3, atol=1e-2)
    # as dgrad: dX of a conv with padding p is the halo conv of dY with the flipped weights and pad' = 2 - p
    xf = x.float().requires_grad_(True)
    out = F.conv3d(xf, w, None, stride=1, padding=pad)
    dy = torch.randn_like(out).bfloat16()
    out.backward(dy.float())
    if Cout % 64 == 0 and V._lib().nidt_conv3d_halo_supported(N, D, H, W, Cout, Cin, 2 - pad, 64):
        dx = V.conv3d_halo_fprop(_cl(dy), wd, None, 2 - pad, block_n=64)
        _close(_cf(dx), xf.grad, what="halo dgrad")
