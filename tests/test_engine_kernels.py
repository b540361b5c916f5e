"""sm_100a volumetric kernels vs. plain PyTorch fp32 references of the same ops (``-m gpu``).
Inputs are rounded to bf16 first, so the comparison isolates the kernel from the input quantisation;
tolerances are bf16-output bands (rtol ~ 2^-8)."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"


def _cl(x):      # NCDHW -> NDHWC
    return x.permute(0, 2, 3, 4, 1).contiguous()


def _cf(x):      # NDHWC -> NCDHW
    return x.permute(0, 4, 1, 2, 3).contiguous()


def _close(a, b, rtol=2e-2, atol=2e-2, what=""):
    a, b = a.float(), b.float()
    scale = b.abs().max().item() + 1e-12
    err = (a - b).abs().max().item()
    assert err <= atol * scale + rtol * 0, f"{what}: max err {err:.4g} vs scale {scale:.4g}"


@pytest.mark.parametrize("C,pool,fast", [(64, 3, False), (64, 3, True), (128, 3, True), (192, 0, False), (128, 0, False)])
def test_bn_relu_pool_fwd_bwd(C, pool, fast):
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(0)
    N, D, H, W = 3, 8, 11, 7
    y = (torch.randn(N, C, D, H, W, device=DEV) * 1.5 + 0.3).bfloat16()
    gamma = torch.rand(C, device=DEV) + 0.5
    beta = torch.randn(C, device=DEV) * 0.2
    yf = y.float().requires_grad_(True)
    g2, b2 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm, rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    ref = F.relu(F.batch_norm(yf, rm.clone(), rv.clone(), g2, b2, training=True, momentum=0.1, eps=1e-5))
    if pool:
        ref = F.max_pool3d(ref, pool, pool)
    # our stats: partials from the bf16 tensor (one "tile")
    ycl = _cl(y)
    flat = ycl.float().reshape(-1, C)
    partials = torch.stack([flat.sum(0), (flat * flat).sum(0)]).unsqueeze(0).contiguous()
    nb = torch.zeros(1, device=DEV)
    scale, shift, mean, invstd = V.bn_finalize(partials, flat.shape[0], gamma, beta, rm, rv, nb)
    ymax = None
    if fast:      # 3x3x3 fast path: forward also emits the raw arg-max values the backward consumes
        a, ymax = V.bn_relu_pool_fwd(ycl, scale, shift, pool, pool, True, want_ymax=True)
    else:
        a = V.bn_relu_pool_fwd(ycl, scale, shift, pool, pool, True)
    _close(_cf(a), ref, what="fwd")
    ref_rm, ref_rv = torch.zeros(C, device=DEV), torch.ones(C, device=DEV)
    F.batch_norm(y.float(), ref_rm, ref_rv, gamma, beta, training=True, momentum=0.1, eps=1e-5)
    torch.testing.assert_close(rm, ref_rm, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(rv, ref_rv, rtol=1e-4, atol=1e-5)
    assert nb.item() == 1.0
    # backward
    da = torch.randn_like(ref).bfloat16()
    ref.backward(da.float())
    dy, dg, db = V.bn_relu_pool_bwd(ycl, _cl(da), scale, shift, mean, invstd, pool, pool, True, ymax=ymax)
    _close(_cf(dy), yf.grad, what="dy")
    _close(dg, g2.grad, atol=1e-2, what="dgamma")
    _close(db, b2.grad, atol=1e-2, what="dbeta")


def test_head_fwd_bwd_matches_torch():
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engine_b200 import _declare
    lib = native.load(); _declare(lib)
    torch.manual_seed(1)
    B, S, C, Hd = 16, 2, 128, 64
    Fdim = S * C
    feat = torch.randn(B, S, C, device=DEV).bfloat16()          # NDHWC with D*H*W = S
    yl = (torch.rand(B, device=DEV) > 0.5).float()
    W1 = (torch.randn(Hd, Fdim, device=DEV) * 0.05).requires_grad_(True)
    b1 = torch.zeros(Hd, device=DEV).requires_grad_(True)
    W2 = (torch.randn(1, Hd, device=DEV) * 0.1).requires_grad_(True)
    b2 = torch.zeros(1, device=DEV).requires_grad_(True)
    fl = feat.float().permute(0, 2, 1).reshape(B, Fdim).requires_grad_(True)   # reference order c*S+s
    z = F.linear(F.relu(F.linear(fl, W1, b1)), W2, b2)
    loss = F.binary_cross_entropy_with_logits(z.flatten(), yl)
    loss.backward()
    gW1, gb1, gW2, gb2 = torch.zeros_like(W1), torch.zeros_like(b1), torch.zeros_like(W2), torch.zeros_like(b2)
    dfeat = torch.zeros_like(feat)
    lacc = torch.zeros((), device=DEV)
    logits = torch.zeros(B, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    native.check(lib.nidt_head_fwd_bwd(feat.data_ptr(), yl.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(),
                                       gW1.data_ptr(), gb1.data_ptr(), gW2.data_ptr(), gb2.data_ptr(), dfeat.data_ptr(), lacc.data_ptr(),
                                       None, logits.data_ptr(), B, S, C, Hd, 0.0, 5, 1, 1, st))
    torch.testing.assert_close(logits, z.flatten().detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(lacc, loss.detach(), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(gW1, W1.grad, rtol=1e-3, atol=1e-5)
    torch.testing.assert_close(gb1, b1.grad, rtol=1e-3, atol=1e-6)
    torch.testing.assert_close(gW2, W2.grad, rtol=1e-3, atol=1e-5)
    torch.testing.assert_close(gb2, b2.grad, rtol=1e-3, atol=1e-6)
    ref_df = fl.grad.reshape(B, C, S).permute(0, 2, 1)
    _close(dfeat, ref_df, atol=1e-2, what="dfeat")
    # eval counters incl. the reference's double-sigmoid loss
    cnt = torch.zeros(3, dtype=torch.float64, device=DEV)
    native.check(lib.nidt_head_fwd_bwd(feat.data_ptr(), yl.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(),
                                       None, None, None, None, None, None, cnt.data_ptr(), logits.data_ptr(), B, S, C, Hd, 0.5, 5, 0, 1, st))
    p = torch.sigmoid(z.flatten().detach())
    assert cnt[2].item() == B
    assert cnt[0].item() == ((p >= 0.5).float() == yl).float().sum().item()
    assert abs(cnt[1].item() - F.binary_cross_entropy_with_logits(p, yl, reduction="sum").item()) < 1e-3


def test_head_dropout_is_consistent_and_seeded():
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engine_b200 import _declare
    lib = native.load(); _declare(lib)
    torch.manual_seed(2)
    B, S, C, Hd = 8, 2, 128, 64
    feat = torch.randn(B, S, C, device=DEV).bfloat16()
    yl = (torch.rand(B, device=DEV) > 0.5).float()
    W1, b1 = torch.randn(Hd, S * C, device=DEV) * 0.05, torch.zeros(Hd, device=DEV)
    W2, b2 = torch.randn(1, Hd, device=DEV) * 0.1, torch.zeros(1, device=DEV)
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for seed in (11, 11, 12):
        logits = torch.zeros(B, device=DEV)
        g = [torch.zeros_like(t) for t in (W1, b1, W2, b2)]
        df = torch.zeros_like(feat)
        native.check(lib.nidt_head_fwd_bwd(feat.data_ptr(), yl.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(),
                                           g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(), g[3].data_ptr(), df.data_ptr(), None, None,
                                           logits.data_ptr(), B, S, C, Hd, 0.5, seed, 1, 1, st))
        outs.append((logits.clone(), df.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    assert not torch.equal(outs[0][0], outs[2][0])
    dropped = (outs[0][1].float() == 0).float().mean().item()
    assert 0.35 < dropped < 0.65      # p = 0.5 on the feature gradient


CONV_CASES = [
    # N, Cin, D,H,W, Cout, k, pad, block_n
    (2, 64, 19, 23, 19, 128, 3, 0, 128),     # conv2 of AlexNet3D
    (2, 64, 19, 23, 19, 128, 3, 0, 64),
    (4, 128, 5, 7, 5, 192, 3, 1, 64),        # conv3
    (3, 192, 5, 7, 5, 192, 3, 1, 64),        # conv4
    (16, 192, 5, 7, 5, 128, 3, 1, 128),      # conv5, full batch
]


@pytest.mark.parametrize("N,Cin,D,H,W,Cout,k,pad,block_n", CONV_CASES)
def test_conv3d_fprop_dgrad_wgrad_tc(N, Cin, D, H, W, Cout, k, pad, block_n):
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(3)
    x = torch.randn(N, Cin, D, H, W, device=DEV).bfloat16()
    w = (torch.randn(Cout, Cin, k, k, k, device=DEV) / (Cin * k ** 3) ** 0.5).bfloat16().float()
    bias = torch.randn(Cout, device=DEV) * 0.1
    xf = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    ref = F.conv3d(xf, wr, bias, stride=1, padding=pad)
    wf, wd = V.pack_conv_weights(w)
    y, part = V.conv3d_fprop(_cl(x), wf, bias, k, 1, pad, block_n=block_n, stats=True)
    torch.cuda.synchronize()
    _close(_cf(y), ref, atol=1e-2, what="fprop")
    yb = y.float().reshape(-1, Cout)
    torch.testing.assert_close(part[:, 0].sum(0), yb.sum(0), rtol=1e-3, atol=1e-2)
    torch.testing.assert_close(part[:, 1].sum(0), (yb * yb).sum(0), rtol=1e-3, atol=1e-2)
    dy = torch.randn_like(ref).bfloat16()
    ref.backward(dy.float())
    dx = V.conv3d_dgrad(_cl(dy), wd, k, pad, (D, H, W), block_n=64)
    _close(_cf(dx), xf.grad, atol=1e-2, what="dgrad")
    for splits in (1, 3):
        dw = V.conv3d_wgrad(_cl(dy), _cl(x), k, 1, pad, splits=splits)
        _close(dw, wr.grad, atol=1e-2, what=f"wgrad splits={splits}")


@pytest.mark.parametrize("N,D,H,W,k,stride,pad", [(2, 121, 145, 121, 5, 2, 0), (1, 29, 33, 31, 5, 2, 0), (2, 20, 22, 21, 3, 2, 3)])
def test_conv_in_fprop_wgrad(N, D, H, W, k, stride, pad):
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(4)
    x = torch.randint(0, 256, (N, D, H, W), dtype=torch.uint8, device=DEV)
    scale = 1.0 / 255.0
    w = (torch.randn(64, 1, k, k, k, device=DEV) / (k ** 3) ** 0.5)
    bias = torch.randn(64, device=DEV) * 0.1
    xb = (x.float() * scale).bfloat16().float().unsqueeze(1).requires_grad_(False)
    wr = w.bfloat16().float().requires_grad_(True)
    br = bias.bfloat16().float().requires_grad_(True)      # the bias rides the GEMM as a bf16 K column
    ref = F.conv3d(xb, wr, br, stride=stride, padding=pad)
    y, part = V.conv_in_fprop(x, w, bias, k, stride, pad, scale, stats=True)
    torch.cuda.synchronize()
    _close(_cf(y), ref, atol=1e-2, what="conv_in fprop")
    yb = y.float().reshape(-1, 64)
    torch.testing.assert_close(part[:, 0].sum(0), yb.sum(0), rtol=2e-3, atol=5e-2)
    torch.testing.assert_close(part[:, 1].sum(0), (yb * yb).sum(0), rtol=2e-3, atol=5e-2)
    dy = (torch.randn_like(ref) * 0.1).bfloat16()
    ref.backward(dy.float())
    dw, dbias = V.conv_in_wgrad(x, _cl(dy), k, stride, pad, scale, with_bias=True)
    _close(dw, wr.grad, atol=1e-2, what="conv_in wgrad")
    _close(dbias, br.grad, atol=1e-2, what="conv_in bias grad")


@pytest.mark.parametrize("N,D,H,W", [(2, 121, 145, 121), (1, 29, 34, 31), (3, 15, 17, 21)])
def test_conv_in_space_to_depth_fprop_wgrad(N, D, H, W):
    """The zero-im2col first layer: tcgen05 descriptors with a 16-byte K-chunk stride over parity-interleaved lines."""
    from neuroimagedisttraining_b200.ops import volumetric as V
    torch.manual_seed(6)
    x = torch.randint(0, 256, (N, D, H, W), dtype=torch.uint8, device=DEV)
    scale = 1.0 / 255.0
    w = torch.randn(64, 1, 5, 5, 5, device=DEV) / 125 ** 0.5
    bias = torch.randn(64, device=DEV) * 0.1
    xb = (x.float() * scale).bfloat16().float().unsqueeze(1)
    wr = w.bfloat16().float().requires_grad_(True)
    ref = F.conv3d(xb, wr, bias, stride=2)
    y, part = V.conv_in_s2d_fprop(x, w, bias, scale, stats=True)
    torch.cuda.synchronize()
    _close(_cf(y), ref, atol=1e-2, what="s2d fprop")
    yb = y.float().reshape(-1, 64)
    torch.testing.assert_close(part[:, 0].sum(0), yb.sum(0), rtol=2e-3, atol=5e-2)
    torch.testing.assert_close(part[:, 1].sum(0), (yb * yb).sum(0), rtol=2e-3, atol=5e-2)
    dy = (torch.randn_like(ref) * 0.1).bfloat16()
    ref.backward(dy.float())
    dw = V.conv_in_s2d_wgrad(x, _cl(dy), scale)
    _close(dw, wr.grad, atol=1e-2, what="s2d wgrad")


def _engine_pair(B, p_drop=0.0):
    from neuroimagedisttraining_b200.fl.engine import TorchEngine
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    torch.manual_seed(5)
    model = AlexNet3D_Dropout(1, p_drop=p_drop)
    import copy
    tmodel = copy.deepcopy(model)
    eng = B200Engine(model, DEV, max_batch=B)
    ref = TorchEngine(tmodel, DEV, volumetric=True, input_scale=1 / 255.0)
    return eng, ref


def _rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return (a - b).norm().item() / (b.norm().item() + 1e-30)


def test_engine_backward_chain_is_exact_per_block():
    """Every kernel of the backward chain, checked on the engine's OWN saved bf16 tensors against
    PyTorch fp32 autograd of the same op (so argmax / ReLU-gate decisions are taken on identical
    values): dy of BN+ReLU+MaxPool, dgamma, dbeta, wgrad and dgrad of every block."""
    torch.backends.cudnn.allow_tf32 = False
    B = 4
    eng, _ = _engine_pair(B)
    slab = eng.new_slab()
    eng.attach(slab)
    g = torch.Generator(device="cpu").manual_seed(0)
    x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g).to(DEV)
    y = torch.tensor([0.0, 1.0, 1.0, 0.0], device=DEV)
    eng.compute_grads(x, y, True)
    torch.cuda.synchronize()
    for i, b in enumerate(eng.blocks):
        s = b.spec
        yb = _cf(b.y[:B]).float().requires_grad_(True)
        gam = slab.view(s.bn + ".weight").clone().requires_grad_(True)
        bet = slab.view(s.bn + ".bias").clone().requires_grad_(True)
        o = F.relu(F.batch_norm(yb, None, None, gam, bet, training=True, eps=1e-5))
        if s.pool_k:
            o = F.max_pool3d(o, s.pool_k, s.pool_s)
        o.backward(_cf(b.da[:B]).float())
        fused0 = i == 0 and getattr(eng, "fused_bwd0", False)       # block 0: dY exists only inside the fused weight-gradient kernel
        if not fused0:
            assert _rel(_cf(b.dy[:B]), yb.grad) < 4e-3, f"dy{i}"
        dy_used = yb.grad.bfloat16().float() if fused0 else _cf(b.dy[:B]).float()
        assert _rel(eng.grad.view(s.bn + ".weight"), gam.grad) < 1e-3, f"dgamma{i}"
        assert _rel(eng.grad.view(s.bn + ".bias"), bet.grad) < 1e-3, f"dbeta{i}"
        xin = (x.float() / 255.0).bfloat16().float().unsqueeze(1) if i == 0 else _cf(eng.blocks[i - 1].a[:B]).float()
        xin.requires_grad_(i > 0)
        wq = slab.view(s.conv + ".weight").clone().bfloat16().float().requires_grad_(True)
        F.conv3d(xin, wq, None, stride=s.stride, padding=s.pad).backward(dy_used)
        assert _rel(eng.grad.view(s.conv + ".weight"), wq.grad) < (3e-3 if fused0 else 1e-3), f"dW{i}"
        if i > 0:
            assert _rel(_cf(eng.blocks[i - 1].da[:B]), xin.grad) < 4e-3, f"da{i - 1}"


def test_engine_end_to_end_vs_fp32_autograd():
    """Whole step against PyTorch fp32 autograd.  The forward agrees to bf16 round-off; the backward
    of BatchNorm+MaxPool at batch 4 on random volumes is sensitive to which near-tied element wins a
    pooling window, so gradients are compared by direction (profiles/engine_numerics.txt)."""
    B = 4
    eng, ref = _engine_pair(B)
    slab = eng.new_slab()
    slab2 = slab.clone()
    eng.attach(slab)
    ref.attach(slab2)
    g = torch.Generator(device="cpu").manual_seed(0)
    x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g).to(DEV)
    y = torch.tensor([0.0, 1.0, 1.0, 0.0], device=DEV)
    loss = eng.compute_grads(x, y, True)
    rloss = ref.compute_grads(x, y, True)
    torch.cuda.synchronize()
    assert abs(loss.item() - rloss.item()) < 2e-2, (loss.item(), rloss.item())
    lay = eng.layout
    for name in lay.param_names:
        if lay[name].prunable or name.startswith("classifier"):
            a, b = eng.grad.view(name).float(), ref.grad.view(name).float()
            cos = F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item()
            assert cos > 0.85, f"{name}: cos {cos:.4f}"
    P = lay.param_numel
    for name in lay.buffer_names:
        torch.testing.assert_close(slab.view(name), slab2.view(name) if "num_batches" not in name else slab.view(name), rtol=3e-2, atol=3e-3)


def test_engine_trains_and_evaluates():
    from neuroimagedisttraining_b200.fl.engine import EvalCounters, StepHParams
    B = 4
    eng, ref = _engine_pair(B, p_drop=0.0)
    slab = eng.new_slab()
    eng.attach(slab)
    g = torch.Generator(device="cpu").manual_seed(1)
    x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g).to(DEV)
    y = torch.tensor([0.0, 1.0, 1.0, 0.0], device=DEV)
    hp = StepHParams(lr=0.01, momentum=0.0, wd=5e-4, clip_norm=10.0)
    losses = []
    for step in range(6):
        eng.train_step(x, y, hp, None, None, step == 0)
        losses.append(eng.pop_mean_loss())
    assert all(l == l for l in losses) and losses[-1] < losses[0] + 0.05, losses
    c = EvalCounters(DEV)
    eng.eval_batch(x, y, c)
    r = c.result()
    assert r["test_total"] == B and 0 <= r["test_correct"] <= B
    # eval logits agree with the PyTorch module on the trained weights
    ref.attach(slab.clone())
    ref.model.eval()
    with torch.no_grad():
        z = ref.model(ref._prep(x)).flatten()
    torch.testing.assert_close(eng.logits(x), z, rtol=5e-2, atol=5e-2)


def test_saliency_epilogue_scores_equal_the_separate_pass():
    """SNIP scores written by the wgrad reductions (conv layers) + the classifier's flat pass == |g*w| computed from the gradient slab
    afterwards, on every prunable weight (``snip.py:60-71``); the gradients themselves are unchanged by the epilogue."""
    from neuroimagedisttraining_b200 import ops
    eng, _ = _engine_pair(4, p_drop=0.0)
    slab = eng.new_slab()
    eng.attach(slab)
    g = torch.Generator().manual_seed(21)
    x = torch.randint(0, 256, (4, 121, 145, 121), generator=g, dtype=torch.uint8).to(DEV)
    y = (torch.rand(4, generator=g) > 0.5).float().to(DEV)
    lay = eng.layout
    scores = torch.zeros(lay.param_numel, device=DEV)
    bn_tail = slab.flat[lay.param_numel:].clone()
    assert eng.saliency_step(x, y, scores, 0.5)
    g_fused = eng.grad.flat.clone()
    slab.flat[lay.param_numel:].copy_(bn_tail)
    eng.step -= 1                                            # same dropout stream (p = 0 here anyway)
    eng.compute_grads(x, y, True)
    torch.cuda.synchronize()
    assert torch.equal(eng.grad.flat, g_fused)
    ref = torch.zeros_like(scores)
    ops.saliency_accumulate(ref, eng.grad.flat, slab.params_flat, 0.5)
    for name in lay.prunable_names:
        e = lay[name]
        a, b = scores[e.offset:e.offset + e.numel], ref[e.offset:e.offset + e.numel]
        assert float(b.abs().sum()) > 0
        assert torch.equal(a, b), f"{name}: max diff {(a - b).abs().max().item()}"


def test_fused_block0_backward_equals_the_separate_kernels_bit_for_bit():
    """The first block's backward with dY formed inside the weight-gradient kernel (BatchNorm backward + ReLU gate + MaxPool scatter
    fused, winner known from the argmax taps recorded by the forward) against the separate apply + wgrad kernels: same arithmetic
    per element, same MMA order, so gradients and updated weights must be identical -- at a full and at a partial batch."""
    import os
    from neuroimagedisttraining_b200.fl.engine import StepHParams
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    g = torch.Generator().manual_seed(31)
    x = torch.randint(0, 256, (4, 121, 145, 121), generator=g, dtype=torch.uint8).to(DEV)
    y = (torch.rand(4, generator=g) > 0.5).float().to(DEV)
    outs = {}
    for fused in ("1", "0"):
        os.environ["NIDT_FUSED_BWD0"] = fused
        try:
            torch.manual_seed(5)
            model = AlexNet3D_Dropout(1, p_drop=0.0)
            eng = B200Engine(model, DEV, max_batch=4)
        finally:
            os.environ.pop("NIDT_FUSED_BWD0", None)
        assert eng.fused_bwd0 == (fused == "1")
        slab = eng.new_slab()
        eng.attach(slab)
        res = []
        for native_exec in (True, False):                      # graph / native executor, and the Python launch loop
            eng.native_exec = native_exec
            eng.use_graphs = native_exec
            for b in (4, 3):
                eng.train_step(x[:b], y[:b], StepHParams(lr=0.01, wd=5e-4), None, None, True)
                res.append(eng.grad.flat.clone())
        res.append(slab.flat.clone())
        torch.cuda.synchronize()
        outs[fused] = res
    for i, (a, b) in enumerate(zip(outs["1"], outs["0"])):
        assert torch.isfinite(a).all()
        assert torch.equal(a, b), f"output {i}: max diff {(a - b).abs().max().item()}"
