"""First-layer weight gradient without a pass over y / dY (csrc/conv_in_alg.cu): per-sample Gram matrices of the input patches
against an fp64 unfold, the whole gradient against fp32 autograd through conv -> BatchNorm -> ReLU -> MaxPool on the engine's own
saved tensors and against the dY-based kernels, cache behaviour, graph replays == the Python launch loop."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _rel(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-30))


def _patches(xt):
    """[D,H,W] (float64) -> [P, 125] patches of the 5^3 / stride-2 convolution, tap order (td, th, tw)."""
    u = xt.unfold(0, 5, 2).unfold(1, 5, 2).unfold(2, 5, 2)          # [Do,Ho,Wo,5,5,5]
    return u.reshape(-1, 125)


@pytest.mark.parametrize("shape", [(21, 25, 23), (13, 30, 121)])
def test_gram_matrices_match_an_fp64_unfold(shape):
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engine_b200 import _declare
    lib = native.load()
    _declare(lib)
    g = torch.Generator().manual_seed(1)
    n = 3
    x = torch.randint(0, 256, (n,) + shape, generator=g, dtype=torch.uint8).to(DEV)
    scale = 1 / 255.0
    fl = lib.nidt_conv_in_gram_floats()
    cache = torch.zeros((8, fl), dtype=torch.float32, device=DEV)
    scratch = torch.empty(lib.nidt_conv_in_gram_scratch_floats(n, shape[0]), dtype=torch.float32, device=DEV)
    rows, slots = (C.c_int * n)(2, 0, 1), (C.c_int * n)(5, 1, 7)
    st = torch.cuda.current_stream().cuda_stream
    native.check(lib.nidt_conv_in_gram_samples(x.data_ptr(), rows, slots, n, cache.data_ptr(), scratch.data_ptr(), *shape, scale, st), "gram")
    gb = torch.zeros(fl, dtype=torch.float64, device=DEV)
    native.check(lib.nidt_conv_in_gram_sum(cache.data_ptr(), (C.c_int * 2)(5, 7), 2, gb.data_ptr(), st), "sum")
    torch.cuda.synchronize()
    want = {}
    for row, slot in ((2, 5), (0, 1), (1, 7)):
        xt = (x[row].float() * scale).bfloat16().double()
        p = _patches(xt)
        p = torch.cat([p, torch.ones_like(p[:, :1]), torch.zeros_like(p[:, :2])], 1)     # tap 125 = 1, 126 / 127 = 0
        want[slot] = p.t() @ p
        got = cache[slot].view(128, 128).double()
        assert _rel(got, want[slot]) < 1e-6, (row, slot, _rel(got, want[slot]))
        assert float(got[125, 125]) == p.shape[0]
    assert _rel(gb.view(128, 128), want[5] + want[7]) < 1e-6
    assert float(cache[0].abs().max()) == 0.0                      # untouched slots stay untouched


def _engine(B=4, p_drop=0.0):
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    torch.manual_seed(5)
    eng = B200Engine(AlexNet3D_Dropout(1, p_drop=p_drop), DEV, max_batch=B)
    slab = eng.new_slab()
    # non-trivial BatchNorm parameters, one negative gamma (the pooling winner is then the window's MINIMUM of y)
    torch.manual_seed(6)
    slab.view("features.1.weight").uniform_(0.5, 1.5)
    slab.view("features.1.weight")[3] = -0.8
    slab.view("features.1.bias").uniform_(-0.3, 0.3)
    eng.attach(slab)
    return eng, slab


def _batch(B=4, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g)
    x[:, :, :, :30] = 0                                            # background, as in masked brain volumes
    y = (torch.arange(B) % 2).float()
    return x.to(DEV), y.to(DEV)


def test_gradient_matches_autograd_and_the_dy_based_kernels():
    torch.backends.cudnn.allow_tf32 = False
    B = 4
    eng, slab = _engine(B)
    assert eng.alg_bwd0 and eng.blocks[0].amax is not None
    x, y = _batch(B)
    ids = [100, 7, 42, 9]
    outs = {}
    for tag, use_ids, native_exec in (("alg_graph", True, True), ("alg_py", True, False), ("std", False, True), ("alg_again", True, True)):
        eng.native_exec = eng.use_graphs = native_exec
        eng.batch_ids = ids if use_ids else None
        eng.compute_grads(x, y, True)
        torch.cuda.synchronize()
        outs[tag] = eng.grad.flat.clone()
        assert eng._plan.blocks[0].bwd0_mode == (2 if use_ids else 0) or not native_exec
    assert eng.gram.built == B                                     # computed once, then served from the cache
    lay = eng.layout
    w = lay["features.0.weight"]
    sl = slice(w.offset, w.offset + w.numel)
    # the executor / graph replay and the Python launch loop issue the same kernels
    assert torch.equal(outs["alg_graph"], outs["alg_py"]) and torch.equal(outs["alg_graph"], outs["alg_again"])
    # everything but the first conv's own gradient does not depend on the variant at all
    other = torch.ones(lay.param_numel, dtype=torch.bool, device=DEV)
    other[sl] = False
    bo = lay["features.0.bias"]
    other[bo.offset:bo.offset + bo.numel] = False
    assert torch.equal(outs["alg_graph"][other], outs["std"][other])
    assert float(outs["alg_graph"][bo.offset:bo.offset + bo.numel].abs().max()) == 0.0       # exactly zero (BN removes the mean)
    # fp32 autograd of BN -> ReLU -> MaxPool on the engine's saved conv output (so the pooling winners / ReLU gates are decided on
    # the same values), then the conv's weight gradient from that dY; bf16 inputs / weights as the forward kernel multiplies them
    b0 = eng.blocks[0]
    yb = b0.y[:B].permute(0, 4, 1, 2, 3).float().requires_grad_(True)
    gam, bet = slab.view("features.1.weight").clone(), slab.view("features.1.bias").clone()
    o = F.max_pool3d(F.relu(F.batch_norm(yb, None, None, gam, bet, training=True, eps=1e-5)), 3, 3)
    o.backward(b0.da[:B].permute(0, 4, 1, 2, 3).float())
    xin = (x.float() / 255.0).bfloat16().float().unsqueeze(1)
    wq = slab.view("features.0.weight").clone().bfloat16().float().requires_grad_(True)
    F.conv3d(xin, wq, None, stride=2).backward(yb.grad)
    e_alg, e_std = _rel(outs["alg_graph"][sl], wq.grad), _rel(outs["std"][sl], wq.grad)
    assert e_alg < 3e-3, (e_alg, e_std)
    assert e_alg < e_std + 1e-3, (e_alg, e_std)                    # y un-rounded, dY never rounded to bf16: not worse than the dY path
    assert _rel(outs["alg_graph"][sl], outs["std"][sl]) < 6e-3


def test_training_with_cached_gram_matrices_and_partial_batches():
    from neuroimagedisttraining_b200.fl.engine import StepHParams
    B = 4
    eng, slab = _engine(B, p_drop=0.5)
    x, y = _batch(B, seed=3)
    hp = StepHParams(lr=0.01, momentum=0.9, wd=5e-4, clip_norm=10.0)
    mom = torch.zeros(eng.layout.param_numel, device=DEV)
    losses = []
    for step in range(6):
        b = 4 if step % 3 else 3                                   # partial batches take their own graphs
        eng.batch_ids = [11, 12, 13, 14][:b]
        eng.train_step(x[:b], y[:b], hp, None, mom, step == 0)
        assert eng._plan.blocks[0].bwd0_mode == 2 and eng.batch_ids is None
        losses.append(eng.pop_mean_loss())
    eng.train_step(x, y, hp, None, mom, False)                     # no ids: the dY-based kernels
    assert eng._plan.blocks[0].bwd0_mode == 0
    assert eng.gram.built == 4 and all(l == l for l in losses) and torch.isfinite(slab.flat).all()
    assert losses[-1] < losses[0]
    # a sample repeated inside one batch, and ids the cache has not seen
    eng.batch_ids = [11, 11, 99, 12]
    xx = torch.stack([x[0], x[0], x[3], x[1]])
    eng.train_step(xx, y, hp, None, mom, False)
    assert eng.gram.built == 5
    # ids are per volume store: the same ids under another store key are new samples
    eng.batch_ids, eng.batch_ids_key = [11, 12, 13, 14], "another store"
    eng.train_step(x, y, hp, None, mom, False)
    assert eng.gram.built == 9 and eng._plan.blocks[0].bwd0_mode == 2


def test_forward_with_the_conv_epilogue_pooling_along_w_is_bit_identical(monkeypatch):
    """NIDT_POOLW=1 (training steps too; the default "eval" uses it for evaluation batches only): when nobody reads the full-resolution conv output (evaluation; training with the backward from pooled tensors),
    the conv epilogue keeps the pooling winner of every 3 output columns and the pooling kernel finishes from those.  Same winners, same
    values: pooled activation, winner values / taps, every gradient, evaluation counters and logits must not change by a bit --
    including the channel with a negative BatchNorm weight (its winner is the window's MINIMUM of y)."""
    from neuroimagedisttraining_b200.fl.engine import EvalCounters
    B = 4
    monkeypatch.setenv("NIDT_POOLW", "1")
    eng_p, _ = _engine(B)
    monkeypatch.setenv("NIDT_POOLW", "0")
    eng_s, _ = _engine(B)
    assert eng_p.poolw and eng_p.poolw_train and not eng_s.poolw
    x, y = _batch(B, seed=7)
    ids = [3, 1, 4, 2]
    for native_exec in (True, False):                      # executor / graph replay, and the Python launch loop
        res = []
        for eng in (eng_p, eng_s):
            eng.native_exec = eng.use_graphs = native_exec
            for b in (B, B - 1):                           # a partial batch too
                eng.batch_ids = ids[:b]
                eng.compute_grads(x[:b], y[:b], True)
                torch.cuda.synchronize()
                b0 = eng.blocks[0]
                res.append((eng.grad.flat.clone(), b0.a[:b].clone(), b0.ymax[:b].clone(), b0.amax[:b].clone()))
            c = EvalCounters(DEV)
            eng.eval_batch(x, y, c)
            res.append((c.buf.clone(), eng.logits(x)))
        n = len(res) // 2
        for i in range(n):
            for j, (p, s) in enumerate(zip(res[i], res[n + i])):
                assert torch.equal(p, s), f"native_exec={native_exec} item {i}.{j}: {(p.double() - s.double()).abs().max().item()}"
