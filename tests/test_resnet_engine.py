"""Native 3-D ResNet engine vs. PyTorch fp32 autograd of the same module (``-m gpu``)."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

DEV = "cuda"
SHAPE = (45, 49, 45)        # stem 25x27x25 -> pool 13x14x13 -> layer2 7x7x7 -> layer3 4x4x4 -> avgpool 1


def _cf(x):
    return x.permute(0, 4, 1, 2, 3).contiguous()


def _rel(a, b):
    a, b = a.float().flatten(), b.float().flatten()
    return (a - b).norm().item() / (b.norm().item() + 1e-30)


KINDS = ["basic", "bottleneck"]


def _pair(B, shape=SHAPE, kind="basic"):
    from neuroimagedisttraining_b200.fl.engine import TorchEngine
    from neuroimagedisttraining_b200.models.volumetric import resnet3d_18, resnet3d_bottleneck
    from neuroimagedisttraining_b200.ops.engines import make_native_engine
    torch.manual_seed(11)
    model = resnet3d_18(1, shape) if kind == "basic" else resnet3d_bottleneck(1, shape, layers=(2, 1, 1))
    # non-trivial BN affine parameters so the test sees them
    for m in model.modules():
        if isinstance(m, torch.nn.BatchNorm3d):
            torch.nn.init.uniform_(m.weight, 0.5, 1.5)
            torch.nn.init.uniform_(m.bias, -0.2, 0.2)
    tmodel = copy.deepcopy(model)
    eng = make_native_engine(model, DEV, max_batch=B)
    ref = TorchEngine(tmodel, DEV, volumetric=True, input_scale=1 / 255.0)
    return eng, ref


def _batch(B, shape=SHAPE, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    x = torch.randint(0, 256, (B,) + shape, dtype=torch.uint8, generator=g).to(DEV)
    y = (torch.arange(B, device=DEV) % 2).float()
    return x, y


@pytest.mark.parametrize("kind", KINDS)
def test_resnet_backward_chain_is_exact_per_op(kind):
    """Each backward kernel on the engine's own saved bf16 tensors against fp32 autograd of the same op."""
    torch.backends.cudnn.allow_tf32 = False
    B = 4
    eng, _ = _pair(B, kind=kind)
    slab = eng.new_slab()
    eng.attach(slab)
    x, y = _batch(B)
    eng.compute_grads(x, y, True)
    torch.cuda.synchronize()

    def check_conv_bn(c, xin, da, relu, want_dx, tag):
        yb = _cf(c.y[:B]).float().requires_grad_(True)
        gam = slab.view(c.bn + ".weight").clone().requires_grad_(True)
        bet = slab.view(c.bn + ".bias").clone().requires_grad_(True)
        o = F.batch_norm(yb, None, None, gam, bet, training=True, eps=1e-5)
        if relu:
            o = F.relu(o)
        o.backward(_cf(da[:B]).float())
        assert _rel(_cf(c.dy[:B]), yb.grad) < 5e-3, f"dy {tag}"
        assert _rel(eng.grad.view(c.bn + ".weight"), gam.grad) < 2e-3, f"dgamma {tag}"
        assert _rel(eng.grad.view(c.bn + ".bias"), bet.grad) < 2e-3, f"dbeta {tag}"
        xi = xin.clone().requires_grad_(want_dx is not None)
        wq = slab.view(c.conv + ".weight").clone().bfloat16().float().requires_grad_(True)
        F.conv3d(xi, wq, None, stride=c.stride, padding=c.pad).backward(_cf(c.dy[:B]).float())
        assert _rel(eng.grad.view(c.conv + ".weight"), wq.grad) < 2e-3, f"dW {tag}"
        return xi.grad

    for i in range(len(eng.blocks) - 1, -1, -1):
        b = eng.blocks[i]
        xin = _cf(b.x_in[:B]).float()
        # gate
        ref_dz = torch.where(b.out[:B] > 0, b.dout[:B], torch.zeros_like(b.dout[:B]))
        assert torch.equal(ref_dz, b.dz[:B]), f"gate {b.name}"
        n_main = len(b.main)
        assert n_main == (2 if kind == "basic" else 3)
        for j in range(n_main - 1, 0, -1):
            g_a = check_conv_bn(b.main[j], _cf(b.acts[j - 1][:B]).float(), b.dz if j == n_main - 1 else b.dacts[j], j < n_main - 1, True,
                                f"{b.name}.c{j + 1}")
            assert _rel(_cf(b.dacts[j - 1][:B]), g_a) < 5e-3, f"da{j} {b.name}"
        g_x = check_conv_bn(b.main[0], xin, b.dacts[0], True, True, b.name + ".c1")
        if b.ds is not None:
            g_x = g_x + check_conv_bn(b.ds, xin, b.dz, False, True, b.name + ".ds")
        else:
            g_x = g_x + _cf(b.dz[:B]).float()
        dx = eng.blocks[i - 1].dout if i > 0 else eng.stem.dp
        assert _rel(_cf(dx[:B]), g_x) < 6e-3, f"dx {b.name}"
    # stem: max-pool backward, BN+ReLU backward, first-layer wgrad
    s = eng.stem
    r = _cf(s.r[:B]).float().requires_grad_(True)
    F.max_pool3d(r, 3, 2, 1).backward(_cf(s.dp[:B]).float())
    assert _rel(_cf(s.dr[:B]), r.grad) < 5e-3, "maxpool bwd"
    xin = (x.float() / 255.0).bfloat16().float().unsqueeze(1)
    check_conv_bn(s, xin, s.dr, True, None, "stem")
    # head
    feat = F.avg_pool3d(_cf(eng.blocks[-1].out[:B]).float(), 3).flatten(1)
    torch.testing.assert_close(eng.feat[:B], feat, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("kind", KINDS)
def test_resnet_end_to_end_vs_fp32_autograd(kind):
    B = 4
    eng, ref = _pair(B, kind=kind)
    slab = eng.new_slab()
    slab2 = slab.clone()
    eng.attach(slab)
    ref.attach(slab2)
    x, y = _batch(B)
    loss = eng.compute_grads(x, y, True)
    rloss = ref.compute_grads(x, y, True)
    torch.cuda.synchronize()
    assert abs(loss.item() - rloss.item()) < 3e-2, (loss.item(), rloss.item())
    lay = eng.layout
    worst = 1.0
    for name in lay.param_names:
        if name.endswith(".weight") and lay[name].prunable:
            a, b = eng.grad.view(name).float(), ref.grad.view(name).float()
            cos = F.cosine_similarity(a.flatten(), b.flatten(), dim=0).item()
            worst = min(worst, cos)
            assert cos > 0.8, f"{name}: cos {cos:.4f}"
    for name in lay.buffer_names:
        if "num_batches" not in name:
            torch.testing.assert_close(slab.view(name), slab2.view(name), rtol=3e-2, atol=3e-3)
    # eval-mode logits agree with the module
    ref.model.eval()
    with torch.no_grad():
        z = ref.model(ref._prep(x)).flatten()
    torch.testing.assert_close(eng.logits(x), z, rtol=5e-2, atol=5e-2)


@pytest.mark.parametrize("kind", KINDS)
def test_resnet_trains_evaluates_and_handles_partial_batches(kind):
    from neuroimagedisttraining_b200.fl.engine import EvalCounters, StepHParams
    B = 4
    eng, _ = _pair(B, kind=kind)
    slab = eng.new_slab()
    eng.attach(slab)
    x, y = _batch(B, seed=1)
    hp = StepHParams(lr=0.01, momentum=0.0, wd=5e-4, clip_norm=10.0)
    losses = []
    for step in range(6):
        eng.train_step(x, y, hp, None, None, step == 0)
        losses.append(eng.pop_mean_loss())
    assert all(l == l for l in losses) and losses[-1] < losses[0], losses
    eng.train_step(x[:3], y[:3], hp, None, None, False)          # ragged last batch of a client
    l3 = eng.pop_mean_loss()
    assert l3 == l3 and l3 < 10.0
    c = EvalCounters(DEV)
    eng.eval_batch(x, y, c)
    eng.eval_batch(x[:2], y[:2], c)
    r = c.result()
    assert r["test_total"] == B + 2 and 0 <= r["test_correct"] <= B + 2
    assert torch.isfinite(slab.flat).all()


@pytest.mark.parametrize("kind", KINDS)
def test_resnet_abcd_shape_step_runs(kind):
    """One step at the ABCD shape (the FC is sized from the input: 3072 features, 4x that with bottleneck blocks)."""
    from neuroimagedisttraining_b200.fl.engine import StepHParams
    B = 2
    eng, _ = _pair(B, (121, 145, 121), kind=kind)
    assert eng.F == (3072 if kind == "basic" else 4 * 3072)
    slab = eng.new_slab()
    eng.attach(slab)
    x, y = _batch(B, (121, 145, 121))
    eng.train_step(x, y, StepHParams(lr=0.01), None, None, True)
    l = eng.pop_mean_loss()
    assert l == l and 0.0 < l < 5.0


def test_salientgrads_round_with_native_resnet_engine():
    """BASELINE config 4 in miniature: SalientGrads, 3-D ResNet, 90 % sparsity, rescale partition, native engine."""
    import logging
    from helpers import make_args
    from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd_rescale
    from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
    from neuroimagedisttraining_b200.models import create_model
    from neuroimagedisttraining_b200.ops import native
    from neuroimagedisttraining_b200.ops.engines import make_native_engine
    n_clients = 6
    store = SyntheticABCD(n_clients * 10, SHAPE, n_sites=3, seed=1, pin=True)
    args = make_args(client_num_in_total=n_clients, batch_size=4, epochs=1, comm_round=2, lr=0.01, dense_ratio=0.1, model="3DResNet")
    log = logging.getLogger("t"); log.setLevel(logging.ERROR)
    ds = load_partition_data_abcd_rescale(store, client_number=n_clients, batch_size=4)
    torch.manual_seed(0)
    model = create_model("3DResNet", 1, SHAPE)
    eng = make_native_engine(model, DEV, max_batch=4, input_scale=store.scale)
    tr = SlabTrainer(model, args, log, device=torch.device(DEV), store=store, engine=eng)
    api = SalientGradsAPI(ds, torch.device(DEV), args, tr, log)
    n0 = native.LAUNCH_COUNT
    st = api.train()
    assert native.LAUNCH_COUNT - n0 > 500                     # the native kernels ran, not a PyTorch fallback
    lay = api.layout
    kept = sum(float(api.mask.view(n).sum()) for n in lay.prunable_names)
    assert kept == int(lay.prunable_numel() * 0.1)
    for n in lay.prunable_names:
        assert float((api.w_global.view(n)[api.mask.view(n) == 0]).abs().max()) == 0.0
    assert len(st["global_test_acc"]) >= 2 and all(0.0 <= a <= 1.0 for a in st["global_test_acc"])
    assert torch.isfinite(api.w_global.flat).all()
