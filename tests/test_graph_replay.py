"""CUDA-graph replays of the native step (csrc/executor.cu: nidt_plan_train_step_graph / nidt_plan_eval_graph) against the plain
launch sequence: same kernels, same order, so weights, gradients, losses and evaluation counters must be bit-identical -- over
several steps with changing learning rate, dropout seed, slab and batch size."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda:0")


def _run(use_graphs: bool):
    from neuroimagedisttraining_b200.fl.engine import EvalCounters, StepHParams
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    torch.manual_seed(3)
    model = AlexNet3D_Dropout(1, p_drop=0.5)
    eng = B200Engine(model, DEV, max_batch=4)
    eng.use_graphs = use_graphs
    g = torch.Generator().manual_seed(0)
    xs = [torch.randint(0, 256, (4, 121, 145, 121), generator=g, dtype=torch.uint8).to(DEV) for _ in range(2)]
    ys = [(torch.rand(4, generator=g) > 0.5).float().to(DEV) for _ in range(2)]
    xbuf, ybuf = torch.empty_like(xs[0]), torch.empty_like(ys[0])      # stable batch buffers, as the stagers provide
    slabs = [eng.new_slab(), eng.new_slab()]
    mask = (torch.rand(eng.layout.param_numel, generator=g) > 0.3).float().to(DEV)
    mom = torch.zeros(eng.layout.param_numel, device=DEV)
    out = []
    for rnd in range(2):
        for si, slab in enumerate(slabs):
            eng.attach(slab)
            for step in range(3):
                b = 4 if step < 2 else 3                                 # a partial batch takes its own graph
                xbuf.copy_(xs[(step + si) % 2]); ybuf.copy_(ys[(step + si) % 2])
                hp = StepHParams(lr=0.01 * (0.9 ** rnd), momentum=0.9 if si else 0.0, wd=5e-4, clip_norm=10.0)
                eng.train_step(xbuf[:b], ybuf[:b], hp, mask, mom, step == 0)
            out.append(eng.pop_mean_loss(sync=False).sum.clone())
            out.append(eng.grad.flat.clone())
        acc = torch.zeros((2, 3), dtype=torch.float64, device=DEV)
        for si, slab in enumerate(slabs):
            eng.attach(slab)
            c = EvalCounters(DEV, acc[si])
            for k in range(2):
                xbuf.copy_(xs[k]); ybuf.copy_(ys[k])
                eng.eval_batch(xbuf, ybuf, c, double_sigmoid=True)
        out.append(acc.clone())
    out.extend(s.flat.clone() for s in slabs)
    torch.cuda.synchronize()
    n_graphs = eng.lib.nidt_plan_graph_count()
    del eng
    return out, n_graphs


def test_graph_replays_equal_plain_launches_bit_for_bit():
    ref, n0 = _run(False)
    got, n1 = _run(True)
    assert n0 == 0
    # 2 slabs x {batch 4, batch 3} x {momentum buffer or not is tied to the slab} training graphs + 2 eval graphs
    assert n1 == 2 * 2 + 2, n1
    assert len(ref) == len(got)
    for i, (a, b) in enumerate(zip(ref, got)):
        assert torch.equal(a, b), f"output {i}: max diff {(a.double() - b.double()).abs().max().item()}"
