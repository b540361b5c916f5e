"""Flat-slab ops: CPU restatements vs. hand formulas (always) and sm_100a kernels vs. the fp32
PyTorch restatements (``-m gpu``)."""

import numpy as np
import pytest
import torch

from neuroimagedisttraining_b200 import ops
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout, Tiny3DCNN
from neuroimagedisttraining_b200.ops import reference as R
from neuroimagedisttraining_b200.runtime.slab import SlabLayout


def _rand(n, seed=0, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(n, generator=g).to(device)


def test_sgd_reference_matches_torch_optimizer():
    torch.manual_seed(0)
    lin = torch.nn.Linear(8, 4)
    w0 = torch.cat([p.detach().flatten() for p in lin.parameters()]).clone()
    opt = torch.optim.SGD(lin.parameters(), lr=0.1, momentum=0.9, weight_decay=5e-4)
    w = w0.clone()
    mom = torch.zeros_like(w)
    for step in range(3):
        x = torch.randn(5, 8)
        lin.zero_grad()
        (lin(x) ** 2).sum().backward()
        g = torch.cat([p.grad.flatten() for p in lin.parameters()]).clone()
        torch.nn.utils.clip_grad_norm_(lin.parameters(), 10)
        opt.step()
        R.masked_sgd_step(w, g, mom, None, 0.1, 0.9, 5e-4, 10.0, first_step=(step == 0))
        torch.testing.assert_close(w, torch.cat([p.detach().flatten() for p in lin.parameters()]), rtol=1e-5, atol=1e-6)


def test_aggregate_reference_order_is_bit_exact():
    xs = [_rand(1000, s) for s in range(5)]
    ns = [3, 7, 11, 2, 9]
    ws = [n / sum(ns) for n in ns]
    out = torch.empty(1000)
    R.weighted_aggregate(out, xs, ws)
    ref = xs[0] * ws[0]
    for x, w in zip(xs[1:], ws[1:]):
        ref += x * w
    assert torch.equal(out, ref)


def test_topk_reference_semantics():
    m = Tiny3DCNN(1)
    layout = SlabLayout.from_module(m)
    scores = torch.rand(layout.param_numel)
    mask = torch.empty(layout.param_numel)
    _, info = ops.global_topk_mask(scores, layout, 0.3, mask)
    kept = sum(float(mask[layout[n].offset:layout[n].offset + layout[n].numel].sum()) for n in layout.prunable_names)
    assert kept == info["k"] == int(layout.prunable_numel() * 0.3)
    for n in layout.param_names:
        if not layout[n].prunable:
            assert bool((mask[layout[n].offset:layout[n].offset + layout[n].numel] == 1).all())


# ------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("momentum,use_mask,first", [(0.0, True, True), (0.9, True, False), (0.9, False, True), (0.0, False, False)])
def test_masked_sgd_kernel(momentum, use_mask, first):
    n = 2570244
    dev = "cuda"
    w, g, mom = _rand(n, 1, dev), _rand(n, 2, dev) * 0.01, _rand(n, 3, dev)
    mask = (torch.rand(n, device=dev) > 0.5).float() if use_mask else None
    w_ref, mom_ref = w.clone().cpu(), mom.clone().cpu()
    tn = R.masked_sgd_step(w_ref, g.cpu(), mom_ref, mask.cpu() if use_mask else None, 0.01, momentum, 5e-4, 10.0, first)
    norm = torch.zeros((), device=dev)
    ops.masked_sgd_step(w, g, mom, mask, 0.01, momentum, 5e-4, 10.0, first, norm_out=norm)
    torch.testing.assert_close(w.cpu(), w_ref, rtol=1e-5, atol=1e-6)
    torch.testing.assert_close(norm.cpu(), tn, rtol=1e-4, atol=1e-5)
    if momentum:
        torch.testing.assert_close(mom.cpu(), mom_ref, rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
def test_masked_sgd_kernel_clips():
    n = 4096
    w, g = _rand(n, 1, "cuda"), _rand(n, 2, "cuda") * 5.0
    w_ref = w.clone().cpu()
    R.masked_sgd_step(w_ref, g.cpu(), None, None, 0.1, 0.0, 0.0, 10.0, True)
    ops.masked_sgd_step(w, g, None, None, 0.1, 0.0, 0.0, 10.0, True)
    torch.testing.assert_close(w.cpu(), w_ref, rtol=1e-5, atol=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 3, 21, 70])
def test_weighted_aggregate_kernel_bit_exact(k):
    n, nm = 2571672, 2570244
    xs = [_rand(n, s, "cuda") for s in range(k)]
    ns = [5 + 3 * i for i in range(k)]
    ws = [x / sum(ns) for x in ns]
    mask = (torch.rand(nm, device="cuda") > 0.5).float()
    out = torch.empty(n, device="cuda")
    ops.weighted_aggregate(out, xs, ws, mask)
    ref = torch.empty(n)
    R.weighted_aggregate(ref, [x.cpu() for x in xs], ws, mask.cpu())
    assert torch.equal(out.cpu(), ref), float((out.cpu() - ref).abs().max())


@pytest.mark.gpu
def test_saliency_accumulate_kernel():
    n = 2570244
    s, g, w = torch.rand(n, device="cuda"), _rand(n, 1, "cuda"), _rand(n, 2, "cuda")
    ref = R.saliency_accumulate(s.clone().cpu(), g.cpu(), w.cpu(), 0.25)
    ops.saliency_accumulate(s, g, w, 0.25)
    torch.testing.assert_close(s.cpu(), ref, rtol=1e-6, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("keep", [0.5, 0.1, 0.9])
def test_global_topk_mask_kernel(keep):
    layout = SlabLayout.from_module(AlexNet3D_Dropout(1))
    torch.manual_seed(0)
    scores = (torch.rand(layout.param_numel) ** 4) * 1e-3
    scores[::7] = scores[3]          # plenty of exact ties
    ref_mask = torch.empty(layout.param_numel)
    _, info_ref = ops.global_topk_mask(scores.clone(), layout, keep, ref_mask)
    mask = torch.empty(layout.param_numel, device="cuda")
    _, info = ops.global_topk_mask(scores.cuda(), layout, keep, mask)
    diff = int((mask.cpu() != ref_mask).sum())
    # the only tolerated disagreement is at the threshold value itself (normalising-sum round-off)
    assert diff <= 8, (diff, info, info_ref)
    assert abs(info["threshold"] - info_ref["threshold"]) <= 1e-6 * abs(info_ref["threshold"]) + 1e-12
    assert info["kept"] >= info["k"]


@pytest.mark.gpu
def test_count_hamming_sqdist_kernels():
    n = 1000003
    a = (torch.rand(n, device="cuda") > 0.3).float() * _rand(n, 1, "cuda")
    b = (torch.rand(n, device="cuda") > 0.6).float()
    assert int(ops.count_nonzero(a)) == int(torch.count_nonzero(a))
    assert int(ops.hamming_distance(a, b)) == int(R.hamming_distance(a.cpu(), b.cpu()))
    torch.testing.assert_close(ops.squared_distance(a, b).cpu(), R.squared_distance(a.cpu(), b.cpu()), rtol=1e-4, atol=1e-3)
    w, ref = _rand(4096, 5, "cuda"), _rand(4096, 6, "cuda")
    w2 = w.clone().cpu()
    R.prox_sgd_pull(w2, ref.cpu(), 0.1, 0.5)
    ops.prox_pull(w, ref, 0.1, 0.5)
    torch.testing.assert_close(w.cpu(), w2, rtol=1e-6, atol=1e-7)


@pytest.mark.gpu
@pytest.mark.parametrize("row_shape", [(121, 145, 121), (5, 7, 3), (4, 4, 4), (33,)])
def test_gather_rows_u8_matches_index_select(row_shape):
    """Odd row lengths: neither source nor destination rows are 16-byte aligned."""
    from neuroimagedisttraining_b200 import ops
    g = torch.Generator().manual_seed(3)
    n = 9
    src = torch.randint(0, 256, (n,) + row_shape, dtype=torch.uint8, generator=g).cuda()
    for idx in ([0], [8, 8, 1], [3, 0, 7, 7, 2, 5, 1]):
        ix = torch.tensor(idx, dtype=torch.int64, device="cuda")
        out = ops.gather_rows_u8(src, ix)
        assert torch.equal(out, src.index_select(0, ix))
    # destination that does not start on a 16-byte boundary
    buf = torch.zeros(4 * src[0].numel() + 64, dtype=torch.uint8, device="cuda")
    view = buf[5:5 + 3 * src[0].numel()].view((3,) + row_shape)
    ix = torch.tensor([6, 2, 4], dtype=torch.int64, device="cuda")
    ops.gather_rows_u8(src, ix, out=view)
    assert torch.equal(view, src.index_select(0, ix)) and int(buf[:5].sum()) == 0 and int(buf[5 + 3 * src[0].numel():].sum()) == 0


@pytest.mark.gpu
def test_batch_stager_delivers_the_stored_volumes_with_lookahead():
    """Pinned 8-bit store -> device ring through the native row-copy call; the lookahead stream is resumed by the next iterate()."""
    from neuroimagedisttraining_b200.data import SyntheticABCD
    from neuroimagedisttraining_b200.data.loaders import IndexLoader
    from neuroimagedisttraining_b200.runtime.staging import BatchStager
    store = SyntheticABCD(24, (9, 11, 7), n_sites=2, seed=3, pin=True)
    st = BatchStager(store, torch.device("cuda"), 4)
    a = IndexLoader(np.arange(0, 10), store.labels[:10], store.sites[:10], 4, shuffle=False)
    b = IndexLoader(np.arange(10, 24), store.labels[10:24], store.sites[10:24], 4, shuffle=False)
    seen = []
    for x, y in st.iterate(a, lookahead=b):
        seen.append((x.clone(), y.clone()))
    assert st._carry is not None
    for x, y in st.iterate(b):
        seen.append((x.clone(), y.clone()))
    ids = np.concatenate([np.arange(0, 10), np.arange(10, 24)])
    got = torch.cat([x for x, _ in seen]).cpu()
    want, labels = store.fetch(ids)
    assert torch.equal(got, torch.as_tensor(want).cpu())
    assert torch.allclose(torch.cat([y for _, y in seen]).cpu(), torch.as_tensor(labels).float().cpu())
    assert [x.shape[0] for x, _ in seen] == [4, 4, 2, 4, 4, 4, 2]


def test_mask_select_k_cpu_semantics():
    torch.manual_seed(0)
    w = torch.randn(1000)
    m = (torch.rand(1000) > 0.4).float()
    n_act = int(m.sum())
    m2 = ops.mask_select_k(w, m.clone(), 37, "fire")
    assert int(m2.sum()) == n_act - 37
    removed = (m == 1) & (m2 == 0)
    assert w.abs()[removed].max() <= w.abs()[(m2 == 1)].min()
    g = torch.randn(1000)
    m3 = ops.mask_select_k(g, m2.clone(), 37, "regrow")
    assert int(m3.sum()) == n_act
    grown = (m2 == 0) & (m3 == 1)
    assert g.abs()[grown].min() >= g.abs()[(m3 == 0)].max()


@pytest.mark.gpu
@pytest.mark.parametrize("n,k", [(1000, 37), (1_769_472, 300_000), (257, 256), (4096, 0), (50_000, 49_999)])
def test_mask_select_k_matches_sort_reference(n, k):
    """Radix-select prune / regrow against the sort-based restatement (no ties: identical masks; with ties: identical counts)."""
    g = torch.Generator().manual_seed(n)
    w = torch.randn(n, generator=g)
    m = (torch.rand(n, generator=g) > 0.4).float()
    k_f = min(k, int(m.sum()))
    ref = ops.mask_select_k(w, m.clone(), k_f, "fire")
    got = ops.mask_select_k(w.cuda(), m.clone().cuda(), k_f, "fire")
    assert torch.equal(got.cpu(), ref)
    k_r = min(k, int((ref == 0).sum()))
    gr = torch.randn(n, generator=g)
    ref2 = ops.mask_select_k(gr, ref.clone(), k_r, "regrow")
    got2 = ops.mask_select_k(gr.cuda(), got.clone(), k_r, "regrow")
    assert torch.equal(got2.cpu(), ref2)
    # heavy ties: quantised values; index-ordered tie-breaking = stable sort
    wq = torch.round(w * 2) / 2
    ref3 = ops.mask_select_k(wq, m.clone(), k_f, "fire")
    got3 = ops.mask_select_k(wq.cuda(), m.clone().cuda(), k_f, "fire")
    assert int(got3.sum()) == int(ref3.sum()) == int(m.sum()) - k_f
    assert torch.equal(got3.cpu(), ref3)
    # k larger than the number of candidates: every candidate changes
    got4 = ops.mask_select_k(w.cuda(), m.clone().cuda(), n, "fire")
    assert int(got4.sum()) == 0
