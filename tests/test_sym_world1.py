"""The fused symmetric-memory collectives (csrc/sym_agg.cu) at world size 1 on one GPU, where every "peer" pointer aliases the
local buffer: the reduce-scatter + mask + all-gather kernel (W1) and the score all-reduce + global top-k + mask kernel (W2) against
plain fp32 PyTorch.  The multi-rank forms are tested in test_multi_gpu.py (>= 2 GPUs) and, for the flag protocol, in
test_sym_model.py (host model)."""
import os
import socket

import pytest
import torch

from neuroimagedisttraining_b200 import ops
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.runtime.slab import SlabLayout

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def world1():
    import torch.distributed as dist
    from neuroimagedisttraining_b200.parallel.context import DistContext
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    created = False
    if not dist.is_initialized():
        s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
        created = True
    yield DistContext(0, 1, 0, dev, "nccl", dist.group.WORLD)
    if created:
        dist.destroy_process_group()


def _paths(agg):
    return [False, True] if agg.has_multicast else [False]


def test_reduce_mask_broadcast_world1_equals_masked_copy(world1):
    from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator
    dev = world1.device
    n, nm = 2571672, 2570244
    agg = SymmetricAggregator(world1, n, dev)
    g = torch.Generator().manual_seed(3)
    part = torch.randn(n, generator=g).to(dev)
    mask = (torch.rand(nm, generator=g) > 0.5).float().to(dev)
    ref = part.clone()
    ref[:nm] *= mask
    for mm in _paths(agg):
        agg.use_multimem = mm
        for rep in range(2):                       # two launches: the epoch / flag protocol must carry over
            agg.local_partial.copy_(part)
            out = torch.full((n,), float("nan"), device=dev)
            agg.reduce_mask_broadcast(out, mask)
            torch.cuda.synchronize()
            assert torch.equal(out, ref), f"multimem={mm} rep={rep}: max diff {(out - ref).abs().max().item()}"
        out = torch.empty(n, device=dev)
        agg.local_partial.copy_(part)
        agg.reduce_mask_broadcast(out, None)
        torch.cuda.synchronize()
        assert torch.equal(out, part)
    # every data path, with the weighting of K client slabs inside the kernel, written in place (adopted slab) and to another tensor
    agg.use_multimem = None
    locs = [torch.randn(n, generator=g).to(dev) for _ in range(3)]
    wts = [0.2, 0.3, 0.5]
    want = torch.empty(n, device=dev)
    ops.weighted_aggregate(want, locs, wts, None)
    want[:nm] *= mask
    cases = [("p2p", 0.0), ("bulk", 0.0)] + ([("multimem", 0.0), ("bulk", 0.5), ("p2p", 0.3)] if agg.has_multicast else [])
    for mode, frac in cases:
        for target in (torch.full((n,), float("nan"), device=dev), agg.local_result):
            agg.aggregate(target, list(zip(wts, locs)), mask, mode=mode, mm_frac=frac)
            torch.cuda.synchronize()
            assert torch.equal(target, want), f"mode={mode} mm_frac={frac}: max diff {(target - want).abs().max().item()}"


@pytest.mark.parametrize("replicate", ["0", "1"])
def test_saliency_topk_mask_world1_equals_torch_topk(world1, replicate, monkeypatch):
    """replicate = 1: the variant that all-gathers the reduced scores and selects over the whole slab on every rank (what runs from
    4 ranks on for L2-sized slabs); 0: sharded histograms exchanged per radix pass."""
    from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator
    monkeypatch.setenv("NIDT_SAL_REPLICATE", replicate)
    dev = world1.device
    layout = SlabLayout.from_module(AlexNet3D_Dropout(1))
    P = layout.param_numel
    agg = SymmetricAggregator(world1, layout.total_numel, dev)
    flags = ops.prunable_flags(layout, dev)
    g = torch.Generator().manual_seed(11)
    for keep in (0.5, 0.1):
        sc = (torch.rand(P, generator=g) ** 3).to(dev) * 1e-3
        # fp32 PyTorch statement of snip.py:80-116: normalise the prunable scores by their sum, k-th largest, >= threshold
        prunable = (flags[:P] == 2)
        k = int(layout.prunable_numel() * keep)
        for mm in _paths(agg):
            agg.use_multimem = mm
            mask = torch.empty(P, device=dev)
            s_in = sc.clone()
            _, info = agg.saliency_topk_mask(s_in, layout, keep, mask, flags)
            torch.cuda.synchronize()
            assert info["k"] == k
            assert abs(info["norm"] - float(sc[prunable].double().sum())) <= 1e-5 * info["norm"]
            # the kernel divides by ITS sum (a fixed-order block reduction); use that value so ties fall the same way
            q = sc[prunable] / torch.tensor(info["norm"], dtype=torch.float32, device=dev)
            thr = torch.topk(q, k, sorted=True)[0][-1]
            assert float(thr) == info["threshold"]
            ref = torch.ones(P, device=dev)
            ref[prunable] = (q >= thr).float()
            real = flags[:P] != 0
            assert torch.equal(mask[real], ref[real]), f"keep={keep} multimem={mm}: {(mask[real] != ref[real]).sum().item()} entries differ"
            assert info["kept"] == int(ref[prunable].sum().item()) >= k
            torch.testing.assert_close(s_in, sc, rtol=0, atol=0)      # one rank: the reduced scores are the input
