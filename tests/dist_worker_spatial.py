"""torchrun worker (gloo, CPU): a feature extractor run on depth slabs with halo exchange must reproduce the un-split module --
outputs, every parameter gradient, the input gradient and the BatchNorm running statistics."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn as nn

from neuroimagedisttraining_b200.models import Tiny3DCNN
from neuroimagedisttraining_b200.parallel.spatial import (SlabParallelFeatures, SpatialParallelModel, allreduce_gradients, halo_bytes,
                                                          scatter_depth)


def resnet_like():
    return nn.Sequential(nn.Conv3d(1, 6, 3, 2, 3), nn.BatchNorm3d(6), nn.ReLU(inplace=True), nn.MaxPool3d(3, 2, 1),
                         nn.Conv3d(6, 8, 3, 1, 1, bias=False), nn.BatchNorm3d(8), nn.ReLU(inplace=True),
                         nn.Conv3d(8, 8, 1, 2, 0), nn.AvgPool3d(2))


def check(features, x, tag, world, rank):
    torch.manual_seed(1)
    head_w = torch.randn(features(x[:1]).flatten(1).shape[1], dtype=x.dtype)
    ref = copy.deepcopy(features)
    xr = x.clone().requires_grad_(True)
    yr = (ref(xr).flatten(1) @ head_w).tanh().sum()
    yr.backward()

    xs, part = scatter_depth(x, world, rank)
    xs.requires_grad_(True)
    sp = SlabParallelFeatures(features)
    out = sp(xs, part)
    y = (out.flatten(1) @ head_w).tanh().sum()
    y.backward()
    allreduce_gradients(list(features.parameters()))
    a, b = part.own(rank)
    errs = {"out": float((y - yr).detach().abs()), "dx": float((xs.grad - xr.grad[:, :, a:b]).abs().max()) if b > a else 0.0}
    for (n, p), (_, q) in zip(features.named_parameters(), ref.named_parameters()):
        errs["d" + n] = float((p.grad - q.grad).abs().max())
    for (n, p), (_, q) in zip(features.named_buffers(), ref.named_buffers()):
        errs[n] = float((p.double() - q.double()).abs().max())
    worst = max(errs.values())
    t = torch.tensor([worst], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{tag}: world={world} worst abs err {float(t):.3e}", flush=True)
    return float(t) < 1e-9, errs


def main():
    torch.set_num_threads(2)
    dist.init_process_group("gloo")
    world, rank = dist.get_world_size(), dist.get_rank()
    torch.manual_seed(0)
    ok = True
    x = torch.randn(2, 1, 32, 30, 29, dtype=torch.float64)
    tiny = Tiny3DCNN(1, in_shape=(32, 30, 29)).double().features          # depth 32 -> 15 -> 5 -> 5 -> 1: some ranks end up empty
    good, errs = check(tiny, x, "tiny3dcnn", world, rank)
    ok = ok and good
    if not good:
        print(rank, errs, flush=True)
    x2 = torch.randn(2, 1, 29, 17, 16, dtype=torch.float64)
    good, errs = check(resnet_like().double(), x2, "resnet-like", world, rank)
    ok = ok and good
    if not good:
        print(rank, errs, flush=True)
    # three SGD steps of the whole model in spatial-parallel mode == three single-process steps (weights and BN statistics)
    torch.manual_seed(3)
    ref_model = Tiny3DCNN(1, in_shape=(32, 30, 29), p_drop=0.0).double()
    sp_model = copy.deepcopy(ref_model)
    spm = SpatialParallelModel(sp_model)
    opt_r = torch.optim.SGD(ref_model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    opt_s = torch.optim.SGD(sp_model.parameters(), lr=0.05, momentum=0.9, weight_decay=5e-4)
    crit = nn.BCEWithLogitsLoss()
    for step in range(3):
        xb = torch.randn(4, 1, 32, 30, 29, dtype=torch.float64)
        yb = (torch.rand(4, 1) > 0.5).double()
        opt_r.zero_grad(); crit(ref_model(xb), yb).backward(); opt_r.step()
        opt_s.zero_grad(); crit(spm(xb), yb).backward(); spm.reduce_gradients(); opt_s.step()
    worst = max(float((a.double() - b.double()).abs().max()) for a, b in zip(sp_model.state_dict().values(), ref_model.state_dict().values()))
    t = torch.tensor([worst], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"spatial-parallel training: world={world} worst abs err after 3 steps {float(t):.3e}", flush=True)
    ok = ok and float(t) < 1e-9
    # the 3-D ResNet (residual blocks, strided 1x1x1 projections, padded max-pool, average pool) in spatial-parallel mode
    from neuroimagedisttraining_b200.models.volumetric import resnet3d_18
    torch.manual_seed(4)
    ref_res = resnet3d_18(1, in_shape=(45, 41, 43)).double()
    sp_res = copy.deepcopy(ref_res)
    spr = SpatialParallelModel(sp_res)
    xb = torch.randn(2, 1, 45, 41, 43, dtype=torch.float64)
    yb = (torch.rand(2, 1) > 0.5).double()
    crit(ref_res(xb), yb).backward()
    crit(spr(xb), yb).backward()
    spr.reduce_gradients()
    worst = max(float((p.grad - q.grad).abs().max()) for p, q in zip(sp_res.parameters(), ref_res.parameters()))
    worst = max(worst, max(float((a.double() - b.double()).abs().max()) for a, b in zip(sp_res.buffers(), ref_res.buffers())))
    t = torch.tensor([worst], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"spatial-parallel 3-D ResNet: world={world} worst abs err (gradients, BN statistics) {float(t):.3e}", flush=True)
    ok = ok and float(t) < 1e-9
    hb = halo_bytes([(1, 5, 2, 0), (64, 3, 3, 0), (64, 3, 1, 0)], (121, 145, 121), 16, max(world, 2))
    ok = ok and len(hb) == 3 and hb[0] > 0
    t = torch.tensor([1.0 if ok else 0.0])
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"RESULT ok={bool(t.item())}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
