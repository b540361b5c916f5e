"""torchrun worker (gloo, CPU): SynchronizedBatchNorm on a batch split over ranks == nn.BatchNorm3d on the whole batch (outputs,
input / affine gradients after summing over ranks, running statistics)."""
import copy
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.nn as nn

from neuroimagedisttraining_b200.parallel import convert_sync_batchnorm


def main():
    dist.init_process_group("gloo")
    world, rank = dist.get_world_size(), dist.get_rank()
    torch.manual_seed(0)
    net = nn.Sequential(nn.Conv3d(2, 4, 3, padding=1), nn.BatchNorm3d(4), nn.ReLU(), nn.Conv3d(4, 3, 1), nn.BatchNorm3d(3)).double()
    ref = copy.deepcopy(net)
    x = torch.randn(2 * world + 1, 2, 5, 6, 4, dtype=torch.float64)          # uneven split: the last rank gets one sample more
    w = torch.randn(3, dtype=torch.float64)
    xr = x.clone().requires_grad_(True)
    (ref(xr) * w.view(1, 3, 1, 1, 1)).sum().backward()

    sync = convert_sync_batchnorm(net)
    lo, hi = 2 * rank, (2 * rank + 2 if rank < world - 1 else x.shape[0])
    xs = x[lo:hi].clone().requires_grad_(True)
    (sync(xs) * w.view(1, 3, 1, 1, 1)).sum().backward()
    errs = [float((xs.grad - xr.grad[lo:hi]).abs().max())]
    for (n, p), (_, q) in zip(sync.named_parameters(), ref.named_parameters()):
        g = p.grad.clone()
        dist.all_reduce(g)
        errs.append(float((g - q.grad).abs().max()))
    for (n, p), (_, q) in zip(sync.named_buffers(), ref.named_buffers()):
        errs.append(float((p.double() - q.double()).abs().max()))
    sync.eval(); ref.eval()
    with torch.no_grad():
        errs.append(float((sync(x) - ref(x)).abs().max()))
    t = torch.tensor([max(errs)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"syncbn world={world} worst abs err {float(t):.3e}")
        print(f"RESULT ok={float(t) < 1e-10}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
