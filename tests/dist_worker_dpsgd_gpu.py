"""torchrun worker (NCCL, >= 2 GPUs): D-PSGD with neighbour averaging over the peer-mapped model arena (W11: the mixing kernel reads
the neighbours' models out of their owners' memory) == the replicate-and-broadcast form, for ring, random and mixing-matrix topologies."""
import logging
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from neuroimagedisttraining_b200.data import load_partition_data_images
from neuroimagedisttraining_b200.fl import DPSGDAPI, SlabTrainer
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.parallel import init_distributed, shutdown


def run(ctx, cs, arena):
    torch.manual_seed(0); np.random.seed(0)
    torch.backends.cudnn.deterministic = True
    logger = logging.getLogger("w"); logger.setLevel(logging.ERROR)
    n = 6
    args = types.SimpleNamespace(client_num_in_total=n, client_num_per_round=2, comm_round=3, epochs=1, lr=0.05, lr_decay=0.998, momentum=0.0,
                                 wd=5e-4, client_optimizer="sgd", batch_size=16, ci=0, frequency_of_the_test=1, dataset="cifar10", cs=cs, seed=0,
                                 peer_arena=arena)
    ds = load_partition_data_images("cifar10", None, "dir", 0.5, n, 16, synthetic_size=(300, 100))
    model = create_model("lenet5_cifar", 10)
    api = DPSGDAPI(ds, ctx.device, args, SlabTrainer(model, args, logger, device=ctx.device, loss_kind="ce"), logger, dist=ctx)
    api.record_information = lambda *a, **k: None
    st = api.train()
    return api, st


def main():
    ctx = init_distributed()
    ok = True
    for cs in ("ring", "random", "topology_sym"):
        a, sa = run(ctx, cs, True)
        b, sb = run(ctx, cs, False)
        good = a.arena is not None and b.arena is None and a.arena.bytes_pulled > 0
        good = good and torch.allclose(a.w_global.flat, b.w_global.flat, rtol=1e-4, atol=1e-5)
        for c in range(6):
            if ctx.owns(c):
                good = good and torch.allclose(a.client_list[c].slab.flat, b.client_list[c].slab.flat, rtol=1e-4, atol=1e-5)
        good = good and all(abs(float(p) - float(q)) < 2e-2 for p, q in zip(sa["global_test_acc"], sb["global_test_acc"]))
        t = torch.tensor([1.0 if good else 0.0], device=ctx.device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN)
        if ctx.rank == 0:
            print(f"dpsgd cs={cs}: arena pulled {a.arena.bytes_pulled / 1e6:.1f} MB on rank 0, match={bool(t.item())}", flush=True)
        ok = ok and bool(t.item())
    if ctx.rank == 0:
        print(f"RESULT ok={ok}", flush=True)
    shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
