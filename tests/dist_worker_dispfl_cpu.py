"""torchrun worker (gloo, CPU): DisPFL over 2-3 ranks must reproduce the single-process run -- every client's personal model and
mask, both accuracy series and the mask Hamming matrix -- with and without client drop-out / neighbour aggregation."""
import logging
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd
from neuroimagedisttraining_b200.fl import DisPFLAPI, SlabTrainer
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.parallel import DistContext, init_distributed, shutdown


def run(ctx, **kw):
    torch.manual_seed(0); np.random.seed(0)
    logger = logging.getLogger("w"); logger.setLevel(logging.ERROR)
    n = 5
    store = SyntheticABCD(10 * n, (32, 32, 32), n_sites=n, seed=1, site_sizes=[10] * n)
    a = dict(client_num_in_total=n, client_num_per_round=2, comm_round=3, epochs=1, lr=0.02, lr_decay=0.998, momentum=0.0, wd=5e-4,
             client_optimizer="sgd", batch_size=4, dense_ratio=0.5, anneal_factor=0.4, cs="random", active=0.7, static=False, uniform=False,
             different_initial=True, diff_spa=False, dis_gradient_check=False, save_masks=True, ci=0, frequency_of_the_test=1, dataset="ABCD",
             erk_power_scale=1.0, neighbor_aggregation=False)
    a.update(kw)
    args = types.SimpleNamespace(**a)
    ds = load_partition_data_abcd(store, client_number=n, batch_size=4)
    model = create_model("tiny3DCNN", 1, (32, 32, 32))
    tr = SlabTrainer(model, args, logger, device="cpu", store=store)
    api = DisPFLAPI(ds, "cpu", args, tr, logger, dist=ctx)
    st = api.train()
    return api, st


def same(a, sa, b, sb):
    ok = all(torch.allclose(x.slab.flat, y.slab.flat, rtol=1e-5, atol=1e-6) for x, y in zip(a.client_list, b.client_list))
    ok = ok and all(torch.equal(x.flat, y.flat) for x, y in zip(a.masks_per_client, b.masks_per_client))
    ok = ok and sa["mask_dis_matrix"] == sb["mask_dis_matrix"]
    for key in ("old_mask_test_acc", "new_mask_test_acc"):
        ok = ok and len(sa[key]) == len(sb[key]) == 3 and all(abs(float(p) - float(q)) < 1e-9 for p, q in zip(sa[key], sb[key]))
    return ok


def main():
    torch.set_num_threads(2)
    ctx = init_distributed(device="cpu")
    ok = True
    for kw in ({}, {"neighbor_aggregation": True, "active": 1.0}):
        multi, sm = run(ctx, **kw)
        single, ss = run(DistContext(device=torch.device("cpu")), **kw)
        good = same(multi, sm, single, ss)
        if ctx.rank == 0:
            print(f"dispfl {kw or 'default'}: world={ctx.world_size} match={good}", flush=True)
        ok = ok and good
    t = torch.tensor([1.0 if ok else 0.0])
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN)
    if ctx.rank == 0:
        print(f"RESULT ok={bool(t.item())}", flush=True)
    shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
