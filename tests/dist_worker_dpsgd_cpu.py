"""torchrun worker (gloo, CPU): D-PSGD (ring and seeded-random neighbours) and FedFomo with clients spread over ranks == the
single-process runs."""
import logging
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from neuroimagedisttraining_b200.data import load_partition_data_images
from neuroimagedisttraining_b200.fl import DPSGDAPI, FedFomoAPI, SlabTrainer
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.parallel import DistContext, init_distributed, shutdown


def run(ctx, cs, API=DPSGDAPI):
    import random
    torch.manual_seed(0); np.random.seed(0); random.seed(0)
    logger = logging.getLogger("w"); logger.setLevel(logging.ERROR)
    n = 5
    args = types.SimpleNamespace(client_num_in_total=n, client_num_per_round=2, comm_round=3, epochs=1, lr=0.05, lr_decay=0.998, momentum=0.0,
                                 wd=5e-4, client_optimizer="sgd", batch_size=16, cs=cs, ci=0, frequency_of_the_test=1, dataset="cifar10", type="epoch")
    ds = load_partition_data_images("cifar10", None, "dir", 0.5, n, 16, synthetic_size=(300, 100), val_split=0.1 if API is FedFomoAPI else 0.0)
    model = create_model("lenet5_cifar", 10)
    api = API(ds, "cpu", args, SlabTrainer(model, args, logger, device="cpu", loss_kind="ce"), logger, dist=ctx)
    api.record_information = lambda *a, **k: None
    st = api.train()
    return api, st


def main():
    torch.set_num_threads(2)
    ctx = init_distributed(device="cpu")
    ok = True
    for cs in ("ring", "random"):
        multi, sm = run(ctx, cs)
        single, ss = run(DistContext(device=torch.device("cpu")), cs)
        good = torch.allclose(multi.w_global.flat, single.w_global.flat, rtol=1e-5, atol=1e-6)
        good = good and all(torch.allclose(a.slab.flat, b.slab.flat, rtol=1e-5, atol=1e-6) for a, b in zip(multi.client_list, single.client_list))
        for key in ("global_test_acc", "person_test_acc"):
            good = good and len(sm[key]) == len(ss[key]) == 3 and all(abs(float(p) - float(q)) < 1e-9 for p, q in zip(sm[key], ss[key]))
        if ctx.rank == 0:
            print(f"dpsgd {cs}: world={ctx.world_size} match={good}", flush=True)
        ok = ok and good
    multi, sm = run(ctx, "random", FedFomoAPI)
    single, ss = run(DistContext(device=torch.device("cpu")), "random", FedFomoAPI)
    good = all(torch.allclose(a.slab.flat, b.slab.flat, rtol=1e-4, atol=1e-5) for a, b in zip(multi.client_list, single.client_list))
    for key in ("old_mask_test_acc", "new_mask_test_acc"):
        good = good and len(sm[key]) == len(ss[key]) == 3 and all(abs(float(p) - float(q)) < 1e-9 for p, q in zip(sm[key], ss[key]))
    good = good and np.allclose(multi._fomo_state[0], single._fomo_state[0], rtol=1e-3, atol=1e-6) and np.allclose(multi._fomo_state[1], single._fomo_state[1], rtol=1e-3, atol=1e-6)
    if ctx.rank == 0:
        print(f"fedfomo: world={ctx.world_size} match={good}", flush=True)
    ok = ok and good
    t = torch.tensor([1.0 if ok else 0.0])
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN)
    if ctx.rank == 0:
        print(f"RESULT ok={bool(t.item())}", flush=True)
    shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
