"""torchrun worker (gloo, CPU): Local, Ditto, FedAvg, TurboAggregate and Sub-FedAvg on an image dataset with clients spread over ranks == the
single-process runs (global model and accuracy series)."""
import logging
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from neuroimagedisttraining_b200.data import load_partition_data_images
from neuroimagedisttraining_b200.fl import DittoAPI, FedAvgAPI, LocalAPI, SlabTrainer, SubAvgAPI, TurboAggregateAPI
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.parallel import DistContext, init_distributed, shutdown


def run(ctx, API, rounds=2, ck=None, resume=None):
    torch.manual_seed(0); np.random.seed(0)
    logger = logging.getLogger("w"); logger.setLevel(logging.ERROR)
    n = 5
    args = types.SimpleNamespace(client_num_in_total=n, client_num_per_round=3, comm_round=rounds, epochs=1, lr=0.05, lr_decay=0.998, momentum=0.0,
                                 wd=5e-4, client_optimizer="sgd", batch_size=16, ci=0, frequency_of_the_test=1, dataset="cifar10", lamda=0.5,
                                 local_epochs=1, final_finetune=False, ta_group_size=2, ta_scale=1 << 16, each_prune_ratio=0.2, dense_ratio=0.3,
                                 dist_thresh=1e-4, acc_thresh=-1.0)
    ds = load_partition_data_images("cifar10", None, "dir", 0.5, n, 16, synthetic_size=(300, 100))
    model = create_model("lenet5_cifar", 10)
    api = API(ds, "cpu", args, SlabTrainer(model, args, logger, device="cpu", loss_kind="ce"), logger, dist=ctx)
    api.record_information = lambda *a, **k: None
    if ck is not None:
        args.checkpoint_every, args.checkpoint_path = 1, ck
    if resume is not None:
        api.resume(resume)
    st = api.train()
    return api, st


def resume_check(ctx):
    """Local-only training over 2 ranks, killed after round 1 and resumed from the checkpoint, against the uninterrupted 3-round run:
    the checkpoint must hold EVERY client's personal model (fetched from its owner), not rank 0's untrained clones of them."""
    import tempfile
    ck = os.path.join(tempfile.gettempdir(), f"nidt_ck_local_{os.environ.get('MASTER_PORT', '0')}.pt")
    full, sf = run(ctx, LocalAPI, rounds=3)
    run(ctx, LocalAPI, rounds=2, ck=ck)
    torch.distributed.barrier()
    saved = torch.load(ck, weights_only=False)
    good = sorted(saved["personal"]) == list(range(5)) and saved["round"] == 1
    res, sr = run(ctx, LocalAPI, rounds=3, resume=ck)
    for c in range(5):
        if ctx.owns(c):
            good = good and torch.allclose(res.client_list[c].slab.flat, full.client_list[c].slab.flat, rtol=1e-5, atol=1e-6)
    good = good and abs(sr["person_test_acc"][-1] - sf["person_test_acc"][-1]) < 1e-9 and sr["sum_training_flops"] == sf["sum_training_flops"]
    torch.distributed.barrier()
    if ctx.rank == 0:
        os.remove(ck)
        print(f"LocalAPI save/resume over {ctx.world_size} ranks: match={good}", flush=True)
    return good


def main():
    torch.set_num_threads(2)
    ctx = init_distributed(device="cpu")
    ok = True
    for API in (LocalAPI, DittoAPI, FedAvgAPI, TurboAggregateAPI, SubAvgAPI):
        multi, sm = run(ctx, API)
        single, ss = run(DistContext(device=torch.device("cpu")), API)
        tol = dict(rtol=1e-3, atol=1e-4) if API is TurboAggregateAPI else dict(rtol=1e-5, atol=1e-6)     # fixed-point field arithmetic
        good = torch.allclose(multi.w_global.flat, single.w_global.flat, **tol)
        for key in ("global_test_acc", "person_test_acc", "test_acc"):
            a, b = sm.get(key, []), ss.get(key, [])
            good = good and len(a) == len(b) and all(abs(float(p) - float(q)) < (2e-2 if API is TurboAggregateAPI else 1e-9) for p, q in zip(a, b))
        good = good and (len(sm.get("person_test_acc", [])) > 0)
        if API is SubAvgAPI:
            good = good and all(torch.equal(a.flat, b.flat) for a, b in zip(multi.mask_pers, single.mask_pers))
            good = good and any(float(m.flat.sum()) < m.flat.numel() for m in single.mask_pers)       # pruning did happen
        if ctx.rank == 0:
            print(f"{API.__name__}: world={ctx.world_size} match={good}", flush=True)
        ok = ok and good
    ok = resume_check(ctx) and ok
    t = torch.tensor([1.0 if ok else 0.0])
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MIN)
    if ctx.rank == 0:
        print(f"RESULT ok={bool(t.item())}", flush=True)
    shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
