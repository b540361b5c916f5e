"""torchrun worker (gloo, CPU): 2-rank SalientGrads must reproduce the single-process run."""
import logging
import os
import sys
import types

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from neuroimagedisttraining_b200.data import SyntheticABCD, load_partition_data_abcd
from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
from neuroimagedisttraining_b200.models import create_model
from neuroimagedisttraining_b200.parallel import DistContext, init_distributed, shutdown


def run(ctx, ck=None, sizes=(10, 10, 10, 10), per_round=None, balance_eval=False):
    torch.manual_seed(0)
    logger = logging.getLogger("w"); logger.setLevel(logging.ERROR)
    n = len(sizes)
    store = SyntheticABCD(sum(sizes), (32, 32, 32), n_sites=n, seed=1, site_sizes=list(sizes))
    args = types.SimpleNamespace(client_num_in_total=n, client_num_per_round=per_round or n, comm_round=2, epochs=1, lr=0.01, lr_decay=0.998,
                                 momentum=0.9, wd=5e-4, client_optimizer="sgd", batch_size=8, dense_ratio=0.5, itersnip_iteration=2,
                                 stratified_sampling=False, snip_mask=True, ci=0, frequency_of_the_test=1, dataset="ABCD", balance_eval=balance_eval)
    ds = load_partition_data_abcd(store, client_number=n, batch_size=8)
    model = create_model("tiny3DCNN", 1, (32, 32, 32))
    tr = SlabTrainer(model, args, logger, device="cpu", store=store)
    if ck is not None:                 # every rank contributes the shuffle state of the loaders it iterates
        args.checkpoint_every, args.checkpoint_path = 1, ck
    api = SalientGradsAPI(ds, "cpu", args, tr, logger, dist=ctx)
    st = api.train()
    st["_foreign_personal_jobs"] = sum(1 for r, jobs in enumerate(api._eval_plan()) for c, k in jobs if k == "p" and c in api._trained and api.dist.owner(c) != r)
    st["_owners"] = [api.dist.owner(c) for c in range(n)]
    return api.w_global.flat.clone(), api.mask.flat.clone(), st


def main():
    torch.set_num_threads(2)
    ctx = init_distributed(device="cpu")
    import tempfile
    ck = os.path.join(tempfile.gettempdir(), f"nidt_ck_{os.environ.get('MASTER_PORT', '0')}.pt")
    w2, m2, st2 = run(ctx, ck)
    w1, m1, st1 = run(DistContext(device=torch.device("cpu")))     # same thing, single process
    ok = torch.equal(m1, m2) and torch.allclose(w1, w2, rtol=1e-5, atol=1e-6)
    ok = ok and abs(st1["global_test_acc"][-1] - st2["global_test_acc"][-1]) < 1e-9
    ok = ok and st1["sum_comm_params"] == st2["sum_comm_params"]
    # uneven sites, 4 of 5 clients sampled per round: size-aware placement, evaluation jobs moved to the lighter rank (personal
    # models copied owner -> evaluator), a client that never trained evaluated with the shared fall-back model
    w4, m4, st4 = run(ctx, None, sizes=(20, 10, 10, 10, 30), per_round=4, balance_eval=True)
    w3, m3, st3 = run(DistContext(device=torch.device("cpu")), None, sizes=(20, 10, 10, 10, 30), per_round=4)
    ok = ok and torch.equal(m3, m4) and torch.allclose(w3, w4, rtol=1e-5, atol=1e-6)
    for key in ("global_test_acc", "person_test_acc", "global_test_loss", "person_test_loss"):
        ok = ok and len(st3[key]) == len(st4[key]) and all(abs(a - b) < 1e-6 for a, b in zip(st3[key], st4[key]))
    ok = ok and st4["_foreign_personal_jobs"] > 0 and st4["_owners"] != [c % 2 for c in range(5)]
    if ctx.rank == 0:
        print(f"uneven: foreign personal evaluations={st4['_foreign_personal_jobs']} owners={st4['_owners']} person_acc={st4['person_test_acc']} vs {st3['person_test_acc']}")
        saved = torch.load(ck, weights_only=False)
        ok = ok and sorted(saved["loaders"]) == [0, 1, 2, 3] and saved["round"] == 1
        os.remove(ck)
        print(f"RESULT ok={ok} maxdiff={(w1 - w2).abs().max().item():.3e} comm={st1['sum_comm_params']},{st2['sum_comm_params']}")
    shutdown()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
