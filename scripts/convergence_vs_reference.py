"""Does the bf16 sm_100a engine train like the fp32 reference?  N SalientGrads rounds, same clients / batches / initial weights /
hyper-parameters, dropout off, on a synthetic task with 15 % label noise (accuracy cannot reach 1): this framework's native
engine against the UNMODIFIED reference (baseline/_ref through baseline/ref_harness.py) on the same GPU.

    python scripts/convergence_vs_reference.py [rounds] [out.json]

Prints / writes per round: global_test_loss and global_test_acc of both, the mask Hamming distance, and the final differences.
"""
import json, logging, os, sys, types
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch


def run(rounds=10, clients=4, per_site=60, batch=16, dense_ratio=0.5, lr=0.01, label_noise=0.15, signal_gain=0.6, device="cuda:0"):
    from baseline import ref_harness
    from neuroimagedisttraining_b200.data import SyntheticABCD, site_partition
    from neuroimagedisttraining_b200.data.loaders import IndexLoader
    from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
    dev = torch.device(device)
    store = SyntheticABCD(clients * per_site, n_sites=clients, seed=5, pool=4, pin=True, site_sizes=[per_site] * clients,
                          label_noise=label_noise, signal_gain=signal_gain)
    train, test, _ = site_partition(store.sites)
    args = ref_harness.make_args(client_num_in_total=clients, batch_size=batch, epochs=2, comm_round=rounds, dense_ratio=dense_ratio, lr=lr)
    torch.manual_seed(2024)
    init = AlexNet3D_Dropout(num_classes=1).state_dict()
    # ---- the reference, verbatim (fp32, TF32 convolutions as torch defaults give it), batches in index order
    torch.backends.cudnn.deterministic = True
    class Grab(logging.Handler):          # the reference only LOGS its test losses ({'global_test_acc': .., 'global_test_loss': ..})
        def __init__(self):
            super().__init__()
            self.rows = []

        def emit(self, record):
            if isinstance(record.msg, dict):
                self.rows.append(dict(record.msg))
    grab = Grab()
    ref_logger = logging.getLogger("conv_ref")
    ref_logger.setLevel(logging.INFO)
    ref_logger.propagate = False
    ref_logger.handlers = [grab]
    ref_api, ref_model, _ = ref_harness.build_reference_api("salientgrads", store, train, test, args, dev, logger=ref_logger, shuffle=False,
                                                           model_state={k: v.clone() for k, v in init.items()}, p_drop=0.0)
    masks = []
    orig = ref_api.generate_global_mask_snip
    ref_api.generate_global_mask_snip = lambda: (masks.append(orig()), masks[-1])[1]
    ref_api.train()
    ref = {"loss": [float(r["global_test_loss"]) for r in grab.rows if "global_test_loss" in r],
           "acc": [float(x) for x in ref_api.stat_info["global_test_acc"]],
           "person_acc": [float(x) for x in ref_api.stat_info["person_test_acc"]]}
    ref_mask = masks[0]
    # ---- this framework, native engine (bf16 activations, fp32 master weights)
    logger = logging.getLogger("conv"); logger.setLevel(logging.ERROR)
    model = AlexNet3D_Dropout(num_classes=1, p_drop=0.0)
    model.load_state_dict(init)
    tr = SlabTrainer(model, args, logger, device=dev, store=store, engine=B200Engine(model, dev, max_batch=batch, input_scale=store.scale))
    trl = {c: IndexLoader(t, store.labels[t], store.sites[t], batch, shuffle=False) for c, t in enumerate(train)}
    tel = {c: IndexLoader(t, store.labels[t], store.sites[t], batch, shuffle=False) for c, t in enumerate(test)}
    api = SalientGradsAPI([None, None, None, None, {c: len(t) for c, t in enumerate(train)}, trl, tel, 2], dev, args, tr, logger)
    st = api.train()
    ours = {"loss": [float(x) for x in st["global_test_loss"]], "acc": [float(x) for x in st["global_test_acc"]],
            "person_acc": [float(x) for x in st["person_test_acc"]]}
    total = diff = 0
    for name, m in ref_mask.items():
        if name in api.layout and api.layout[name].prunable:
            d = int((api.mask.view(name) != m.to(dev)).sum())
            diff += d
            total += m.numel()
    n_test = sum(len(t) for t in test)
    out = {"config": {"rounds": rounds, "clients": clients, "subjects_per_client": per_site, "test_samples": n_test, "batch": batch, "lr": lr,
                      "dense_ratio": dense_ratio, "label_noise": label_noise, "signal_gain": signal_gain, "dropout": 0.0,
                      "reference": "unmodified SailentGradsAPI.train(), fp32 (TF32 conv), cudnn.deterministic",
                      "ours": "B200Engine: bf16 activations / fp32 master weights, hand-written sm_100a kernels"},
           "reference": ref, "ours": ours,
           "mask_hamming_fraction": diff / max(1, total), "mask_entries": total,
           "max_abs_loss_diff": max(abs(a - b) for a, b in zip(ref["loss"], ours["loss"])),
           "final_acc_diff": abs(ref["acc"][-1] - ours["acc"][-1]), "one_test_sample_in_acc": 1.0 / n_test}
    return out


if __name__ == "__main__":
    rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 10
    res = run(rounds)
    print("round  ref_loss  our_loss   ref_acc  our_acc")
    for r, (a, b, c, d) in enumerate(zip(res["reference"]["loss"], res["ours"]["loss"], res["reference"]["acc"], res["ours"]["acc"])):
        print(f"{r:5d}  {a:8.4f}  {b:8.4f}   {c:7.4f}  {d:7.4f}")
    print({k: res[k] for k in ("mask_hamming_fraction", "max_abs_loss_diff", "final_acc_diff", "one_test_sample_in_acc")})
    if len(sys.argv) > 2:
        with open(sys.argv[2], "w") as f:
            json.dump(res, f, indent=1)
