"""Determinism / equivalence check at the benchmark's scale: per-round checksums of the global model for the sequential schedule
(twice) and for 2 and 3 concurrent client lanes.   python scripts/check_lanes.py [rounds]"""
import logging, os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
from neuroimagedisttraining_b200.data import load_partition_data_abcd
from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.parallel import DistContext

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 4
a = types.SimpleNamespace(clients=21, per_site=40, batch=16, epochs=2, dense_ratio=0.5, steps=rounds, warmup=0)
dev = torch.device("cuda:0")
store = bench.build_store(a, (121, 145, 121), pin=True)


def run(lanes, staging="client"):
    fl_args = bench.make_fl_args(a, 1)
    logger = logging.getLogger("p"); logger.setLevel(logging.ERROR)
    torch.manual_seed(fl_args.seed)
    ds = load_partition_data_abcd(store, client_number=a.clients, batch_size=a.batch)
    model = AlexNet3D_Dropout(1)
    tr = SlabTrainer(model, fl_args, logger, device=dev, store=store, engine=B200Engine(model, dev, max_batch=a.batch, input_scale=store.scale), staging=staging)
    for _ in range(lanes - 1):
        tr.add_lane(B200Engine(model, dev, max_batch=a.batch, input_scale=store.scale))
    api = SalientGradsAPI(ds, dev, fl_args, tr, logger, dist=DistContext(device=dev))
    api.w_init = api.w_global.clone()
    api.mask = api.generate_global_mask_snip()
    sums = [float(api.mask.flat.double().sum())]
    for r in range(rounds):
        api.run_round(r, mask=api.mask)
        torch.cuda.synchronize()
        sums.append(float(api.w_global.flat.double().sum()))
    return sums, api.w_global.flat.clone(), api.stat_info["global_test_acc"]


ref, w_ref, acc = run(1)
print("lanes=1 A", ref, acc)
for tag, lanes, staging in (("lanes=1 B", 1, "client"), ("lanes=1 resident", 1, "resident"), ("lanes=2", 2, "client"), ("lanes=3", 3, "client")):
    s, w, acc = run(lanes, staging)
    print(tag, "identical" if s == ref and torch.equal(w, w_ref) else f"DIFFERS first at index {next((i for i, (x, y) in enumerate(zip(s, ref)) if x != y), None)} maxdiff {(w - w_ref).abs().max().item():.3e}", acc)
