"""cProfile of one FL round through the public API (where does host time go?)."""
import cProfile, pstats, io, os, sys, logging, types, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench
a = types.SimpleNamespace(clients=6, per_site=40, batch=16, epochs=2, dense_ratio=0.5, steps=1, warmup=1)
from neuroimagedisttraining_b200.data import load_partition_data_abcd
from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.parallel import DistContext
dev = torch.device("cuda:0")
fl_args = bench.make_fl_args(a, 1)
logger = logging.getLogger("p"); logger.setLevel(logging.ERROR)
store = bench.build_store(a, (121, 145, 121), pin=True)
ds = load_partition_data_abcd(store, client_number=a.clients, batch_size=a.batch)
model = AlexNet3D_Dropout(1)
eng = B200Engine(model, dev, max_batch=a.batch, input_scale=store.scale)
tr = SlabTrainer(model, fl_args, logger, device=dev, store=store, engine=eng)
api = SalientGradsAPI(ds, dev, fl_args, tr, logger, dist=DistContext(device=dev))
api.w_init = api.w_global.clone()
api.mask = api.generate_global_mask_snip()
api.run_round(0, mask=api.mask)
torch.cuda.synchronize()
for mode in ("pinned-host", "device-resident"):
    if mode == "device-resident":
        tr.set_staging("resident")
        api.run_round(1, mask=api.mask); torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    api.run_round(2, mask=api.mask)
    pr.disable()
    t_cpu = time.perf_counter() - t0
    torch.cuda.synchronize()
    t_all = time.perf_counter() - t0
    s = io.StringIO()
    pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(22)
    print(f"==== {mode}: host issue time {t_cpu*1e3:.1f} ms, until GPU idle {t_all*1e3:.1f} ms  ({a.clients} clients, {a.clients*4} steps)")
    print("\n".join(s.getvalue().splitlines()[4:40]))
