"""GPU timeline of one FL round (torch.profiler / CUPTI): busy time per kernel name, idle gaps and what precedes them.

    python scripts/trace_round.py [clients] [device|pinned]
"""
import collections, logging, os, sys, types
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import profile, ProfilerActivity
import bench
from neuroimagedisttraining_b200.data import load_partition_data_abcd
from neuroimagedisttraining_b200.fl import SalientGradsAPI, SlabTrainer
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.parallel import DistContext

clients = int(sys.argv[1]) if len(sys.argv) > 1 else 6
mode = sys.argv[2] if len(sys.argv) > 2 else "device"
a = types.SimpleNamespace(clients=clients, per_site=40, batch=16, epochs=2, dense_ratio=0.5, steps=1, warmup=1)
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
if mode == "device":
    tr.set_staging("resident")
for r in range(3):
    api.run_round(r, mask=api.mask)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    api.run_round(3, mask=api.mask)
    torch.cuda.synchronize()
ev = []
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA and e.time_range is not None:
        ev.append((e.time_range.start, e.time_range.end, e.name))
ev.sort()
t0, t1 = ev[0][0], max(e[1] for e in ev)
busy = collections.Counter(); cnt = collections.Counter()
gaps = []
cur_end = ev[0][0]
union = 0.0
for s, e, n in ev:
    short = n.split("(")[0].split("<")[0][-60:]
    busy[short] += e - s; cnt[short] += 1
    if s > cur_end:
        gaps.append((s - cur_end, prev, short))
        union += e - s
        cur_end = e
    elif e > cur_end:
        union += e - cur_end
        cur_end = e
    prev = short
print(f"mode={mode} clients={clients}: span {(t1-t0)/1e3:.2f} ms, GPU busy (union) {union/1e3:.2f} ms, idle {(t1-t0-union)/1e3:.2f} ms, {len(ev)} activities")
def union(iv):
    iv = sorted(iv); tot = 0.0; ce = None; cs = None; out = []
    for a, b in iv:
        if ce is None or a > ce:
            if ce is not None: out.append((cs, ce))
            cs, ce = a, b
        else:
            ce = max(ce, b)
    if ce is not None: out.append((cs, ce))
    return out
def length(u): return sum(b - a for a, b in u)
def inter(u, v):
    i = j = 0; tot = 0.0
    while i < len(u) and j < len(v):
        a, b = max(u[i][0], v[j][0]), min(u[i][1], v[j][1])
        if b > a: tot += b - a
        if u[i][1] < v[j][1]: i += 1
        else: j += 1
    return tot
ku = union([(s_, e_) for s_, e_, n_ in ev if not n_.startswith("Memcpy") and not n_.startswith("Memset")])
mu = union([(s_, e_) for s_, e_, n_ in ev if n_.startswith("Memcpy HtoD")])
print(f"kernels busy {length(ku)/1e3:.2f} ms, H2D copies busy {length(mu)/1e3:.2f} ms, of which under a kernel {inter(ku, mu)/1e3:.2f} ms")
# where are the exposed copies?  walk the timeline: for every H2D interval part not under a kernel, name the last kernel that ended before it
kern_ev = sorted([(s_, e_, n_) for s_, e_, n_ in ev if not n_.startswith("Mem")])
import bisect
starts = [k[0] for k in kern_ev]
exp = collections.Counter(); expn = collections.Counter()
for a, b in mu:
    # subtract kernel union
    segs = [(a, b)]
    for ka, kb in ku:
        if kb <= a or ka >= b: continue
        ns = []
        for x, y_ in segs:
            if ka > x: ns.append((x, min(y_, ka)))
            if kb < y_: ns.append((max(x, kb), y_))
        segs = [(x, y_) for x, y_ in ns if y_ > x]
    for x, y_ in segs:
        i = bisect.bisect_right(starts, x) - 1
        prevk = kern_ev[i][2].split("(")[0].split("<")[0][-40:] if i >= 0 else "<start>"
        nxt = kern_ev[i + 1][2].split("(")[0].split("<")[0][-40:] if i + 1 < len(kern_ev) else "<end>"
        exp[(prevk, nxt)] += y_ - x; expn[(prevk, nxt)] += 1
print("--- exposed H2D time by (kernel before -> kernel after) (ms, count)")
for k_, v in exp.most_common(8):
    print(f"{v/1e3:9.3f}  {expn[k_]:5d}  {k_[0]}  ->  {k_[1]}")
print("--- busy time by kernel (ms, launches)")
for n, t in busy.most_common(28):
    print(f"{t/1e3:9.3f}  {cnt[n]:5d}  {n}")
print("--- idle by (previous -> next) pair (ms, count)")
pair = collections.Counter(); pc = collections.Counter()
for g, p, n in gaps:
    pair[(p, n)] += g; pc[(p, n)] += 1
for (p, n), g in pair.most_common(16):
    print(f"{g/1e3:9.3f}  {pc[(p,n)]:5d}  {p}  ->  {n}")
