"""Per-kernel CUDA-event timing of one local SGD step of the B200 engine (AlexNet3D, ABCD shape)."""
import os, sys, collections
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = "cuda"
torch.manual_seed(0)
eng = B200Engine(AlexNet3D_Dropout(1), dev, max_batch=B)
slab = eng.new_slab(); eng.attach(slab)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, device=dev)
y = (torch.arange(B, device=dev) % 2).float()
hp = StepHParams(lr=0.01, wd=5e-4)
mask = torch.ones(eng.layout.param_numel, device=dev)
IDS = list(range(B)) if os.environ.get("PROFILE_IDS", "1") != "0" else None     # sample ids -> the Gram-cached first-layer backward
_step = eng.train_step
def _train_step(*a, **k):
    eng.batch_ids = IDS
    return _step(*a, **k)
eng.train_step = _train_step
for i in range(3):
    eng.train_step(x, y, hp, mask, None, i == 0)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for i in range(steps):
    eng.train_step(x, y, hp, mask, None, False)
e1.record(); torch.cuda.synchronize()
print(f"B={B}: {e0.elapsed_time(e1)/steps:.3f} ms/step (back-to-back, incl. launch gaps)")
agg = collections.OrderedDict()
for i in range(steps):
    eng.start_profile()
    eng.train_step(x, y, hp, mask, None, False)
    for n, ms in eng.stop_profile():
        agg.setdefault(n, []).append(ms)
tot = 0
print(f"{'kernel':32s} {'ms(min)':>9s} {'ms(med)':>9s}")
for n, v in agg.items():
    v.sort(); tot += v[0]
    print(f"{n:32s} {v[0]:9.4f} {v[len(v)//2]:9.4f}")
print(f"sum of mins {tot:.3f} ms (event-bracketed, includes launch latency per kernel)")
# eval batch
from neuroimagedisttraining_b200.fl.engine import EvalCounters
c = EvalCounters(dev)
eng.eval_batch(x, y, c); torch.cuda.synchronize()
e0.record()
for i in range(steps): eng.eval_batch(x, y, c)
e1.record(); torch.cuda.synchronize()
print(f"eval: {e0.elapsed_time(e1)/steps:.3f} ms/batch")
