"""Per-kernel CUDA-event timing of one SGD step of the native 3-D ResNet engine at the ABCD shape, and the same step
through the PyTorch engine (cuDNN, bf16 autocast + fp32) for comparison."""
import os, sys, collections, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams, TorchEngine
from neuroimagedisttraining_b200.models.volumetric import resnet3d_18
from neuroimagedisttraining_b200.ops.engines import make_native_engine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
dev = "cuda"
torch.manual_seed(0)
model = resnet3d_18(1)
tmodel = copy.deepcopy(model)
eng = make_native_engine(model, dev, max_batch=B)
slab = eng.new_slab(); eng.attach(slab)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, device=dev)
y = (torch.arange(B, device=dev) % 2).float()
hp = StepHParams(lr=0.01, wd=5e-4)
mask = torch.ones(eng.layout.param_numel, device=dev)


def timeit(e, n):
    for i in range(3):
        e.train_step(x, y, hp, mask, None, i == 0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        e.train_step(x, y, hp, mask, None, False)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print(f"native ResNet3D engine  B={B}: {timeit(eng, steps):.3f} ms/step")
agg = collections.OrderedDict()
for i in range(steps):
    eng.start_profile()
    eng.train_step(x, y, hp, mask, None, False)
    for n, ms in eng.stop_profile():
        agg.setdefault(n, []).append(ms)
tot = 0
by_kind = collections.Counter()
for n, v in agg.items():
    v.sort(); tot += v[0]
    by_kind[n.split("[")[0]] += v[0]
    print(f"{n:44s} {v[0]:9.4f}")
print(f"sum of mins {tot:.3f} ms")
for k, v in by_kind.most_common():
    print(f"  {k:24s} {v:8.3f} ms")
for amp in (True, False):
    te = TorchEngine(copy.deepcopy(tmodel), dev, volumetric=True, input_scale=1 / 255.0, amp=amp)
    te.attach(te.new_slab())
    print(f"TorchEngine (cuDNN, {'bf16 autocast' if amp else 'fp32/tf32'}) B={B}: {timeit(te, steps):.3f} ms/step")
