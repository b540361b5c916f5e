"""Timing of the first-layer backward without y / dY passes (csrc/conv_in_alg.cu) against the dY-based kernels: whole SGD step in
both modes (CUDA-graph replays), the per-sample Gram matrices cold, the per-batch sum, and the block-0 kernels through the Python
launch loop.  NIDT_ALG_DEBUG=1 / 2 drops the sparse pass / the combine kernel (isolates the other one); NIDT_ALG_CTAS=2|3.

    python scripts/time_alg_bwd0.py
"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
DEV = torch.device("cuda:0")
B = 16
torch.manual_seed(0)
eng = B200Engine(AlexNet3D_Dropout(1), DEV, max_batch=B)
eng.attach(eng.new_slab())
g = torch.Generator().manual_seed(0)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g).to(DEV)
y = (torch.arange(B) % 2).float().to(DEV)
hp = StepHParams(lr=0.01, momentum=0.0, wd=5e-4, clip_norm=10.0)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
ids = list(range(B))
def step_ids():
    eng.batch_ids = ids
    eng.train_step(x, y, hp, None, None, False)
def step_std():
    eng.train_step(x, y, hp, None, None, False)
print("step mode0 ms", round(timeit(step_std), 4))
print("step mode2 ms", round(timeit(step_ids), 4), "built", eng.gram.built)
st = torch.cuda.current_stream().cuda_stream
def gram_cold():
    eng.gram.slots.clear()
    eng.gram.sum_for(ids, x, st)
print("gram 16 samples cold ms", round(timeit(gram_cold, 5), 4))
print("gram sum ms", round(timeit(lambda: eng.gram.sum_for(ids, x, st)), 4))
# per-kernel through the python loop
eng.native_exec = False
eng.use_graphs = False
eng.batch_ids = ids
eng.start_profile()
eng.batch_ids = ids
eng.train_step(x, y, hp, None, None, False)
for n, ms in eng.stop_profile():
    if "[0]" in n or "alg" in n or "stats" in n:
        print(f"  {n:32s} {ms:.4f}")
