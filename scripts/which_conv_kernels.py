import os, sys
sys.path.insert(0, "/root/repo")
import torch
from torch.profiler import profile, ProfilerActivity
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.fl.engine import StepHParams
eng = B200Engine(AlexNet3D_Dropout(1), "cuda", max_batch=16)
eng.use_graphs = False
slab = eng.new_slab(); eng.attach(slab)
x = torch.randint(0, 256, (16, 121, 145, 121), dtype=torch.uint8, device="cuda")
y = (torch.arange(16, device="cuda") % 2).float()
hp = StepHParams(lr=0.01, wd=5e-4)
for i in range(3): eng.train_step(x, y, hp, None, None, i == 0)
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    eng.train_step(x, y, hp, None, None, False); torch.cuda.synchronize()
for e in prof.events():
    if "conv3d" in e.name: print(e.name[:70], round(e.cuda_time if hasattr(e,'cuda_time') else e.device_time, 1))
