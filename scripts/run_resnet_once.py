"""Two native ResNet3D training steps at the ABCD shape (ncu target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams
from neuroimagedisttraining_b200.models.volumetric import resnet3d_18
from neuroimagedisttraining_b200.ops.engines import make_native_engine
B = int(sys.argv[1]) if len(sys.argv) > 1 else 8
torch.manual_seed(0)
eng = make_native_engine(resnet3d_18(1), "cuda", max_batch=B)
slab = eng.new_slab(); eng.attach(slab)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, device="cuda")
y = (torch.arange(B, device="cuda") % 2).float()
for i in range(2):
    eng.train_step(x, y, StepHParams(lr=0.01, wd=5e-4), None, None, i == 0)
torch.cuda.synchronize()
print("ok", eng.pop_mean_loss())
