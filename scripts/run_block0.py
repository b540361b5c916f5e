"""Runs the block-0 kernels (conv_in fprop/wgrad, BN+ReLU+pool fwd/bwd) a few times -- ncu target."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.ops import volumetric as V
B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
dev = "cuda"
torch.manual_seed(0)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, device=dev)
w = torch.randn(64, 1, 5, 5, 5, device=dev) * 0.09
bias = torch.zeros(64, device=dev)
gamma, beta = torch.ones(64, device=dev), torch.zeros(64, device=dev)
for _ in range(reps):
    y, part = V.conv_in_s2d_fprop(x, w, bias, 1 / 255.0, stats=True)
    scale, shift, mean, invstd = V.bn_finalize(part, y.numel() / 64, gamma, beta)
    a, ymax = V.bn_relu_pool_fwd(y, scale, shift, 3, 3, True, want_ymax=True)
    da = torch.randn_like(a)
    dy, dg, db = V.bn_relu_pool_bwd(y, da, scale, shift, mean, invstd, 3, 3, True, ymax=ymax)
    dw = V.conv_in_s2d_wgrad(x, dy, 1 / 255.0)
torch.cuda.synchronize()
print("ok", float(dw.abs().sum()))
