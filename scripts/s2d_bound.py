import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.ops import volumetric as V
dev = "cuda"
x = torch.randint(0, 256, (16, 121, 145, 121), dtype=torch.uint8, device=dev)
w = torch.randn(64, 1, 5, 5, 5, device=dev) * 0.09
dy = torch.randn(16, 59, 71, 59, 64, device=dev).bfloat16()
def t(fn):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / 5
print(os.environ.get("NIDT_S2D_DEBUG", "0"), "wgrad ms", round(t(lambda: V.conv_in_s2d_wgrad(x, dy, 1 / 255.0)), 4), "fprop ms", round(t(lambda: V.conv_in_s2d_fprop(x, w, None, 1 / 255.0, stats=True)), 4))
