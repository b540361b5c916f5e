"""Numerics + timing probe of the halo-resident conv kernel against F.conv3d and the per-tap kernel."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from neuroimagedisttraining_b200.ops import volumetric as V
dev = "cuda"
cl = lambda x: x.permute(0, 2, 3, 4, 1).contiguous()
cf = lambda x: x.permute(0, 4, 1, 2, 3).contiguous()
def t_ms(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
CASES = [  # N, Cin, D, H, W, Cout, pad, block_n
    (1, 64, 4, 5, 6, 64, 1, 64),
    (2, 64, 6, 14, 32, 64, 1, 64),
    (2, 64, 9, 13, 11, 128, 0, 128),
    (2, 128, 7, 9, 9, 64, 2, 64),
    (8, 64, 32, 38, 32, 64, 1, 64),       # ResNet layer1
    (16, 64, 19, 23, 19, 128, 0, 128),    # AlexNet conv2 fprop
    (16, 128, 17, 21, 17, 64, 2, 64),     # AlexNet conv2 dgrad
    (8, 128, 16, 19, 16, 128, 1, 128),    # ResNet layer2
    (8, 256, 8, 10, 8, 256, 1, 128),      # ResNet layer3
]
only = int(sys.argv[1]) if len(sys.argv) > 1 else None
for ci, (N, Cin, D, H, W, Cout, pad, bn) in enumerate(CASES):
    if only is not None and ci != only: continue
    torch.manual_seed(ci)
    x = torch.randn(N, Cin, D, H, W, device=dev).bfloat16()
    w = (torch.randn(Cout, Cin, 3, 3, 3, device=dev) / (Cin * 27) ** 0.5).bfloat16().float()
    bias = torch.randn(Cout, device=dev) * 0.1
    ref = F.conv3d(x.float(), w, bias, stride=1, padding=pad)
    wf, wd = V.pack_conv_weights(w)
    xc = cl(x)
    for flags in ((0, 2) if os.environ.get('HALO_PROF') else (0,)):
        y, part = V.conv3d_halo_fprop(xc, wf, bias, pad, block_n=bn, stats=True, debug_flags=flags)
        torch.cuda.synchronize()
        err = (cf(y).float() - ref).abs().max().item() / ref.abs().max().item()
        yb = y.float().reshape(-1, Cout)
        serr = (part[:, 0].sum(0) - yb.sum(0)).abs().max().item() / (yb.sum(0).abs().max().item() + 1e-6)
        s2err = (part[:, 1].sum(0) - (yb * yb).sum(0)).abs().max().item() / (yb * yb).sum(0).abs().max().item()
        print(f"case {ci} {(N,Cin,D,H,W,Cout,pad,bn)} base_offset={'off' if flags == 0 else 'on'}: rel err {err:.4f} stats err {serr:.2e} {s2err:.2e}", flush=True)
    if os.environ.get("HALO_ZERO"): xc = torch.zeros_like(xc)
    if os.environ.get("HALO_NOSTATS"):
        print(f"      halo without stats/bias {t_ms(lambda: V.conv3d_halo_fprop(xc, wf, None, pad, block_n=bn, stats=False))*1e3:.1f} us", flush=True)
    t_new = t_ms(lambda: V.conv3d_halo_fprop(xc, wf, bias, pad, block_n=bn, stats=True))
    t_old = t_ms(lambda: V.conv3d_fprop(xc, wf, bias, 3, 1, pad, block_n=bn, stats=True))
    flops = 2.0 * ref.numel() * Cin * 27
    print(f"      halo {t_new*1e3:.1f} us ({flops/t_new/1e9:.0f} TFLOP/s)   per-tap {t_old*1e3:.1f} us ({flops/t_old/1e9:.0f} TFLOP/s)   (wrappers allocate outputs)", flush=True)
