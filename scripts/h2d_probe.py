"""Host->device copy bandwidth from pinned memory for the transfer sizes the stager uses."""
import torch
dev = "cuda"
vol = 121 * 145 * 121
src = torch.randint(0, 255, (64, vol), dtype=torch.uint8).pin_memory()
dst = torch.empty((16, vol), dtype=torch.uint8, device=dev)
s = torch.cuda.Stream()
def run(fn, n=20):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(s):
        e0.record(s)
        for _ in range(n): fn()
        e1.record(s)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
def per_sample():
    with torch.cuda.stream(s):
        for j in range(16): dst[j].copy_(src[(7 * j) % 64], non_blocking=True)
def one_block():
    with torch.cuda.stream(s):
        dst.copy_(src[:16], non_blocking=True)
for name, fn in (("16 copies of 2.1 MB (scattered rows)", per_sample), ("one 34 MB copy", one_block)):
    ms = run(fn)
    print(f"{name}: {ms:.3f} ms per batch = {16 * vol / ms / 1e6:.1f} GB/s")
