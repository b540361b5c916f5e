"""BASELINE.json config 5: aggregation-only sweep, weighted masked reduce of 1 MB .. 1 GB at N GPUs.
Fused symmetric-memory kernel (P2P and NVLS multimem paths) vs. the NCCL path (all_reduce + mask multiply + copy).
Launch with torchrun; rank 0 prints one JSON line per size."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from neuroimagedisttraining_b200.parallel import init_distributed, shutdown
from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator

ctx = init_distributed()
dev = ctx.device
sizes_mb = [1, 4, 16, 64, 256, 1024]
if len(sys.argv) > 1:
    sizes_mb = [int(x) for x in sys.argv[1].split(",")]
LINK_GBS = 770.0   # measured peer copy, per direction per GPU (B200_PROFILING.md)


def timeit(fn, reps):
    for _ in range(3):
        fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record(); torch.cuda.synchronize()
    x = torch.tensor([e0.elapsed_time(e1) / reps], device=dev)
    dist.all_reduce(x, op=dist.ReduceOp.MAX)
    return x.item()


for mb in sizes_mb:
    n = mb * (1 << 20) // 4
    agg = SymmetricAggregator(ctx, n, dev)
    mask = (torch.rand(n, device=dev) > 0.5).float()
    part = torch.randn(n, device=dev)
    agg.local_partial.copy_(part)
    out = torch.empty(n, device=dev)
    reps = 20 if mb <= 64 else 5
    agg.use_multimem = False
    t_p2p = timeit(lambda: agg.reduce_mask_broadcast(out, mask), reps)
    t_mm = None
    if agg.mc_partial:
        agg.use_multimem = True
        t_mm = timeit(lambda: agg.reduce_mask_broadcast(out, mask), reps)

    def nccl():
        dist.all_reduce(part)
        part.mul_(mask)
        out.copy_(part)
    t_nccl = timeit(nccl, reps)
    P = ctx.world_size
    bytes_link = n * 4 * (P - 1) / P            # per direction per GPU: reduce-scatter in, all-gather out (they overlap directions)
    roof_ms = bytes_link / (LINK_GBS * 1e9) * 1e3
    best = min(t for t in (t_p2p, t_mm) if t is not None)
    if ctx.is_main:
        print(json.dumps({"size_mb": mb, "n_gpus": P, "fused_p2p_ms": round(t_p2p, 4), "fused_multimem_ms": None if t_mm is None else round(t_mm, 4),
                          "nccl_allreduce_mask_copy_ms": round(t_nccl, 4), "nvlink_roofline_ms": round(roof_ms, 4),
                          "fraction_of_roofline": round(roof_ms / best, 3), "speedup_vs_nccl": round(t_nccl / best, 2)}))
    del agg, mask, part, out
    torch.cuda.empty_cache()
shutdown()
