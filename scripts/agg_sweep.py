"""BASELINE.json config 5: aggregation-only sweep, weighted masked reduce of 1 MB .. 1 GB at N GPUs, with the SM clocks sampled
while every number is taken.

Fused symmetric-memory kernel (csrc/sym_agg.cu) on each of its data paths -- P2P through registers, P2P through shared memory
(bulk-async copies), NVLS multimem, and splits of a chunk between the in-switch and the P2P path -- against the NCCL sequence
(all_reduce + mask multiply + copy).  The fused kernel writes the result where the global slab lives (no write-back pass), the
ranks are lined up by the timing loop itself (``align=False``).  Also times the fused saliency reduce + top-k + mask kernel (W2)
on the FL model's 2.57 M scores.  Launch with torchrun; rank 0 prints one JSON line per size / kernel.

    torchrun ... scripts/agg_sweep.py [sizes_mb] [--k K]      # K > 0: weight K local client slabs inside the kernel
"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import bench
from neuroimagedisttraining_b200 import ops
from neuroimagedisttraining_b200.parallel import init_distributed, shutdown
from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator

ctx = init_distributed()
dev = ctx.device
argv = [a for a in sys.argv[1:] if not a.startswith("--")]
sizes_mb = [int(x) for x in argv[0].split(",")] if argv else [1, 10, 64, 256, 1024]
K = int(sys.argv[sys.argv.index("--k") + 1]) if "--k" in sys.argv else 0
LINK_GBS = 770.0   # measured peer copy, per direction per GPU (B200_PROFILING.md)
P = ctx.world_size


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
    return round(x.item(), 4)


clk = bench.ClockSampler(ctx.local_rank)
if ctx.is_main:
    clk.start()
for mb in sizes_mb:
    n = mb * (1 << 20) // 4
    agg = SymmetricAggregator(ctx, n, dev)
    mask = (torch.rand(n, device=dev) > 0.5).float()
    part = torch.randn(n, device=dev)
    agg.local_partial.copy_(part)
    locs = [(1.0 / (K * P), torch.randn(n, device=dev)) for _ in range(K)]
    out = agg.local_result                 # in place: the all-gather is the write of the global slab
    reps = 30 if mb <= 64 else 8
    row = {"kernel": "W1 agg_weighted_reduce_mask_bcast", "size_mb": mb, "n_gpus": P, "local_clients_weighted_in_kernel": K}
    variants = [("p2p", 0.0, 0), ("bulk", 0.0, 0)]
    if agg.has_multicast:
        variants += [("multimem", 0.0, 0), ("multimem", 0.0, 148), ("bulk", 0.33, 0), ("bulk", 0.5, 0)]
    times = {}
    for mode, frac, ctas in variants:
        name = mode + (f"+mm{frac:g}" if frac else "") + (f"@{ctas}" if ctas else "")
        times[name] = timeit(lambda: agg.aggregate(out, locs, mask, align=False, mode=mode, mm_frac=frac, ctas=ctas, use_partial=(K == 0)), reps)
    times["auto"] = timeit(lambda: agg.aggregate(out, locs, mask, align=False, use_partial=(K == 0)), reps)
    row["fused_ms"] = times
    row["auto_choice"] = agg.pick_mode(n)

    other = torch.empty(n, device=dev)

    def nccl():
        if K:
            ops.weighted_aggregate(part, [t for _, t in locs], [w for w, _ in locs], None)
        dist.all_reduce(part)
        part.mul_(mask)
        other.copy_(part)
    row["nccl_allreduce_mask_copy_ms"] = timeit(nccl, reps)
    bytes_link = n * 4 * (P - 1) / P            # per direction per GPU with in-switch reduction: reduce-scatter out, all-gather in
    roof = bytes_link / (LINK_GBS * 1e9) * 1e3
    best = min(times.values())
    row.update({"nvlink_roofline_ms": round(roof, 4), "best": min(times, key=times.get), "fraction_of_roofline": round(roof / best, 3),
                "fraction_of_p2p_roofline": round(2 * roof / min(v for k_, v in times.items() if not k_.startswith("multimem")), 3),
                "speedup_vs_nccl": round(row["nccl_allreduce_mask_copy_ms"] / best, 2)})
    if ctx.is_main:
        print(json.dumps(row), flush=True)
    del agg, mask, part, out, other, locs
    torch.cuda.empty_cache()

# W2 on the FL model
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.runtime.slab import SlabLayout
layout = SlabLayout.from_module(AlexNet3D_Dropout(1))
Pn = layout.param_numel
agg = SymmetricAggregator(ctx, layout.total_numel, dev)
flags = ops.prunable_flags(layout, dev)
sc = (torch.rand(Pn, device=dev) ** 3) * 1e-3
mask_out = torch.empty(Pn, device=dev)
work = sc.clone()
t_w2 = {}
for rep in ("0", "1"):              # sharded histogram exchange (6 cross-rank barriers) / select on every rank after an in-kernel all-gather (2)
    os.environ["NIDT_SAL_REPLICATE"] = rep
    for mm in ([False, True] if agg.has_multicast else [False]):
        agg.use_multimem = mm
        t_w2[("multimem" if mm else "p2p") + ("+replicated_select" if rep == "1" else "")] = timeit(
            lambda: agg.saliency_topk_mask(work, layout, 0.5, mask_out, flags, read=False, align=False), 30)
os.environ.pop("NIDT_SAL_REPLICATE", None)


def nccl_w2():
    dist.all_reduce(work)
    ops.global_topk_mask(work, layout, 0.5, mask_out, flags)
t_nccl = timeit(nccl_w2, 30)
bytes_link = 4 * Pn * (P - 1) / P + 4 * Pn * (P - 1) / P          # scores in (reduce-scatter) + mask out (all-gather), per direction
roof = 4 * Pn * (P - 1) / P / (LINK_GBS * 1e9) * 1e3
if ctx.is_main:
    clocks = clk.stop()
    print(json.dumps({"kernel": "W2 saliency_reduce_topk_mask", "scores": Pn, "bytes_scores": 4 * Pn, "n_gpus": P, "fused_ms": t_w2,
                      "nccl_allreduce_plus_topk_ms": t_nccl, "nvlink_roofline_ms": round(roof, 4),
                      "fraction_of_roofline": round(roof / min(t_w2.values()), 3), "speedup_vs_nccl": round(t_nccl / min(t_w2.values()), 2),
                      "note": "sharded: 6 cross-rank barriers + 3 radix passes over n/P scores; replicated_select: 2 barriers, passes over all n scores"}))
    print(json.dumps({"clocks": clocks, "applies_to": "every line of this run (sampled over the whole sweep)"}))
shutdown()
