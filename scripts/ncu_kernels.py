"""Launches every kernel family once or twice on ONE GPU so a single `ncu --set full` run can capture them
(ncu must never wrap a multi-rank command: the symmetric-memory kernels run here with world size 1, where
the peer pointers are the rank's own buffers and the cross-rank barrier degenerates to a self hand-shake)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29577")
os.environ.setdefault("RANK", "0"); os.environ.setdefault("WORLD_SIZE", "1"); os.environ.setdefault("LOCAL_RANK", "0")
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.ops import flat as flat_ops
from neuroimagedisttraining_b200.parallel import init_distributed
from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator

B = 16
ctx = init_distributed()
dev = ctx.device
torch.manual_seed(0)
eng = B200Engine(AlexNet3D_Dropout(1), dev, max_batch=B)
slab = eng.new_slab(); eng.attach(slab)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, device=dev)
y = (torch.arange(B, device=dev) % 2).float()
hp = StepHParams(lr=0.01, wd=5e-4)
lay = eng.layout
mask = torch.ones(lay.param_numel, device=dev)
for i in range(2):
    eng.train_step(x, y, hp, mask, None, i == 0)
torch.cuda.synchronize()
torch.cuda.profiler.start()      # ncu --profile-from-start off
eng.train_step(x, y, hp, mask, None, False)
# flat ops: saliency, top-k, aggregation, nnz
scores = torch.zeros(lay.param_numel, device=dev)
flat_ops.saliency_accumulate(scores, eng.grad.flat, slab.params_flat, 1.0)
flags = flat_ops.prunable_flags(lay, dev)
m2, info = flat_ops.global_topk_mask(scores, lay, 0.5, torch.empty_like(scores), flags=flags)
others = [slab.clone() for _ in range(3)]
out = slab.clone()
flat_ops.weighted_aggregate(out.flat, [o.flat for o in others], [0.2, 0.3, 0.5], mask=m2)
flat_ops.count_nonzero(out.flat)
# symmetric-memory kernels at world size 1 (P2P template and, if the platform has a multicast address, multimem)
import torch.distributed as dist
if not dist.is_initialized():
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
agg = SymmetricAggregator(ctx, slab.flat.numel(), dev, use_multimem=False)
agg.local_partial.copy_(slab.flat)
agg.reduce_mask_broadcast(out.flat, m2)
agg.saliency_topk_mask(scores.clone(), lay, 0.5, torch.empty_like(m2), flags)
if agg.mc_partial:
    agg.use_multimem = True
    agg.reduce_mask_broadcast(out.flat, m2)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("ok kept", info.get("kept"))
if dist.is_initialized():
    dist.destroy_process_group()
