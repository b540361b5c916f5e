"""torchrun worker (NCCL, >= 2 GPUs): DataParallelB200Engine over the process group == whole-batch training on one engine
(tests/dp_check.py), every rank ends with bit-identical weights; then a whole `--data_parallel` SalientGrads job from the CLI."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import torch
import torch.distributed as dist

import dp_check
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.ops.engine_dp import DataParallelB200Engine, DistGroup
from neuroimagedisttraining_b200.parallel import init_distributed, shutdown


def main():
    ctx = init_distributed()
    dev = ctx.device
    x, y = dp_check.make_inputs(dev)
    world = ctx.world_size
    plain = B200Engine(dp_check.make_model(), dev, max_batch=8)
    eng = DataParallelB200Engine(dp_check.make_model(), dev, DistGroup(None), max_batch=8)
    # sharp: identical shares on every rank == one engine on one share
    ok_sharp, errs = dp_check.check_sharp(dp_check.run_sharp_dp(eng, x, y, world), dp_check.run_sharp_plain(plain, x, y), world)
    # coarse: ordinary / ragged batches vs whole-batch training
    got = dp_check.programme(eng, x, y, dp_check.BATCHES, 6)
    ok_coarse, errs2 = dp_check.check_coarse(got, dp_check.programme(plain, x, y, dp_check.BATCHES, 6))
    ok = ok_sharp and ok_coarse
    errs = {"sharp": errs, "coarse": errs2}
    w0 = got["w"].clone()
    dist.broadcast(w0, 0)
    same = torch.equal(w0, got["w"])
    t = torch.tensor([1.0 if (ok and same) else 0.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    print(f"rank {ctx.rank}: errs={errs} same_weights_as_rank0={same}", flush=True)
    if ctx.rank == 0:
        print(f"RESULT ok={bool(t.item())}", flush=True)
    shutdown()
    sys.exit(0 if t.item() else 1)


if __name__ == "__main__":
    main()
