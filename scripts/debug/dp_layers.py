import os, sys, threading
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import dp_check
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine
from neuroimagedisttraining_b200.ops.engine_dp import DataParallelB200Engine, ThreadGroup
DEV = torch.device("cuda:0")
x, y = dp_check.make_inputs(DEV)
ref_eng = B200Engine(dp_check.make_model(), DEV, max_batch=8)
ref = dp_check.programme(ref_eng, x, y)
lay = ref_eng.layout
for world in (1, 2, 3):
    group = ThreadGroup(world)
    engines = [DataParallelB200Engine(dp_check.make_model(), DEV, group.member(r), max_batch=8) for r in range(world)]
    outs = [None] * world
    def work(r):
        torch.cuda.set_device(DEV)
        outs[r] = dp_check.programme(engines[r], x, y)
    th = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    [t.start() for t in th]; [t.join() for t in th]
    ok, errs = dp_check.compare(outs[0], ref)
    print("world", world, ok, {k: round(v, 5) for k, v in errs.items()}, "same", all(torch.equal(o["w"], outs[0]["w"]) for o in outs))
    if world > 1:
        for name in lay.param_names:
            e = lay[name]
            a = outs[0]["grad"][e.offset:e.offset + e.numel]; b = ref["grad"][e.offset:e.offset + e.numel]
            print(f"   {name:28s} rel {dp_check.rel(a, b):.5f}  |ref| {float(b.norm()):.4e}")
