"""Which tensor of a training step is not reproducible run to run?  Two identical forward+backward passes at batch B on the same
engine / slab / batch; prints the first differing activation / gradient.   python scripts/check_determinism.py [B] [graphs 0|1]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from neuroimagedisttraining_b200.fl.engine import StepHParams
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = AlexNet3D_Dropout(1)
eng = B200Engine(model, dev, max_batch=B)
eng.use_graphs = (sys.argv[2] == "1") if len(sys.argv) > 2 else True
g = torch.Generator().manual_seed(0)
x = torch.randint(0, 256, (B, 121, 145, 121), generator=g, dtype=torch.uint8).to(dev)
y = (torch.rand(B, generator=g) > 0.5).float().to(dev)
base = eng.new_slab()
hp = StepHParams(lr=0.01, momentum=0.0, wd=5e-4, clip_norm=10.0)


def snap():
    out = {}
    for i, b in enumerate(eng.blocks):
        for name in ("y", "a", "dy", "da", "stats", "scale", "shift", "mean", "invstd", "c12", "bwd_partials"):
            t = getattr(b, name, None)
            if t is not None:
                out[f"block{i}.{name}"] = t.clone()
    out["grad"] = eng.grad.flat.clone()
    out["slab"] = eng.slab.flat.clone()
    return out


runs = []
POISON = os.environ.get("POISON", "0") == "1"
for r in range(3):
    if POISON and r > 0:          # does any kernel read a buffer element no kernel of this step wrote?
        for b in eng.blocks:
            for name in ("y", "a", "dy", "da", "ymax", "wf", "wd"):
                t = getattr(b, name, None)
                if t is not None:
                    t.fill_(float("nan") if r == 1 else 3.0)
            # (stats / c12 / bwd_partials / wg_scratch and the gradient slab are allocated zeroed and some of their entries are never
            # written -- e.g. the identically-zero conv bias gradients in front of a BatchNorm -- so they are not poisoned)
        eng._packed_dirty = True
    s = base.clone()
    eng.attach(s)
    eng.step = 7
    steps = []
    for k in range(2):
        eng.train_step(x, y, hp, None, None, k == 0)
        torch.cuda.synchronize()
        steps.append(snap())
    runs.append(steps)
for r in (1, 2):
    for k in range(2):
        bad = [(n, (runs[r][k][n].float() - runs[0][k][n].float()).abs().max().item()) for n in runs[0][k] if not torch.equal(runs[r][k][n], runs[0][k][n])]
        print(f"run {r} step {k}:", "identical" if not bad else bad[:12])
