"""Per-layer forward/backward error of the B200 engine vs PyTorch fp32 (diagnostic, GPU)."""
import copy, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
from neuroimagedisttraining_b200.fl.engine import TorchEngine
from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
from neuroimagedisttraining_b200.ops.engine_b200 import B200Engine

torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
DEV = "cuda"
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
torch.manual_seed(5)
model = AlexNet3D_Dropout(1, p_drop=0.0)
tmodel = copy.deepcopy(model)
eng = B200Engine(model, DEV, max_batch=B)
ref = TorchEngine(tmodel, DEV, volumetric=True, input_scale=1 / 255.0)
slab = eng.new_slab(); slab2 = slab.clone()
eng.attach(slab); ref.attach(slab2)
g = torch.Generator().manual_seed(0)
x = torch.randint(0, 256, (B, 121, 145, 121), dtype=torch.uint8, generator=g).to(DEV)
y = (torch.arange(B) % 2).float().to(DEV)

# torch forward with hooks capturing block outputs and their grads
acts = {}
def hook(name):
    def f(m, i, o):
        o.retain_grad(); acts[name] = o
    return f
feats = ref.model.features
for i, b in enumerate(eng.blocks):
    conv_idx = int(b.spec.conv.split(".")[1])
    feats[conv_idx].register_forward_hook(hook(f"y{i}"))
    last = conv_idx + (3 if b.spec.pool_k else 2)
    feats[last].register_forward_hook(hook(f"a{i}"))
ref.model.train()
ref.grad.flat.zero_()
for name, p in ref.model.named_parameters():
    p.grad = ref.grad.view(name)
out = ref.model(ref._prep(x))
loss = F.binary_cross_entropy_with_logits(out.flatten(), y)
loss.backward()
eng.compute_grads(x, y, True)
torch.cuda.synchronize()
print("loss", eng.loss_acc.item(), loss.item())
def cmp(tag, a, b):
    a, b = a.float().flatten(), b.float().flatten()
    rel = (a - b).norm().item() / (b.norm().item() + 1e-30)
    cos = F.cosine_similarity(a, b, dim=0).item()
    print(f"{tag:28s} relL2 {rel:9.3e} cos {cos:.5f}  |ref| {b.norm().item():.3e}")
for i, b in enumerate(eng.blocks):
    Bn = x.shape[0]
    cmp(f"y{i} (conv out)", b.y[:Bn].permute(0, 4, 1, 2, 3), acts[f"y{i}"])
    cmp(f"a{i} (block out)", b.a[:Bn].permute(0, 4, 1, 2, 3), acts[f"a{i}"])
for i, b in reversed(list(enumerate(eng.blocks))):
    cmp(f"da{i}", b.da[:B].permute(0, 4, 1, 2, 3), acts[f"a{i}"].grad)
    cmp(f"dy{i}", b.dy[:B].permute(0, 4, 1, 2, 3), acts[f"y{i}"].grad)
for name in eng.layout.param_names:
    cmp(name, eng.grad.view(name), ref.grad.view(name))

print("---- isolated per-block checks (torch fp32 on the engine's own bf16 tensors)")
lay = eng.layout
for i, b in enumerate(eng.blocks):
    s = b.spec
    yb = b.y[:B].permute(0, 4, 1, 2, 3).float().contiguous().requires_grad_(True)
    gam = slab.view(s.bn + ".weight").clone().requires_grad_(True)
    bet = slab.view(s.bn + ".bias").clone().requires_grad_(True)
    o = F.relu(F.batch_norm(yb, None, None, gam, bet, training=True, eps=1e-5))
    if s.pool_k:
        o = F.max_pool3d(o, s.pool_k, s.pool_s)
    o.backward(b.da[:B].permute(0, 4, 1, 2, 3).float())
    cmp(f"[iso] dy{i}", b.dy[:B].permute(0, 4, 1, 2, 3), yb.grad)
    cmp(f"[iso] dgamma{i}", eng.grad.view(s.bn + ".weight"), gam.grad)
    cmp(f"[iso] dbeta{i}", eng.grad.view(s.bn + ".bias"), bet.grad)
    w = slab.view(s.conv + ".weight").clone()
    dyf = b.dy[:B].permute(0, 4, 1, 2, 3).float().contiguous()
    if i == 0:
        xin = (x.float() / 255.0).bfloat16().float().unsqueeze(1)
    else:
        xin = eng.blocks[i - 1].a[:B].permute(0, 4, 1, 2, 3).float().contiguous()
    xin.requires_grad_(i > 0)
    wq = w.bfloat16().float().requires_grad_(True)
    oc = F.conv3d(xin, wq, None, stride=s.stride, padding=s.pad)
    oc.backward(dyf)
    cmp(f"[iso] dW{i}", eng.grad.view(s.conv + ".weight"), wq.grad)
    if i > 0:
        cmp(f"[iso] da{i-1}", eng.blocks[i - 1].da[:B].permute(0, 4, 1, 2, 3), xin.grad)
