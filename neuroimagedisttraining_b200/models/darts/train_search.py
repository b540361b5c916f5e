import torch

from .architect import Architect
from .model_search import Network


def search(train_loader, valid_loader, num_classes=10, C=16, layers=8, epochs=1, lr=0.025, unrolled=False, device="cpu", max_batches=None):
    """Alternating weight / architecture optimisation (``darts/train_search.py`` of the reference); returns (model, genotype)."""
    model = Network(C, num_classes, layers).to(device)
    opt = torch.optim.SGD(model.weight_parameters(), lr, momentum=0.9, weight_decay=3e-4)
    sched = torch.optim.lr_scheduler.CosineAnnealingLR(opt, float(max(epochs, 1)), eta_min=1e-3)
    arch = Architect(model)
    for _ in range(epochs):
        model.train()
        for b, ((x, y), (xv, yv)) in enumerate(zip(train_loader, valid_loader)):
            if max_batches is not None and b >= max_batches:
                break
            x, y, xv, yv = x.to(device), y.to(device).long(), xv.to(device), yv.to(device).long()
            arch.step(x, y, xv, yv, sched.get_last_lr()[0], unrolled)
            opt.zero_grad()
            loss = model._loss(x, y)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.weight_parameters(), 5.0)
            opt.step()
        sched.step()
    return model, model.genotype()
