"""Architecture search loop (``darts/train_search.py:60-315`` of the reference, without wandb / DataParallel): alternating
architecture step on validation data and weight step on training data; ``arch="gdas"`` searches with the Gumbel-softmax network and
anneals its temperature from ``tau_max`` to ``tau_min`` over the epochs."""
import torch

from .architect import Architect
from .model_search import Network
from .model_search_gdas import NetworkGumbelSoftmax


def search(train_loader, valid_loader, num_classes=10, C=16, layers=8, epochs=1, lr=0.025, unrolled=False, device="cpu", max_batches=None,
           arch="darts", tau_max=10.0, tau_min=1.0, logger=None):
    """Returns (model, genotype)."""
    if arch == "gdas":
        model = NetworkGumbelSoftmax(C, num_classes, layers, tau=tau_max).to(device)
    elif arch == "darts":
        model = Network(C, num_classes, layers).to(device)
    else:
        raise ValueError(f"unknown search space relaxation {arch!r} (darts | gdas)")
    opt = torch.optim.SGD(model.weight_parameters(), lr, momentum=0.9, weight_decay=3e-4)
    sched = torch.optim.lr_scheduler.CosineAnnealingLR(opt, float(max(epochs, 1)), eta_min=1e-3)
    architect = Architect(model)
    for epoch in range(epochs):
        if arch == "gdas":
            model.set_tau(tau_max - (tau_max - tau_min) * epoch / max(epochs - 1, 1))
        model.train()
        for b, ((x, y), (xv, yv)) in enumerate(zip(train_loader, valid_loader)):
            if max_batches is not None and b >= max_batches:
                break
            x, y, xv, yv = x.to(device), y.to(device).long(), xv.to(device), yv.to(device).long()
            architect.step(x, y, xv, yv, sched.get_last_lr()[0], unrolled)
            opt.zero_grad()
            loss = model._loss(x, y)
            loss.backward()
            torch.nn.utils.clip_grad_norm_(model.weight_parameters(), 5.0)
            opt.step()
        sched.step()
        if logger is not None:
            logger.info({"epoch": epoch, "genotype": model.genotype(), "tau": getattr(model, "tau", None)})
    return model, model.genotype()
