"""Training / evaluation of a discrete (genotype) network -- the reference's ``darts/train.py:60-214`` without wandb / DataParallel:
SGD + momentum, cosine learning-rate schedule, gradient clipping, auxiliary-head loss, linearly increasing drop-path."""
import torch
import torch.nn as nn

from .model import NetworkCIFAR
from .utils import AvgrageMeter, accuracy


def train_epoch(model, loader, optimizer, criterion=None, device="cpu", auxiliary_weight=0.4, grad_clip=5.0, max_batches=None):
    criterion = criterion or nn.CrossEntropyLoss()
    objs, top1, top5 = AvgrageMeter(), AvgrageMeter(), AvgrageMeter()
    model.train()
    for b, (x, y) in enumerate(loader):
        if max_batches is not None and b >= max_batches:
            break
        x, y = x.to(device), y.to(device).long()
        optimizer.zero_grad()
        out = model(x)
        logits, aux = out if isinstance(out, tuple) else (out, None)
        loss = criterion(logits, y)
        if aux is not None:
            loss = loss + auxiliary_weight * criterion(aux, y)
        loss.backward()
        nn.utils.clip_grad_norm_(model.parameters(), grad_clip)
        optimizer.step()
        k = min(5, logits.shape[1])
        p1, p5 = accuracy(logits.detach(), y, topk=(1, k))
        objs.update(loss.item(), x.size(0)); top1.update(p1.item(), x.size(0)); top5.update(p5.item(), x.size(0))
    return top1.avg, top5.avg, objs.avg


@torch.no_grad()
def infer(model, loader, criterion=None, device="cpu", max_batches=None):
    criterion = criterion or nn.CrossEntropyLoss()
    objs, top1, top5 = AvgrageMeter(), AvgrageMeter(), AvgrageMeter()
    model.eval()
    for b, (x, y) in enumerate(loader):
        if max_batches is not None and b >= max_batches:
            break
        x, y = x.to(device), y.to(device).long()
        logits = model(x)
        k = min(5, logits.shape[1])
        p1, p5 = accuracy(logits, y, topk=(1, k))
        objs.update(criterion(logits, y).item(), x.size(0)); top1.update(p1.item(), x.size(0)); top5.update(p5.item(), x.size(0))
    return top1.avg, top5.avg, objs.avg


def train(train_loader, valid_loader, genotype=None, num_classes=10, C=36, layers=20, auxiliary=True, epochs=600, lr=0.025, momentum=0.9,
          weight_decay=3e-4, drop_path_prob=0.2, auxiliary_weight=0.4, grad_clip=5.0, device="cpu", max_batches=None, logger=None):
    """Returns (model, history) with one ``{"epoch", "train_acc", "valid_acc", "train_loss", "valid_loss", "lr"}`` per epoch."""
    model = NetworkCIFAR(C, num_classes, layers, auxiliary, genotype).to(device)
    opt = torch.optim.SGD(model.parameters(), lr, momentum=momentum, weight_decay=weight_decay)
    sched = torch.optim.lr_scheduler.CosineAnnealingLR(opt, float(max(epochs, 1)))
    history = []
    for epoch in range(epochs):
        model.drop_path_prob = drop_path_prob * epoch / max(epochs, 1)
        tr1, _, trl = train_epoch(model, train_loader, opt, None, device, auxiliary_weight, grad_clip, max_batches)
        va1, _, val = infer(model, valid_loader, None, device, max_batches)
        history.append({"epoch": epoch, "train_acc": tr1, "valid_acc": va1, "train_loss": trl, "valid_loss": val, "lr": sched.get_last_lr()[0]})
        if logger is not None:
            logger.info(history[-1])
        sched.step()
    return model, history
