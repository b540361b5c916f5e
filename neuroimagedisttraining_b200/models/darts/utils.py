"""Helpers of the DARTS scripts (reference: ``fedml_api/model/cv/darts/utils.py:9-100``): running averages, top-k accuracy, Cutout,
parameter counting, checkpoints, drop-path, experiment directories.  ``drop_path`` works on any device (the reference's allocates
a ``torch.cuda.FloatTensor`` and mutates its input)."""
import os
import shutil

import numpy as np
import torch


class AvgrageMeter:          # (sic) the reference's spelling, kept so scripts port over unchanged
    def __init__(self):
        self.reset()

    def reset(self):
        self.avg, self.sum, self.cnt = 0.0, 0.0, 0

    def update(self, val, n=1):
        self.sum += float(val) * n
        self.cnt += n
        self.avg = self.sum / self.cnt


AverageMeter = AvgrageMeter


def accuracy(output, target, topk=(1,)):
    """Top-k accuracies in percent."""
    maxk = max(topk)
    pred = output.topk(maxk, 1, True, True)[1].t()
    correct = pred.eq(target.view(1, -1).expand_as(pred))
    return [correct[:k].reshape(-1).float().sum() * (100.0 / target.size(0)) for k in topk]


class Cutout:
    """Zeroes one ``length`` x ``length`` square (clipped at the border) of a CHW image tensor; draws from numpy's global stream."""

    def __init__(self, length):
        self.length = length

    def __call__(self, img):
        h, w = img.size(1), img.size(2)
        y, x = np.random.randint(h), np.random.randint(w)
        y1, y2 = np.clip([y - self.length // 2, y + self.length // 2], 0, h)
        x1, x2 = np.clip([x - self.length // 2, x + self.length // 2], 0, w)
        mask = torch.ones(h, w, dtype=img.dtype)
        mask[y1:y2, x1:x2] = 0
        return img * mask.expand_as(img)


def count_parameters_in_MB(model):
    return sum(v.numel() for name, v in model.named_parameters() if "auxiliary" not in name) / 1e6


def save_checkpoint(state, is_best, save):
    filename = os.path.join(save, "checkpoint.pth.tar")
    torch.save(state, filename)
    if is_best:
        shutil.copyfile(filename, os.path.join(save, "model_best.pth.tar"))


def save(model, model_path):
    torch.save(model.state_dict(), model_path)


def load(model, model_path):
    model.load_state_dict(torch.load(model_path, map_location="cpu"))


def drop_path(x, drop_prob):
    if drop_prob > 0.0:
        keep = 1.0 - drop_prob
        mask = torch.bernoulli(torch.full((x.size(0),) + (1,) * (x.dim() - 1), keep, device=x.device, dtype=x.dtype))
        x = x / keep * mask
    return x


def create_exp_dir(path, scripts_to_save=None):
    os.makedirs(path, exist_ok=True)
    if scripts_to_save:
        os.makedirs(os.path.join(path, "scripts"), exist_ok=True)
        for script in scripts_to_save:
            shutil.copyfile(script, os.path.join(path, "scripts", os.path.basename(script)))
