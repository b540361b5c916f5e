"""Synchronised BatchNorm over ranks.

The reference's only collective call site is a (dead) SyncBN for ``nn.DataParallel``: per-replica sum and sum-of-squares are
reduce-added to a master thread, mean / inv-std are broadcast back through Python queues
(``fedml_api/model/cv/batchnorm_utils.py:150-237``, ``SyncMaster`` ``:58-133``, ``DataParallelWithCallback`` ``:422-437``).
With one process per GPU that machinery collapses to two small all-reduces: ``[sum(x), sum(x^2), count]`` per channel in the
forward pass and ``[sum(dy), sum(dy*xhat)]`` in the backward pass (2*C+1 and 2*C floats).  Used when ONE client's batch is
split across ranks (data parallel inside a client); client-parallel FL does not need it.
"""
from __future__ import annotations

import torch
import torch.distributed as dist
import torch.nn as nn


def _reduce(t: torch.Tensor, group=None) -> torch.Tensor:
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t


class _SyncBNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps, group):
        C = x.shape[1]
        dims = [0] + list(range(2, x.dim()))
        stat = torch.cat([x.sum(dims), (x * x).sum(dims), x.new_tensor([x.numel() / C])])
        _reduce(stat, group)
        n = stat[-1]
        mean = stat[:C] / n
        var = (stat[C:2 * C] / n - mean * mean).clamp_(min=0)
        invstd = torch.rsqrt(var + eps)
        shp = [1, C] + [1] * (x.dim() - 2)
        xhat = (x - mean.view(shp)) * invstd.view(shp)
        ctx.save_for_backward(xhat, weight, invstd)
        ctx.group, ctx.n = group, n
        return xhat * weight.view(shp) + bias.view(shp), mean, var * n / (n - 1).clamp(min=1)

    @staticmethod
    def backward(ctx, dy, _dm, _dv):
        xhat, weight, invstd = ctx.saved_tensors
        C = dy.shape[1]
        dims = [0] + list(range(2, dy.dim()))
        shp = [1, C] + [1] * (dy.dim() - 2)
        local = torch.cat([dy.sum(dims), (dy * xhat).sum(dims)])
        dbeta_l, dgamma_l = local[:C].clone(), local[C:].clone()
        _reduce(local, ctx.group)
        c1, c2 = local[:C] / ctx.n, local[C:] / ctx.n
        dx = (weight * invstd).view(shp) * (dy - c1.view(shp) - xhat * c2.view(shp))
        return dx, dgamma_l, dbeta_l, None, None


class SynchronizedBatchNorm(nn.modules.batchnorm._BatchNorm):
    """Drop-in BatchNorm{1,2,3}d whose batch statistics span all ranks of ``process_group``."""

    def __init__(self, num_features, eps=1e-5, momentum=0.1, affine=True, process_group=None):
        super().__init__(num_features, eps, momentum, affine, True)
        self.process_group = process_group

    def _check_input_dim(self, input):
        if input.dim() < 2:
            raise ValueError("expected at least 2D input")

    def forward(self, x):
        if not self.training:
            return nn.functional.batch_norm(x, self.running_mean, self.running_var, self.weight, self.bias, False, 0.0, self.eps)
        C = x.shape[1]
        weight = self.weight if self.affine else torch.ones(C, dtype=x.dtype, device=x.device)
        bias = self.bias if self.affine else torch.zeros(C, dtype=x.dtype, device=x.device)
        y, mean, var_unbiased = _SyncBNFn.apply(x, weight, bias, self.eps, self.process_group)
        with torch.no_grad():
            self.num_batches_tracked += 1
            # momentum=None: cumulative moving average, as in torch.nn.BatchNorm
            m = self.momentum if self.momentum is not None else 1.0 / float(self.num_batches_tracked)
            self.running_mean.mul_(1 - m).add_(m * mean)
            self.running_var.mul_(1 - m).add_(m * var_unbiased)
        return y


SynchronizedBatchNorm1d = SynchronizedBatchNorm2d = SynchronizedBatchNorm3d = SynchronizedBatchNorm


def convert_sync_batchnorm(module: nn.Module, process_group=None) -> nn.Module:
    """Replace every BatchNorm in ``module`` by :class:`SynchronizedBatchNorm` (parameters and statistics are kept)."""
    out = module
    if isinstance(module, nn.modules.batchnorm._BatchNorm) and not isinstance(module, SynchronizedBatchNorm):
        out = SynchronizedBatchNorm(module.num_features, module.eps, module.momentum, module.affine, process_group)
        if module.affine:
            out.weight, out.bias = module.weight, module.bias
        out.running_mean, out.running_var, out.num_batches_tracked = module.running_mean, module.running_var, module.num_batches_tracked
    for name, child in module.named_children():
        out.add_module(name, convert_sync_batchnorm(child, process_group))
    return out
