"""The reference's experimental 2-D models (all dead code there), kept for completeness:

* ``GroupNorm2d`` / ``GroupNorm3d`` with running statistics (``group_normalization.py:7-53``: group norm expressed as a
  reshaped batch norm) and the ImageNet-style ``resnet*_gn`` family built on them (``resnet_gn.py``);
* ``per_batch_norm`` -- BatchNorm whose affine parameters are supplied per call -- and the CIFAR bottleneck ResNets
  ``resnet29_ip / 56 / 110`` (``resnet_ip.py``);
* ``Meta_net`` (a small hyper-network mapping a mask to weights, imported by most reference trainers but never used) and
  ``cnn_cifar10_meta`` (``cnn_meta.py``);
* CIFAR bottleneck ``resnet20 / resnet29`` with an optional feature output for distillation (``resnet_meta.py``; the
  reference's ``Meta_net(nn.module)`` typo at ``:148`` is not reproduced) and the slimmable ``ResNet20`` whose stage
  widths are chosen per call by scale ids (``resnet_meta_2.py``).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.nn.modules.batchnorm import _BatchNorm


# ------------------------------------------------------------------------ group norm with running stats
def group_norm(input, group, running_mean, running_var, weight=None, bias=None, use_input_stats=True, momentum=0.1, eps=1e-5):
    """Normalise each sample over (channels-in-group x spatial); channels of a group are adjacent.  Running statistics
    (one value per group) are the batch mean of the per-sample group statistics."""
    if not use_input_stats and (running_mean is None or running_var is None):
        raise ValueError("running statistics are required when use_input_stats=False")
    b, c = input.shape[:2]
    g = c // group                      # number of groups (`group` = channels per group, as in the reference)
    x = input.reshape(b, g, group, *input.shape[2:])
    dims = tuple(range(2, x.dim()))
    if use_input_stats:
        mean = x.mean(dim=dims, keepdim=True)
        var = x.var(dim=dims, unbiased=False, keepdim=True)
        if running_mean is not None:
            with torch.no_grad():
                n = x[0, 0].numel()
                running_mean.mul_(1 - momentum).add_(momentum * mean.reshape(b, g).mean(0))
                running_var.mul_(1 - momentum).add_(momentum * (var.reshape(b, g).mean(0) * n / max(n - 1, 1)))
    else:
        shape = (1, g) + (1,) * (x.dim() - 2)
        mean, var = running_mean.view(shape), running_var.view(shape)
    y = (x - mean) / torch.sqrt(var + eps)
    if weight is not None:              # one scale / shift per group, as in the reference (its batch_norm "channels" are the groups)
        shp = (1, g) + (1,) * (x.dim() - 2)
        y = y * weight.view(shp) + bias.view(shp)
    return y.reshape(input.shape)


class _GroupNorm(_BatchNorm):
    def __init__(self, num_features, num_groups=1, eps=1e-5, momentum=0.1, affine=False, track_running_stats=False):
        self.num_groups = num_groups
        self.track_running_stats = track_running_stats
        # `num_groups` is the number of CHANNELS PER GROUP (the reference's naming); statistics and the affine pair are per group,
        # so parameters / buffers have num_features / num_groups entries and checkpoints are interchangeable with the reference's
        super().__init__(int(num_features / num_groups), eps, momentum, affine, track_running_stats)

    def _check_input_dim(self, input):
        return NotImplemented

    def forward(self, input):
        self._check_input_dim(input)
        return group_norm(input, self.num_groups, self.running_mean, self.running_var, self.weight, self.bias,
                          self.training or not self.track_running_stats, self.momentum, self.eps)


class GroupNorm2d(_GroupNorm):
    def _check_input_dim(self, input):
        if input.dim() != 4:
            raise ValueError(f"expected 4D input (got {input.dim()}D input)")


class GroupNorm3d(_GroupNorm):
    def _check_input_dim(self, input):
        if input.dim() != 5:
            raise ValueError(f"expected 5D input (got {input.dim()}D input)")


def norm2d(planes: int, num_channels_per_group: int = 32) -> nn.Module:
    if num_channels_per_group > 0:
        return GroupNorm2d(planes, num_channels_per_group, affine=True, track_running_stats=False)
    return nn.BatchNorm2d(planes)


# ------------------------------------------------------------------------------------- resnet_gn
class _GNBasic(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None, group_norm=0):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = norm2d(planes, group_norm)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = norm2d(planes, group_norm)
        self.downsample = downsample

    def forward(self, x):
        r = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        return self.relu(self.bn2(self.conv2(y)) + r)


class _GNBottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None, group_norm=0):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = norm2d(planes, group_norm)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = norm2d(planes, group_norm)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = norm2d(planes * 4, group_norm)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        r = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.relu(self.bn2(self.conv2(y)))
        return self.relu(self.bn3(self.conv3(y)) + r)


class ResNetGN(nn.Module):
    def __init__(self, block, layers: Sequence[int], num_classes=1000, group_norm=0):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = norm2d(64, group_norm)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(3, 2, 1)
        self.layer1 = self._make(block, 64, layers[0], 1, group_norm)
        self.layer2 = self._make(block, 128, layers[1], 2, group_norm)
        self.layer3 = self._make(block, 256, layers[2], 2, group_norm)
        self.layer4 = self._make(block, 512, layers[3], 2, group_norm)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(512 * block.expansion, num_classes)

    def _make(self, block, planes, n, stride, gn):
        ds = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            ds = nn.Sequential(nn.Conv2d(self.inplanes, planes * block.expansion, 1, stride, bias=False), norm2d(planes * block.expansion, gn))
        blocks = [block(self.inplanes, planes, stride, ds, gn)]
        self.inplanes = planes * block.expansion
        blocks += [block(self.inplanes, planes, group_norm=gn) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        return self.fc(torch.flatten(self.avgpool(x), 1))


def resnet18_gn(num_classes=1000, group_norm=2):
    return ResNetGN(_GNBasic, [2, 2, 2, 2], num_classes, group_norm)


def resnet34_gn(num_classes=1000, group_norm=2):
    return ResNetGN(_GNBasic, [3, 4, 6, 3], num_classes, group_norm)


def resnet50_gn(num_classes=1000, group_norm=2):
    return ResNetGN(_GNBottleneck, [3, 4, 6, 3], num_classes, group_norm)


# ------------------------------------------------------------------------------------- resnet_ip
class per_batch_norm(nn.BatchNorm2d):
    """BatchNorm2d whose affine weight / bias are given per call (personalised normalisation layers)."""

    def forward(self, input, weight=None, bias=None):
        self._check_input_dim(input)
        w = self.weight if weight is None else weight
        b = self.bias if bias is None else bias
        factor = 0.0 if self.momentum is None else self.momentum
        if self.training and self.track_running_stats and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
            factor = 1.0 / float(self.num_batches_tracked) if self.momentum is None else self.momentum
        bn_training = self.training or (self.running_mean is None and self.running_var is None)
        return F.batch_norm(input, self.running_mean if not self.training or self.track_running_stats else None,
                            self.running_var if not self.training or self.track_running_stats else None, w, b, bn_training, factor, self.eps)


class _CifarBottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None, norm=nn.BatchNorm2d):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 1, bias=False)
        self.bn1 = norm(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, stride, 1, bias=False)
        self.bn2 = norm(planes)
        self.conv3 = nn.Conv2d(planes, planes * 4, 1, bias=False)
        self.bn3 = norm(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        r = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.relu(self.bn2(self.conv2(y)))
        return self.relu(self.bn3(self.conv3(y)) + r)


class CifarBottleneckResNet(nn.Module):
    """16-32-64 CIFAR ResNet with bottleneck blocks; ``KD=True`` also returns the pooled features."""

    def __init__(self, layers: Sequence[int], num_classes=10, norm=nn.BatchNorm2d, KD=False):
        super().__init__()
        self.inplanes = 16
        self.KD = KD
        self.conv1 = nn.Conv2d(3, 16, 3, 1, 1, bias=False)
        self.bn1 = norm(16)
        self.relu = nn.ReLU(inplace=True)
        self.layer1 = self._make(16, layers[0], 1, norm)
        self.layer2 = self._make(32, layers[1], 2, norm)
        self.layer3 = self._make(64, layers[2], 2, norm)
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc = nn.Linear(64 * 4, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")

    def _make(self, planes, n, stride, norm):
        ds = None
        if stride != 1 or self.inplanes != planes * 4:
            ds = nn.Sequential(nn.Conv2d(self.inplanes, planes * 4, 1, stride, bias=False), norm(planes * 4))
        blocks = [_CifarBottleneck(self.inplanes, planes, stride, ds, norm)]
        self.inplanes = planes * 4
        blocks += [_CifarBottleneck(self.inplanes, planes, norm=norm) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = self.relu(self.bn1(self.conv1(x)))
        x = self.layer3(self.layer2(self.layer1(x)))
        f = torch.flatten(self.avgpool(x), 1)
        out = self.fc(f)
        return (f, out) if self.KD else out


def resnet20(class_num=10, **kw):
    return CifarBottleneckResNet([2, 2, 2], class_num, **kw)


def resnet29(class_num=10, **kw):
    return CifarBottleneckResNet([3, 3, 3], class_num, **kw)


def resnet29_ip(class_num=10, **kw):
    return CifarBottleneckResNet([3, 3, 3], class_num, norm=per_batch_norm, **kw)


def resnet56_ip(class_num=10, **kw):
    return CifarBottleneckResNet([6, 6, 6], class_num, norm=per_batch_norm, **kw)


def resnet110_ip(class_num=10, **kw):
    return CifarBottleneckResNet([12, 12, 12], class_num, norm=per_batch_norm, **kw)


# ------------------------------------------------------------------------------------- cnn_meta
class Meta_net(nn.Module):
    """mask (any shape) -> tensor of the same shape through a 50-50 MLP."""

    def __init__(self, mask: torch.Tensor):
        super().__init__()
        size = int(mask.numel())
        self.fc11 = nn.Linear(size, 50)
        self.fc12 = nn.Linear(50, 50)
        self.fc13 = nn.Linear(50, size)
        for m in self.modules():
            if isinstance(m, nn.Linear):
                nn.init.kaiming_uniform_(m.weight)
                nn.init.zeros_(m.bias)

    def forward(self, input):
        x = F.relu(self.fc11(input.flatten()))
        x = F.relu(self.fc12(x))
        return self.fc13(x).view(input.shape)


class cnn_cifar10_meta(nn.Module):
    """The CIFAR CNN with per-conv binary masks (random, ``dense_ratio`` kept) and optional hyper-networks that re-weight
    the masked kernels."""

    def __init__(self, dense_ratio: float = 0.2, used_meta: bool = False, n_cls: int = 10):
        super().__init__()
        self.dense_ratio, self.used_meta = dense_ratio, used_meta
        self.conv1 = nn.Conv2d(3, 64, 5)
        self.conv2 = nn.Conv2d(64, 64, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.fc1 = nn.Linear(64 * 5 * 5, 384)
        self.fc2 = nn.Linear(384, 192)
        self.fc3 = nn.Linear(192, n_cls)
        self.init_masks()
        if used_meta:
            self.meta1, self.meta2 = Meta_net(self.mask1[0, 0]), Meta_net(self.mask2[0, 0])

    def init_conv_masks(self, size, dense_ratio):
        m = torch.zeros(size)
        k = int(dense_ratio * m.numel())
        m.view(-1)[torch.randperm(m.numel())[:k]] = 1
        return m

    def init_masks(self):
        self.register_buffer("mask1", self.init_conv_masks(self.conv1.weight.shape, self.dense_ratio))
        self.register_buffer("mask2", self.init_conv_masks(self.conv2.weight.shape, self.dense_ratio))

    def get_masks(self):
        return {"conv1.weight": self.mask1, "conv2.weight": self.mask2}

    def forward(self, x):
        w1, w2 = self.conv1.weight * self.mask1, self.conv2.weight * self.mask2
        if self.used_meta:
            w1 = w1 * (1 + self.meta1(self.mask1[0, 0]))
            w2 = w2 * (1 + self.meta2(self.mask2[0, 0]))
        x = self.pool(F.relu(F.conv2d(x, w1, self.conv1.bias)))
        x = self.pool(F.relu(F.conv2d(x, w2, self.conv2.bias)))
        x = F.relu(self.fc1(x.flatten(1)))
        return self.fc3(F.relu(self.fc2(x)))


# ------------------------------------------------------------------------ slimmable ResNet20 (resnet_meta_2)
CHANNEL_SCALES = [0.25, 0.5, 0.75, 1.0]


class _SlimBlock(nn.Module):
    def __init__(self, inplanes, planes, stride):
        super().__init__()
        self.conv1 = nn.Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn1 = nn.ModuleList([nn.BatchNorm2d(int(planes * s)) for s in CHANNEL_SCALES])
        self.bn2 = nn.ModuleList([nn.BatchNorm2d(planes) for _ in CHANNEL_SCALES])
        self.short = None if (stride == 1 and inplanes == planes) else nn.Conv2d(inplanes, planes, 1, stride, bias=False)

    def forward(self, x, mid_scale_id: int):
        mid = int(self.conv1.out_channels * CHANNEL_SCALES[mid_scale_id])
        y = F.conv2d(x, self.conv1.weight[:mid], None, self.conv1.stride, 1)
        y = F.relu(self.bn1[mid_scale_id](y))
        y = self.bn2[mid_scale_id](F.conv2d(y, self.conv2.weight[:, :mid], None, 1, 1))
        r = x if self.short is None else self.short(x)
        return F.relu(y + r)


class ResNet20(nn.Module):
    """ResNet-20 whose blocks run at a per-call width: ``forward(x, mid_scale_ids)`` with one scale id (0..3) per block."""

    def __init__(self, num_classes=10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 16, 3, 1, 1, bias=False)
        self.bn1 = nn.BatchNorm2d(16)
        cfg = [(16, 16, 1)] * 3 + [(16, 32, 2)] + [(32, 32, 1)] * 2 + [(32, 64, 2)] + [(64, 64, 1)] * 2
        self.blocks = nn.ModuleList([_SlimBlock(i, o, s) for i, o, s in cfg])
        self.fc = nn.Linear(64, num_classes)

    def forward(self, x, mid_scale_ids: Optional[Sequence[int]] = None):
        ids = list(mid_scale_ids) if mid_scale_ids is not None else [len(CHANNEL_SCALES) - 1] * len(self.blocks)
        x = F.relu(self.bn1(self.conv1(x)))
        for b, i in zip(self.blocks, ids):
            x = b(x, int(i))
        return self.fc(F.adaptive_avg_pool2d(x, 1).flatten(1))
