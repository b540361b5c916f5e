"""The reference's experimental 2-D models (all dead code there), kept for completeness:

* ``GroupNorm2d`` / ``GroupNorm3d`` with running statistics (``group_normalization.py:7-53``: group norm expressed as a
  reshaped batch norm) and the ImageNet-style ``resnet*_gn`` family built on them (``resnet_gn.py``);
* ``per_batch_norm`` -- BatchNorm whose affine parameters are supplied per call -- and the CIFAR bottleneck ResNets
  ``resnet29_ip / 56 / 110`` whose every tensor is a shared ``*_g`` + personal ``*_v`` pair (``resnet_ip.py``);
* ``Meta_net`` (a small hyper-network mapping a mask to weights, imported by most reference trainers but never used) and
  ``cnn_cifar10_meta`` (``cnn_meta.py``);
* CIFAR bottleneck ``resnet20 / resnet29`` with an optional feature output for distillation (``resnet_meta.py``; the
  reference's ``Meta_net(nn.module)`` typo at ``:148`` is not reproduced), the slimmable ``ResNet20`` whose block widths are
  chosen per call by scale ids, and ``MetaPruningResNet20`` -- the working form of ``resnet_meta_2.py``: hyper-networks emit every
  convolution weight from the requested width ratios (31 scales), one statistics-only BatchNorm per width.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

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


# "ip" networks (``resnet_ip.py:98-300``): every learnable tensor exists twice, a shared part ``*_g`` (what federated averaging
# exchanges) and a personal variation ``*_v`` (kept by the client); the layer runs with their SUM.  Normalisation layers keep no
# running statistics (``track_running_stats=False``) and take the summed affine pair per call.
def _pair_conv(g: nn.Conv2d, v: nn.Conv2d, x):
    return F.conv2d(x, g.weight + v.weight, None, g.stride, g.padding, g.dilation, g.groups)


def _pair_bn(g: per_batch_norm, v: per_batch_norm, x):
    return g(x, g.weight + v.weight, g.bias + v.bias)


def _ip_pair(mod: nn.Module, tag: str, cin: int, cout: int, k: int, stride: int = 1, bn_first: bool = False):
    """Registers ``conv<tag>_g/_v`` and ``bn<tag>_g/_v`` on ``mod`` in the reference's order (checkpoints stay interchangeable)."""
    mk_conv = lambda: nn.Conv2d(cin, cout, k, stride, k // 2, bias=False)
    mk_bn = lambda: per_batch_norm(cout, track_running_stats=False)
    order = ([("conv", "g"), ("bn", "g"), ("conv", "v"), ("bn", "v")] if bn_first else [("conv", "g"), ("conv", "v"), ("bn", "g"), ("bn", "v")])
    for kind, part in order:
        mod.add_module(f"{kind}{tag}_{part}", mk_conv() if kind == "conv" else mk_bn())


class _IPBottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        _ip_pair(self, "1", inplanes, planes, 1)
        _ip_pair(self, "2", planes, planes, 3, stride)
        _ip_pair(self, "3", planes, planes * 4, 1)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample

    def forward(self, x):
        r = x
        y = self.relu(_pair_bn(self.bn1_g, self.bn1_v, _pair_conv(self.conv1_g, self.conv1_v, x)))
        y = self.relu(_pair_bn(self.bn2_g, self.bn2_v, _pair_conv(self.conv2_g, self.conv2_v, y)))
        y = _pair_bn(self.bn3_g, self.bn3_v, _pair_conv(self.conv3_g, self.conv3_v, y))
        if self.downsample is not None:
            d = self.downsample
            r = _pair_bn(d.bn1_g, d.bn1_v, _pair_conv(d.conv1_g, d.conv1_v, x))
        return self.relu(y + r)


class ResNet_ip(nn.Module):
    """16-32-64 CIFAR bottleneck ResNet with shared + personal parameter pairs; ``KD=True`` also returns the pooled features.
    ``shared_parameters()`` / ``personal_parameters()`` split the two halves for aggregation."""

    def __init__(self, layers: Sequence[int], num_classes=10, KD=False, zero_init_residual=False):
        super().__init__()
        self.inplanes, self.KD = 16, KD
        _ip_pair(self, "1", 3, 16, 3)
        self.relu = nn.ReLU(inplace=True)
        self.layer1 = self._make(16, layers[0], 1)
        self.layer2 = self._make(32, layers[1], 2)
        self.layer3 = self._make(64, layers[2], 2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc_g = nn.Linear(64 * 4, num_classes)
        self.fc_v = nn.Linear(64 * 4, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, per_batch_norm):        # both halves start at 1: the effective scale is 2, as in the reference
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                nn.init.zeros_(m.weight)
                nn.init.zeros_(m.bias)
        if zero_init_residual:
            for m in self.modules():
                if isinstance(m, _IPBottleneck):
                    nn.init.zeros_(m.bn3_g.weight)
                    nn.init.zeros_(m.bn3_v.weight)

    def _make(self, planes, n, stride):
        ds = None
        if stride != 1 or self.inplanes != planes * 4:
            ds = nn.Sequential()
            _ip_pair(ds, "1", self.inplanes, planes * 4, 1, stride, bn_first=True)
        blocks = [_IPBottleneck(self.inplanes, planes, stride, ds)]
        self.inplanes = planes * 4
        blocks += [_IPBottleneck(self.inplanes, planes) for _ in range(1, n)]
        return nn.Sequential(*blocks)

    def shared_parameters(self):
        return [p for n, p in self.named_parameters() if n.rsplit(".", 1)[0].endswith("_g")]

    def personal_parameters(self):
        return [p for n, p in self.named_parameters() if n.rsplit(".", 1)[0].endswith("_v")]

    def forward(self, x):
        x = self.relu(_pair_bn(self.bn1_g, self.bn1_v, _pair_conv(self.conv1_g, self.conv1_v, x)))
        x = self.layer3(self.layer2(self.layer1(x)))
        f = torch.flatten(self.avgpool(x), 1)
        out = F.linear(f, self.fc_g.weight + self.fc_v.weight, self.fc_g.bias + self.fc_v.bias)
        return (f, out) if self.KD else out


def resnet29_ip(class_num=10, **kw):
    return ResNet_ip([3, 3, 3], class_num, **kw)


def resnet56_ip(class_num=10, **kw):
    return ResNet_ip([6, 6, 6], class_num, **kw)


def resnet110_ip(class_num=10, **kw):
    return ResNet_ip([12, 12, 12], class_num, **kw)


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
    """Bias-free CIFAR CNN whose three learnable layers are named ``meta_*`` (``cnn_meta.py:12-70``): conv 3->64 k5, conv 64->64 k5,
    3x3 / stride-2 max-pools, linear 1024->10.  ``init_masks()`` draws one random binary mask per ``meta_*`` weight keeping
    ``dense_ratio`` of the entries (the reference hard-codes 0.2 there); ``set_masks`` makes the forward pass use ``w * mask``,
    and with ``used_meta`` a :class:`Meta_net` per layer turns the mask into a multiplicative re-weighting of the kept weights."""

    def __init__(self, dense_ratio: float = 0.2, used_meta: bool = False):
        super().__init__()
        self.dense_ratio, self.used_meta = dense_ratio, used_meta
        self.meta_conv1 = nn.Conv2d(3, 64, 5, bias=False)
        self.meta_conv2 = nn.Conv2d(64, 64, 5, bias=False)
        self.pool = nn.MaxPool2d(3, 2)
        self.meta_fc1 = nn.Linear(64 * 4 * 4, 10, bias=False)
        self.raw_conv_masks: Optional[Dict[str, torch.Tensor]] = None
        self.hyper: Optional[nn.ModuleDict] = None        # created by set_masks when used_meta (adds keys to the state_dict)

    def init_conv_masks(self, size, dense_ratio):
        m = torch.zeros(size)
        k = int(dense_ratio * m.numel())
        if k > 0:
            m.view(-1)[torch.randperm(m.numel())[:k]] = 1
        return m

    def init_masks(self):
        return {name + ".weight": self.init_conv_masks(mod.weight.shape, self.dense_ratio)
                for name, mod in self.named_modules() if "meta" in name and hasattr(mod, "weight")}

    def set_masks(self, masks: Optional[Dict[str, torch.Tensor]]):
        self.raw_conv_masks = masks
        if masks is not None and self.used_meta and self.hyper is None:
            # hyper-networks work on one kernel slice (conv) / one row (linear) so their size stays small
            self.hyper = nn.ModuleDict({k.replace(".", "_"): Meta_net(m[0, 0] if m.dim() == 4 else m[0]) for k, m in masks.items()})

    def get_masks(self):
        return self.raw_conv_masks

    def _weight(self, name: str, w: torch.Tensor) -> torch.Tensor:
        if self.raw_conv_masks is None:
            return w
        m = self.raw_conv_masks[name + ".weight"].to(w.device)
        w = w * m
        if self.used_meta and self.hyper is not None:
            probe = m[0, 0] if m.dim() == 4 else m[0]
            w = w * (1 + self.hyper[(name + ".weight").replace(".", "_")](probe))
        return w

    def forward(self, x):
        x = self.pool(F.relu(F.conv2d(x, self._weight("meta_conv1", self.meta_conv1.weight))))
        x = self.pool(F.relu(F.conv2d(x, self._weight("meta_conv2", self.meta_conv2.weight))))
        return F.linear(x.reshape(-1, 64 * 4 * 4), self._weight("meta_fc1", self.meta_fc1.weight))


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


# ------------------------------------------------------------- MetaPruning-style ResNet (resnet_meta_2, hyper-network weights)
# The reference's version (``resnet_meta_2.py:9-200``) cannot be constructed (``stage_repeat[3]`` with a 3-element list, blocks whose
# widths do not chain: 16 -> 32 -> Bottleneck(256, 256) ...).  This is the working form of the same idea: no block owns convolution
# weights; two-layer hyper-networks map the requested (mid, in, out) width ratios to the full weight tensor, of which the leading
# channels are used, and every width has its own statistics-only BatchNorm.  Module / parameter names follow the reference's.
META_CHANNEL_SCALES = [(10 + 3 * i) / 100 for i in range(31)]          # 0.10 .. 1.00


def _scaled_bns(max_c: int) -> nn.ModuleList:
    return nn.ModuleList(nn.BatchNorm2d(max(1, int(max_c * s)), affine=False) for s in META_CHANNEL_SCALES)


class first_conv_block(nn.Module):
    def __init__(self, base_inplanes, base_planes, stride, kernel: int = 3, pool: bool = False):
        super().__init__()
        self.max_inp, self.max_oup, self.stride, self.kernel = base_inplanes, base_planes, stride, kernel
        self.fc11 = nn.Linear(1, 32)
        self.fc12 = nn.Linear(32, base_planes * base_inplanes * kernel * kernel)
        self.first_bn = _scaled_bns(base_planes)
        self.maxpool = nn.MaxPool2d(3, 2, 1) if pool else nn.Identity()

    def forward(self, x, oup_scale_id: int):
        s = META_CHANNEL_SCALES[oup_scale_id]
        oup = max(1, int(self.max_oup * s))
        w = self.fc12(F.relu(self.fc11(x.new_tensor([s])))).view(self.max_oup, self.max_inp, self.kernel, self.kernel)
        y = F.conv2d(x, w[:oup], None, self.stride, self.kernel // 2)
        return self.maxpool(F.relu(self.first_bn[oup_scale_id](y)))


class MetaBottleneck(nn.Module):
    expansion = 4

    def __init__(self, base_inplanes, base_planes, stride=1, is_downsample=False):
        super().__init__()
        self.max_inp, self.max_oup, self.max_mid, self.stride = base_inplanes, base_planes, max(1, base_planes // self.expansion), stride
        mid, inp, oup = self.max_mid, self.max_inp, self.max_oup
        self.fc11, self.fc12 = nn.Linear(3, 32), nn.Linear(32, mid * inp)
        self.fc21, self.fc22 = nn.Linear(3, 32), nn.Linear(32, mid * mid * 9)
        self.fc31, self.fc32 = nn.Linear(3, 32), nn.Linear(32, oup * mid)
        self.bn1, self.bn2, self.bn3 = _scaled_bns(mid), _scaled_bns(mid), _scaled_bns(oup)
        self.is_downsample = is_downsample
        if is_downsample:
            self.fc11_downsample, self.fc12_downsample = nn.Linear(3, 32), nn.Linear(32, oup * inp)
            self.bn_downsample = _scaled_bns(oup)

    def forward(self, x, mid_scale_id: int, inp_scale_id: int, oup_scale_id: int):
        ms, is_, os_ = (META_CHANNEL_SCALES[i] for i in (mid_scale_id, inp_scale_id, oup_scale_id))
        mid, oup = max(1, int(self.max_mid * ms)), max(1, int(self.max_oup * os_))
        inp = x.shape[1]                                    # whatever the previous block produced at its chosen width
        r = x.new_tensor([ms, is_, os_])
        w1 = self.fc12(F.relu(self.fc11(r))).view(self.max_mid, self.max_inp, 1, 1)
        w2 = self.fc22(F.relu(self.fc21(r))).view(self.max_mid, self.max_mid, 3, 3)
        w3 = self.fc32(F.relu(self.fc31(r))).view(self.max_oup, self.max_mid, 1, 1)
        y = F.relu(self.bn1[mid_scale_id](F.conv2d(x, w1[:mid, :inp])))
        y = F.relu(self.bn2[mid_scale_id](F.conv2d(y, w2[:mid, :mid], None, self.stride, 1)))
        y = self.bn3[oup_scale_id](F.conv2d(y, w3[:oup, :mid]))
        if self.is_downsample:
            wd = self.fc12_downsample(F.relu(self.fc11_downsample(r))).view(self.max_oup, self.max_inp, 1, 1)
            x = self.bn_downsample[oup_scale_id](F.conv2d(x, wd[:oup, :inp], None, self.stride))
        return F.relu(y + x)


class MetaPruningResNet20(nn.Module):
    """``forward(x, stage_oup_scale_ids, mid_scale_ids)``: ``stage_oup_scale_ids[i]`` is the output width (index into the 31 scales) of
    layer ``i`` (stem + one per block; blocks without a projection keep their input width, so repeated blocks of a stage must be given
    the stage's id), ``mid_scale_ids[i-1]`` the bottleneck width of block ``i``."""

    def __init__(self, num_classes=10, stage_repeat=(2, 2, 2), widths=(64, 128, 256)):
        super().__init__()
        self.layers = nn.ModuleList([first_conv_block(3, 16, stride=1)])
        inp = 16
        for si, (rep, w) in enumerate(zip(stage_repeat, widths)):
            self.layers.append(MetaBottleneck(inp, w, stride=1 if si == 0 else 2, is_downsample=True))
            self.layers.extend(MetaBottleneck(w, w) for _ in range(1, rep))
            inp = w
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(inp, num_classes)
        self.max_width = inp

    @property
    def num_blocks(self):
        return len(self.layers) - 1

    def forward(self, x, stage_oup_scale_ids: Optional[Sequence[int]] = None, mid_scale_ids: Optional[Sequence[int]] = None):
        full = len(META_CHANNEL_SCALES) - 1
        oup_ids = list(stage_oup_scale_ids) if stage_oup_scale_ids is not None else [full] * len(self.layers)
        mid_ids = list(mid_scale_ids) if mid_scale_ids is not None else [full] * self.num_blocks
        for i, block in enumerate(self.layers):
            x = block(x, oup_ids[0]) if i == 0 else block(x, mid_ids[i - 1], oup_ids[i - 1], oup_ids[i])
        f = self.avgpool(x).flatten(1)
        return F.linear(f, self.fc.weight[:, :f.shape[1]], self.fc.bias)        # narrower last stage -> leading classifier columns
