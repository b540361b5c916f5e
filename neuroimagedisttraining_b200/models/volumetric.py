"""3-D sMRI classifiers, state_dict-key compatible with the reference
(``fedml_api/model/cv/salient_models.py:84-297``) but built from a declarative *plan*:
each model carries ``self.plan`` -- a list of :class:`ConvBlockSpec` / :class:`HeadSpec`
records -- which the sm_100a engine (``ops/engine.py``) walks to launch the fused
conv -> BN -> ReLU -> MaxPool kernels directly on the parameter slab.  ``forward`` below is
the plain PyTorch restatement used on CPU and as the numerics oracle.

Differences from the reference, on purpose:
* ``ResNet3D_l3`` sizes its first FC from the input shape (the reference hard-codes 9216
  features, ``salient_models.py:99``, which does not fit 121x145x121) and returns logits,
  not ``[logits, feat]`` (``:139``); ``return_features=True`` restores the pair.
* ``Tiny3DCNN`` has no counterpart (the reference cannot run 32^3 inputs); it exists for
  the CPU plumbing config of BASELINE.json.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch.nn as nn

ABCD_SHAPE = (121, 145, 121)


@dataclass(frozen=True)
class ConvBlockSpec:
    """conv3d (+bias) -> BatchNorm3d -> ReLU -> optional MaxPool3d, as one fused unit."""
    conv: str            # module path of the conv inside the model, e.g. "features.0"
    bn: str
    cin: int
    cout: int
    k: int
    stride: int
    pad: int
    pool_k: int = 0      # 0 = no pooling
    pool_s: int = 0
    pool_p: int = 0
    bias: bool = True
    relu: bool = True


@dataclass(frozen=True)
class HeadSpec:
    """Dropout -> Linear -> ReLU -> Dropout -> Linear (the AlexNet3D classifier)."""
    fc1: str
    fc2: str
    in_features: int
    hidden: int
    out_features: int
    p_drop: float


def _conv_out(n: int, k: int, s: int, p: int) -> int:
    return (n + 2 * p - k) // s + 1


def trace_shapes(plan: Sequence[ConvBlockSpec], in_shape: Tuple[int, int, int]) -> List[Tuple[Tuple[int, int, int], Tuple[int, int, int]]]:
    """(conv output DHW, block output DHW) for every block of ``plan``."""
    out = []
    cur = tuple(in_shape)
    for b in plan:
        c = tuple(_conv_out(n, b.k, b.stride, b.pad) for n in cur)
        a = tuple(_conv_out(n, b.pool_k, b.pool_s, b.pool_p) for n in c) if b.pool_k else c
        out.append((c, a))
        cur = a
    return out


class _AlexNet3DBase(nn.Module):
    """Shared builder for the three AlexNet3D variants of the reference."""

    # (cin, cout, k, stride, pad, pool)
    CFG: Sequence[Tuple[int, int, int, int, int, bool]] = ()
    HEAD_NAME = "classifier"

    def __init__(self, num_classes: int, in_shape=ABCD_SHAPE, p_drop: float = 0.5):
        super().__init__()
        layers: List[nn.Module] = []
        plan: List[ConvBlockSpec] = []
        for (cin, cout, k, s, p, pool) in self.CFG:
            i = len(layers)
            layers += [nn.Conv3d(cin, cout, kernel_size=k, stride=s, padding=p), nn.BatchNorm3d(cout), nn.ReLU(inplace=True)]
            if pool:
                layers.append(nn.MaxPool3d(kernel_size=3, stride=3))
            plan.append(ConvBlockSpec(f"features.{i}", f"features.{i + 1}", cin, cout, k, s, p,
                                      pool_k=3 if pool else 0, pool_s=3 if pool else 0))
        self.features = nn.Sequential(*layers)
        shapes = trace_shapes(plan, in_shape)
        d, h, w = shapes[-1][1]
        if min(d, h, w) < 1:
            raise ValueError(f"input shape {in_shape} is too small for {type(self).__name__}")
        feat = self.CFG[-1][1] * d * h * w
        head = nn.Sequential(nn.Dropout(p_drop), nn.Linear(feat, 64), nn.ReLU(inplace=True),
                             nn.Dropout(p_drop), nn.Linear(64, num_classes))
        setattr(self, self.HEAD_NAME, head)
        self.plan = plan
        self.head_spec = HeadSpec(f"{self.HEAD_NAME}.1", f"{self.HEAD_NAME}.4", feat, 64, num_classes, p_drop)
        self.in_shape = tuple(in_shape)
        self.num_classes = num_classes
        # the reference's init loop only touches Conv2d (never present) and resets BN affine
        for m in self.modules():
            if isinstance(m, nn.BatchNorm3d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    def head(self) -> nn.Sequential:
        return getattr(self, self.HEAD_NAME)

    def forward_features(self, x):
        return self.features(x)

    def forward(self, x):
        f = self.features(x)
        return self.head()(f.flatten(1))


class AlexNet3D_Dropout(_AlexNet3DBase):
    """The model every ABCD entry point builds (``salient_models.py:142-191``)."""
    CFG = ((1, 64, 5, 2, 0, True), (64, 128, 3, 1, 0, True), (128, 192, 3, 1, 1, False),
           (192, 192, 3, 1, 1, False), (192, 128, 3, 1, 1, True))

    def __init__(self, num_classes: int = 2, in_shape=ABCD_SHAPE, p_drop: float = 0.5):
        super().__init__(num_classes, in_shape, p_drop)


class AlexNet3D_Deeper_Dropout(_AlexNet3DBase):
    """``salient_models.py:194-246``; the reference returns ``[x, x]`` -- we return logits."""
    CFG = ((1, 64, 5, 2, 0, True), (64, 128, 3, 1, 0, True), (128, 192, 3, 1, 1, False),
           (192, 384, 3, 1, 1, False), (384, 256, 3, 1, 1, False), (256, 256, 3, 1, 1, True))

    def __init__(self, num_classes: int = 2, in_shape=ABCD_SHAPE, p_drop: float = 0.5):
        super().__init__(num_classes, in_shape, p_drop)


class AlexNet3D_Dropout_Regression(_AlexNet3DBase):
    """``salient_models.py:248-297``: head is called ``regressor``, output squeezed."""
    CFG = AlexNet3D_Dropout.CFG
    HEAD_NAME = "regressor"

    def __init__(self, num_classes: int = 1, in_shape=ABCD_SHAPE, p_drop: float = 0.5, return_features: bool = False):
        super().__init__(num_classes, in_shape, p_drop)
        self.return_features = return_features

    def forward(self, x):
        xp = self.features(x)
        y = self.head()(xp.flatten(1)).squeeze()
        return [y, xp] if self.return_features else y


class Tiny3DCNN(nn.Module):
    """Two conv blocks + the AlexNet-style head for small volumes (BASELINE.json config 1:
    4 clients x 8 synthetic 32^3 volumes on CPU).  Same key naming as AlexNet3D."""

    def __init__(self, num_classes: int = 1, in_shape=(32, 32, 32), width: int = 8, p_drop: float = 0.0):
        super().__init__()
        cfg = ((1, width, 3, 2, 0, True), (width, 2 * width, 3, 1, 1, True))
        layers: List[nn.Module] = []
        plan: List[ConvBlockSpec] = []
        for (cin, cout, k, s, p, pool) in cfg:
            i = len(layers)
            layers += [nn.Conv3d(cin, cout, k, s, p), nn.BatchNorm3d(cout), nn.ReLU(inplace=True), nn.MaxPool3d(3, 3)]
            plan.append(ConvBlockSpec(f"features.{i}", f"features.{i + 1}", cin, cout, k, s, p, 3, 3))
        self.features = nn.Sequential(*layers)
        d, h, w = trace_shapes(plan, in_shape)[-1][1]
        feat = 2 * width * d * h * w
        self.classifier = nn.Sequential(nn.Dropout(p_drop), nn.Linear(feat, 16), nn.ReLU(inplace=True),
                                        nn.Dropout(p_drop), nn.Linear(16, num_classes))
        self.plan = plan
        self.head_spec = HeadSpec("classifier.1", "classifier.4", feat, 16, num_classes, p_drop)
        self.in_shape = tuple(in_shape)
        self.num_classes = num_classes

    def forward(self, x):
        return self.classifier(self.features(x).flatten(1))


# --------------------------------------------------------------------------- ResNet
def _conv3x3(cin, cout, stride=1):
    return nn.Conv3d(cin, cout, kernel_size=3, stride=stride, padding=1, bias=False)


class BasicBlock3D(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv3x3(inplanes, planes, stride)
        self.bn1 = nn.BatchNorm3d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = _conv3x3(planes, planes)
        self.bn2 = nn.BatchNorm3d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return self.relu(y + idt)


class Bottleneck3D(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = nn.Conv3d(inplanes, planes, kernel_size=1, bias=False)
        self.bn1 = nn.BatchNorm3d(planes)
        self.conv2 = nn.Conv3d(planes, planes, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = nn.BatchNorm3d(planes)
        self.conv3 = nn.Conv3d(planes, planes * 4, kernel_size=1, bias=False)
        self.bn3 = nn.BatchNorm3d(planes * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        idt = x if self.downsample is None else self.downsample(x)
        y = self.relu(self.bn1(self.conv1(x)))
        y = self.relu(self.bn2(self.conv2(y)))
        y = self.bn3(self.conv3(y))
        return self.relu(y + idt)


class ResNet3D_l3(nn.Module):
    """Three-stage 3-D ResNet (``salient_models.py:84-139``); ``ResNet_l3`` is an alias.
    ``layers=[2,2,2]`` with :class:`BasicBlock3D` is the "3-D ResNet-18" of BASELINE config 4."""

    def __init__(self, block=BasicBlock3D, layers=(2, 2, 2), num_classes: int = 1, in_shape=ABCD_SHAPE,
                 return_features: bool = False):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv3d(1, 64, kernel_size=3, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm3d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool3d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.avgpool = nn.AvgPool3d(3)
        feat = self._feature_count(in_shape) * 256 * block.expansion
        self.fc = nn.Linear(feat, 512)
        self.fc2 = nn.Linear(512, num_classes)
        self.return_features = return_features
        self.in_shape = tuple(in_shape)
        self.num_classes = num_classes
        for m in self.modules():
            if isinstance(m, nn.BatchNorm3d):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)

    @staticmethod
    def _feature_count(in_shape) -> int:
        n = 1
        for s in in_shape:
            s = _conv_out(s, 3, 2, 3)      # conv1
            s = _conv_out(s, 3, 2, 1)      # maxpool
            s = _conv_out(s, 3, 2, 1)      # layer2
            s = _conv_out(s, 3, 2, 1)      # layer3
            s = s // 3                     # avgpool(3)
            if s < 1:
                raise ValueError(f"input shape {in_shape} too small for ResNet3D_l3")
            n *= s
        return n

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(
                nn.Conv3d(self.inplanes, planes * block.expansion, kernel_size=1, stride=stride, bias=False),
                nn.BatchNorm3d(planes * block.expansion))
        seq = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        seq += [block(self.inplanes, planes) for _ in range(1, blocks)]
        return nn.Sequential(*seq)

    def forward(self, x):
        x = self.maxpool(self.relu(self.bn1(self.conv1(x))))
        x = self.layer3(self.layer2(self.layer1(x)))
        x = self.avgpool(x).flatten(1)
        x1 = self.fc(x)
        out = self.fc2(x1)
        return [out, x1] if self.return_features else out


ResNet_l3 = ResNet3D_l3


def resnet3d_18(num_classes: int = 1, in_shape=ABCD_SHAPE) -> ResNet3D_l3:
    return ResNet3D_l3(BasicBlock3D, (2, 2, 2), num_classes, in_shape)


def resnet3d_bottleneck(num_classes: int = 1, in_shape=ABCD_SHAPE, layers=(2, 2, 2)) -> ResNet3D_l3:
    """``ResNet_l3(Bottleneck, layers)`` (``salient_models.py:45-81``): 1x1x1 -> 3x3x3 -> 1x1x1 blocks, 4x channel expansion."""
    return ResNet3D_l3(Bottleneck3D, tuple(layers), num_classes, in_shape)
