"""2-D image classifiers used by the CIFAR / Tiny-ImageNet / MNIST-style entry points
(reference: ``fedml_api/model/cv/resnet.py:9-214``, ``vgg.py:14-83``, ``cnn_cifar10.py:12-50``,
``lenet5.py:4-45``, ``cnn.py:6-167``).  state_dict keys follow the reference so checkpoints
interchange.  GroupNorm(32) replaces BatchNorm in the "customized"/"tiny" ResNets and VGG
exactly as the reference does (its post-construction attribute patching is expressed here
as a ``norm`` factory)."""
from __future__ import annotations

from typing import Callable, List, Sequence, Union

import torch
import torch.nn as nn
import torch.nn.functional as F


def _gn(c: int) -> nn.Module:
    return nn.GroupNorm(num_groups=32, num_channels=c)


def _bn(c: int) -> nn.Module:
    return nn.BatchNorm2d(c)


class BasicBlock2D(nn.Module):
    expansion = 1

    def __init__(self, in_planes: int, planes: int, stride: int = 1, norm: Callable[[int], nn.Module] = _bn):
        super().__init__()
        self.conv1 = nn.Conv2d(in_planes, planes, 3, stride, 1, bias=False)
        self.bn1 = norm(planes)
        self.conv2 = nn.Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = norm(planes)
        self.shortcut = nn.Sequential()
        if stride != 1 or in_planes != self.expansion * planes:
            self.shortcut = nn.Sequential(nn.Conv2d(in_planes, self.expansion * planes, 1, stride, bias=False),
                                          norm(self.expansion * planes))

    def forward(self, x):
        y = F.relu(self.bn1(self.conv1(x)))
        y = self.bn2(self.conv2(y))
        return F.relu(y + self.shortcut(x))


class ResNet2D(nn.Module):
    """CIFAR-style ResNet: 3x3 stem, four stages, ``avg_pool2d(4)`` (``adaptive=True`` gives the
    Tiny-ImageNet variant with ``AdaptiveAvgPool2d``)."""

    def __init__(self, block=BasicBlock2D, num_blocks: Sequence[int] = (2, 2, 2, 2), class_num: int = 10,
                 norm: Callable[[int], nn.Module] = _bn, adaptive: bool = False):
        super().__init__()
        self.in_planes = 64
        self.conv1 = nn.Conv2d(3, 64, 3, 1, 1, bias=False)
        self.bn1 = norm(64)
        self.layer1 = self._stage(block, 64, num_blocks[0], 1, norm)
        self.layer2 = self._stage(block, 128, num_blocks[1], 2, norm)
        self.layer3 = self._stage(block, 256, num_blocks[2], 2, norm)
        self.layer4 = self._stage(block, 512, num_blocks[3], 2, norm)
        if adaptive:
            self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.adaptive = adaptive
        self.linear = nn.Linear(512 * block.expansion, class_num)

    def _stage(self, block, planes, n, stride, norm):
        blocks = []
        for s in [stride] + [1] * (n - 1):
            blocks.append(block(self.in_planes, planes, s, norm))
            self.in_planes = planes * block.expansion
        return nn.Sequential(*blocks)

    def forward(self, x):
        x = F.relu(self.bn1(self.conv1(x)))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x) if self.adaptive else F.avg_pool2d(x, 4)
        return self.linear(torch.flatten(x, 1))


def customized_resnet18(class_num: int = 10) -> ResNet2D:
    return ResNet2D(class_num=class_num, norm=_gn)


def original_resnet18(class_num: int = 10) -> ResNet2D:
    return ResNet2D(class_num=class_num, norm=_bn)


def tiny_resnet18(class_num: int = 200) -> ResNet2D:
    return ResNet2D(class_num=class_num, norm=_gn, adaptive=True)


VGG_CFG = {
    "A": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "B": [64, 64, "M", 128, 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "D": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
    "E": [64, 64, "M", 128, 128, "M", 256, 256, 256, 256, "M", 512, 512, 512, 512, "M", 512, 512, 512, 512, "M"],
}


class VGG(nn.Module):
    def __init__(self, cfg: List[Union[int, str]], num_classes: int = 10, group_norm: bool = True):
        super().__init__()
        layers: List[nn.Module] = []
        cin = 3
        for v in cfg:
            if v == "M":
                layers.append(nn.MaxPool2d(2, 2))
            else:
                layers.append(nn.Conv2d(cin, v, 3, padding=1))
                if group_norm:
                    layers.append(_gn(v))
                layers.append(nn.ReLU(inplace=True))
                cin = v
        layers.append(nn.AvgPool2d(1, 1))
        self.features = nn.Sequential(*layers)
        self.classifier = nn.Sequential(nn.Linear(512, num_classes))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                if m.bias is not None:
                    nn.init.zeros_(m.bias)
            elif isinstance(m, nn.GroupNorm):
                nn.init.ones_(m.weight)
                nn.init.zeros_(m.bias)
            elif isinstance(m, nn.Linear):
                nn.init.normal_(m.weight, 0, 0.01)
                nn.init.zeros_(m.bias)

    def forward(self, x):
        return self.classifier(torch.flatten(self.features(x), 1))


def vgg11(num_class: int = 10) -> VGG:
    return VGG(VGG_CFG["A"], num_class)


def vgg16(num_class: int = 10) -> VGG:
    return VGG(VGG_CFG["D"], num_class)


class CifarCNN(nn.Module):
    """conv5-pool-conv5-pool-fc384-fc192-fc (``cnn_cifar10.py:12-50``)."""

    def __init__(self, n_cls: int = 10):
        super().__init__()
        self.n_cls = n_cls
        self.conv1 = nn.Conv2d(3, 64, 5)
        self.conv2 = nn.Conv2d(64, 64, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.fc1 = nn.Linear(64 * 5 * 5, 384)
        self.fc2 = nn.Linear(384, 192)
        self.fc3 = nn.Linear(192, n_cls)

    def forward(self, x):
        x = self.pool(F.relu(self.conv1(x)))
        x = self.pool(F.relu(self.conv2(x)))
        x = F.relu(self.fc1(x.flatten(1)))
        return self.fc3(F.relu(self.fc2(x)))


def cnn_cifar10() -> CifarCNN:
    return CifarCNN(10)


def cnn_cifar100() -> CifarCNN:
    return CifarCNN(100)


class LeNet5(nn.Module):
    """Caffe-style LeNet-5 without first-layer padding (``lenet5.py:4-27``)."""

    def __init__(self, class_num: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 20, 5)
        self.conv2 = nn.Conv2d(20, 50, 5)
        self.fc3 = nn.Linear(50 * 4 * 4, 500)
        self.fc4 = nn.Linear(500, class_num)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.conv1(x)), 2)
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        return self.fc4(F.relu(self.fc3(x.flatten(1))))


class LeNet5_cifar(nn.Module):
    def __init__(self, out_size: int = 10):
        super().__init__()
        self.conv1 = nn.Conv2d(3, 6, 5)
        self.pool = nn.MaxPool2d(2, 2)
        self.conv2 = nn.Conv2d(6, 16, 5)
        self.fc1 = nn.Linear(16 * 5 * 5, 120)
        self.fc2 = nn.Linear(120, 84)
        self.fc3 = nn.Linear(84, out_size)

    def forward(self, x):
        x = self.pool(F.relu(self.conv1(x)))
        x = self.pool(F.relu(self.conv2(x)))
        x = F.relu(self.fc1(x.flatten(1)))
        return self.fc3(F.relu(self.fc2(x)))


class CNN_OriginalFedAvg(nn.Module):
    """The FedAvg-paper EMNIST CNN (``cnn.py:6-74``): 28x28x1 -> 3136 -> 512 -> 10/62."""

    def __init__(self, only_digits: bool = True):
        super().__init__()
        self.conv2d_1 = nn.Conv2d(1, 32, 5, padding=2)
        self.max_pooling = nn.MaxPool2d(2, stride=2)
        self.conv2d_2 = nn.Conv2d(32, 64, 5, padding=2)
        self.flatten = nn.Flatten()
        self.linear_1 = nn.Linear(3136, 512)
        self.linear_2 = nn.Linear(512, 10 if only_digits else 62)
        self.relu = nn.ReLU()

    def forward(self, x):
        if x.dim() == 3:
            x = x.unsqueeze(1)
        x = self.max_pooling(self.relu(self.conv2d_1(x)))
        x = self.max_pooling(self.relu(self.conv2d_2(x)))
        return self.linear_2(self.relu(self.linear_1(self.flatten(x))))


class CNN_DropOut(nn.Module):
    """``cnn.py:77-140``: 28x28x1 -> conv3 -> conv3 -> pool -> 9216 -> 128 -> 10/62."""

    def __init__(self, only_digits: bool = True):
        super().__init__()
        self.conv2d_1 = nn.Conv2d(1, 32, 3)
        self.max_pooling = nn.MaxPool2d(2, stride=2)
        self.conv2d_2 = nn.Conv2d(32, 64, 3)
        self.dropout_1 = nn.Dropout(0.25)
        self.flatten = nn.Flatten()
        self.linear_1 = nn.Linear(9216, 128)
        self.dropout_2 = nn.Dropout(0.5)
        self.linear_2 = nn.Linear(128, 10 if only_digits else 62)
        self.relu = nn.ReLU()

    def forward(self, x):
        if x.dim() == 3:
            x = x.unsqueeze(1)
        x = self.relu(self.conv2d_1(x))
        x = self.dropout_1(self.max_pooling(self.relu(self.conv2d_2(x))))
        x = self.dropout_2(self.relu(self.linear_1(self.flatten(x))))
        return self.linear_2(x)
