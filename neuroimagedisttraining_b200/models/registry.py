"""``create_model`` -- the name -> constructor table every CLI uses (reference:
``main_sailentgrads.py:164-178``, ``main_fedavg.py:156-170``, ``main_dispfl.py:163-177``)."""
from __future__ import annotations

import torch.nn as nn

from . import planar, volumetric


def create_model(model_name: str, class_num: int = 1, in_shape=volumetric.ABCD_SHAPE, dataset: str = "ABCD") -> nn.Module:
    name = model_name.lower()
    if name in ("3dcnn", "alexnet3d", "alexnet3d_dropout"):
        return volumetric.AlexNet3D_Dropout(num_classes=class_num, in_shape=in_shape)
    if name in ("3dcnn_deeper", "alexnet3d_deeper_dropout"):
        return volumetric.AlexNet3D_Deeper_Dropout(num_classes=class_num, in_shape=in_shape)
    if name in ("3dcnn_regression", "alexnet3d_dropout_regression"):
        return volumetric.AlexNet3D_Dropout_Regression(num_classes=class_num, in_shape=in_shape)
    if name in ("3dresnet", "resnet3d", "resnet_l3", "resnet3d_18"):
        return volumetric.resnet3d_18(num_classes=class_num, in_shape=in_shape)
    if name in ("3dresnet_bottleneck", "resnet3d_bottleneck", "resnet_l3_bottleneck"):
        return volumetric.resnet3d_bottleneck(num_classes=class_num, in_shape=in_shape)
    if name in ("tiny3dcnn", "tiny3d"):
        return volumetric.Tiny3DCNN(num_classes=class_num, in_shape=in_shape)
    if name == "resnet18":
        if dataset == "tiny":
            return planar.tiny_resnet18(class_num=class_num)
        return planar.customized_resnet18(class_num=class_num)
    if name == "resnet18_bn":
        return planar.original_resnet18(class_num=class_num)
    if name == "vgg11":
        return planar.vgg11(class_num)
    if name == "vgg16":
        return planar.vgg16(class_num)
    if name == "cnn_cifar10":
        return planar.cnn_cifar10()
    if name == "cnn_cifar100":
        return planar.cnn_cifar100()
    if name == "lenet5":
        return planar.LeNet5(class_num)
    if name == "lenet5_cifar":
        return planar.LeNet5_cifar(class_num)
    if name == "cnn_fedavg":
        return planar.CNN_OriginalFedAvg(only_digits=class_num == 10)
    if name == "cnn_dropout":
        return planar.CNN_DropOut(only_digits=class_num == 10)
    raise ValueError(f"unknown model {model_name!r}")


MODEL_NAMES = ["3DCNN", "3DCNN_deeper", "3DCNN_regression", "3DResNet", "3DResNet_bottleneck", "tiny3DCNN", "resnet18", "resnet18_bn",
               "vgg11", "vgg16", "cnn_cifar10", "cnn_cifar100", "lenet5", "lenet5_cifar", "cnn_fedavg", "cnn_dropout"]
