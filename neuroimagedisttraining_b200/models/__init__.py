from .volumetric import (ABCD_SHAPE, AlexNet3D_Dropout, AlexNet3D_Deeper_Dropout,  # noqa: F401
                         AlexNet3D_Dropout_Regression, Tiny3DCNN, ResNet3D_l3, ResNet_l3,
                         BasicBlock3D, Bottleneck3D, resnet3d_18, resnet3d_bottleneck, ConvBlockSpec, HeadSpec, trace_shapes)
from .registry import create_model, MODEL_NAMES  # noqa: F401
from . import experimental, planar  # noqa: F401
