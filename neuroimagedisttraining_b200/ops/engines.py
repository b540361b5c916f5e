"""Picks the sm_100a compute engine for a model (or raises ImportError when only the PyTorch engine can run it)."""
from __future__ import annotations


def make_native_engine(model, device, max_batch: int, input_scale: float = 1.0 / 255.0, seed: int = 1234):
    """AlexNet3D family / Tiny3DCNN (``model.plan``) -> :class:`B200Engine`; ``ResNet3D_l3`` with basic blocks ->
    :class:`ResNet3DEngine`.  Both consume 8-bit volumes and keep parameters in a :class:`ParamSlab`."""
    from ..models.volumetric import ResNet3D_l3
    if isinstance(model, ResNet3D_l3):
        from .engine_resnet_b200 import ResNet3DEngine
        return ResNet3DEngine(model, device, max_batch=max_batch, input_scale=input_scale, seed=seed)
    from .engine_b200 import B200Engine
    return B200Engine(model, device, max_batch=max_batch, input_scale=input_scale, seed=seed)
