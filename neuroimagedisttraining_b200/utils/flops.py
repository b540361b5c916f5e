"""Sparse-aware FLOP counter.

The reference's counter hooks ``Conv2d``/``Linear`` only and feeds a 1xCxRxR dummy image, so it cannot handle the 3-D
models -- its 3-D clients comment it out and report ``epochs * n_samples`` instead (``fedml_api/utils/main_flops_counter.py:30-163``,
``sailentgrads/client.py:93-99``).  This one hooks Conv1d/2d/3d and Linear, takes the input shape from the model
(``model.in_shape``) or the dataset name, and counts non-zero weights so pruned models report their sparse FLOPs.
Training = 3 x inference, as in the reference (``:30-32``).
"""
from __future__ import annotations

from typing import Optional, Sequence

import numpy as np
import torch
import torch.nn as nn

DATASET_INPUT = {"emnist": (1, 28, 28), "cifar10": (3, 32, 32), "cifar100": (3, 32, 32), "tiny": (3, 64, 64)}


def _input_for(model: nn.Module, dataset: Optional[str], input_shape: Optional[Sequence[int]] = None) -> torch.Tensor:
    if input_shape is None:
        if hasattr(model, "in_shape"):
            input_shape = (1,) + tuple(model.in_shape)
        elif dataset in DATASET_INPUT:
            input_shape = DATASET_INPUT[dataset]
        elif dataset == "ABCD":
            input_shape = (1, 121, 145, 121)
        else:
            raise ValueError(f"cannot infer an input shape for dataset {dataset!r}")
    dev = next(model.parameters()).device
    return torch.rand((1,) + tuple(input_shape), device=dev)


def count_model_param_flops(model: nn.Module, dataset: Optional[str] = None, multiply_adds: bool = True, full: bool = False,
                            input_shape: Optional[Sequence[int]] = None) -> float:
    flops = []

    def conv_hook(m, inp, out):
        out_pos = float(np.prod(out.shape[2:])) * out.shape[0]
        nz = float(m.weight.numel()) if full else float((m.weight.detach() != 0).sum())
        bias = m.out_channels if m.bias is not None else 0
        flops.append((nz * (2 if multiply_adds else 1) + bias) * out_pos)

    def linear_hook(m, inp, out):
        batch = inp[0].shape[0] if inp[0].dim() == 2 else 1
        nz = float(m.weight.numel()) if full else float((m.weight.detach() != 0).sum())
        bias = m.bias.numel() if m.bias is not None else 0
        flops.append(batch * (nz * (2 if multiply_adds else 1) + bias))

    handles = []
    for m in model.modules():
        if isinstance(m, (nn.Conv1d, nn.Conv2d, nn.Conv3d)):
            handles.append(m.register_forward_hook(conv_hook))
        elif isinstance(m, nn.Linear):
            handles.append(m.register_forward_hook(linear_hook))
    was_training = model.training
    model.eval()
    try:
        with torch.no_grad():
            model(_input_for(model, dataset, input_shape))
    finally:
        for h in handles:
            h.remove()
        model.train(was_training)
    return float(sum(flops))


def count_training_flops(model, dataset=None, full: bool = False, input_shape=None) -> float:
    return 3 * count_model_param_flops(model, dataset, full=full, input_shape=input_shape)


def count_inference_flops(model, dataset=None, input_shape=None) -> float:
    return count_model_param_flops(model, dataset, input_shape=input_shape)


def print_model_param_nums(model: nn.Module) -> int:
    total = int(sum((p != 0).sum() for p in model.parameters() if p.dim() >= 2))
    print("  + Number of params: %.2f" % total)
    return total
