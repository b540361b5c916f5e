"""Small helpers of the sparse-training algorithms (reference: ``fedml_api/standalone/DisPFL/slim_util.py:7-19``)."""
from __future__ import annotations

import numpy as np
import torch

from .. import ops


def cosine_annealing(args, round_idx: int) -> float:
    return args.anneal_factor / 2 * (1 + np.cos((round_idx * np.pi) / args.comm_round))


def model_difference(model_a, model_b):
    """``sum ||a-b||^2`` over two state_dicts, two flat tensors or two slabs."""
    if hasattr(model_a, "flat"):
        return ops.squared_distance(model_a.flat, model_b.flat)
    if torch.is_tensor(model_a):
        return ops.squared_distance(model_a, model_b)
    return sum(torch.sum(torch.square(model_a[name] - model_b[name])) for name in model_a)


def hamming_distance(mask_a, mask_b):
    """(number of differing entries, total entries) for two mask dicts / flat tensors / slabs."""
    if hasattr(mask_a, "flat"):
        return ops.hamming_distance(mask_a.flat, mask_b.flat), mask_a.layout.numel_real(params_only=True)
    if torch.is_tensor(mask_a):
        return ops.hamming_distance(mask_a, mask_b), mask_a.numel()
    dis, total = 0, 0
    for key in mask_a:
        dis = dis + torch.sum(mask_a[key].int() ^ mask_b[key].int())
        total += mask_a[key].numel()
    return dis, total
