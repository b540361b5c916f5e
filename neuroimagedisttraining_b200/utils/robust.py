"""Robust aggregation defences: norm-difference clipping of a client's update and weak-DP Gaussian noise
(reference: ``fedml_core/robustness/robust_aggregation.py:32-55``).  Operates on flat slabs (or state_dicts);
only trainable parameters count as "weights" (BN running statistics are excluded, ``is_weight_param`` ``:28-29``)."""
from __future__ import annotations

from typing import Optional

import torch

from .. import ops
from ..runtime.slab import ParamSlab


def is_weight_param(k: str) -> bool:
    return "running_mean" not in k and "running_var" not in k and "num_batches_tracked" not in k


def vectorize_weight(state_dict) -> torch.Tensor:
    return torch.cat([v.flatten().float() for k, v in state_dict.items() if is_weight_param(k)])


class RobustAggregator:
    def __init__(self, args):
        self.defense_type = getattr(args, "defense_type", "norm_diff_clipping")
        self.norm_bound = float(getattr(args, "norm_bound", 5.0))
        self.stddev = float(getattr(args, "stddev", 0.025))

    def norm_diff_clipping(self, local, global_):
        """``w_local <- w_global + (w_local - w_global) / max(1, ||diff|| / norm_bound)`` over the parameters."""
        if isinstance(local, ParamSlab):
            P = local.layout.param_numel
            diff = local.flat[:P] - global_.flat[:P]
            norm = float(torch.sqrt(ops.squared_distance(local.flat[:P].contiguous(), global_.flat[:P].contiguous())))
            local.flat[:P] = global_.flat[:P] + diff / max(1.0, norm / self.norm_bound)
            return local
        vl, vg = vectorize_weight(local), vectorize_weight(global_)
        diff = vl - vg
        clipped = diff / max(1.0, float(torch.norm(diff)) / self.norm_bound)
        out, off = {}, 0
        for k, v in local.items():
            if is_weight_param(k):
                n = v.numel()
                out[k] = (global_[k].float().flatten() + clipped[off:off + n]).view_as(v).to(v.dtype)
                off += n
            else:
                out[k] = v
        return out

    def add_noise(self, local_weight: torch.Tensor, device=None, generator: Optional[torch.Generator] = None) -> torch.Tensor:
        noise = torch.randn(local_weight.size(), device=device or local_weight.device, generator=generator) * self.stddev
        return local_weight + noise
