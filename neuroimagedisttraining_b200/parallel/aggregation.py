"""Cross-rank model aggregation.

Every rank first reduces its *own* sampled clients into a local weighted partial (one flat
kernel, ``ops.weighted_aggregate``); the cross-rank sum + SalientGrads mask multiply + write-back
of the new global model into every rank's slab is then either

* ``fused``  one sm_100a kernel per rank over NVLink symmetric memory (``csrc/sym_agg.cu``):
             reduce-scatter by P2P loads of the peers' partials, mask multiply, all-gather by P2P
             stores into every peer's global slab -- no NCCL launch, no extra passes; or
* ``nccl``   the baseline the fused kernel is measured against: ``all_reduce`` of the partial
             followed by an elementwise mask multiply (what a thin wrapper around the reference's
             ``_aggregate`` would do, ``sailentgrads_api.py:212-227``).
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from .. import ops
from ..runtime.slab import ParamSlab
from .context import DistContext


class Aggregator:
    def __init__(self, ctx: DistContext, layout, device, mode: str = "auto"):
        self.ctx = ctx
        self.layout = layout
        self.device = torch.device(device)
        if mode == "auto":
            mode = "fused" if (ctx.is_distributed and self.device.type == "cuda") else "nccl"
        if not ctx.is_distributed:
            mode = "local"       # one rank: the weighted-aggregate kernel writes the global slab, nothing crosses a link
        self.mode = mode
        self.partial = torch.zeros(layout.total_numel, dtype=torch.float32, device=self.device)
        self._fused = None
        if ctx.is_distributed and mode == "fused":
            from .symmetric import SymmetricAggregator
            self._fused = SymmetricAggregator(ctx, layout.total_numel, self.device)
            self.partial = self._fused.local_partial   # the staging buffer peers read from

    def aggregate(self, out: ParamSlab, local: Sequence[Tuple[float, ParamSlab]], mask: Optional[ParamSlab] = None) -> ParamSlab:
        """``out = mask * sum_{all ranks} sum_{local} weight * slab``; ``local`` holds this rank's
        (already normalised weight n_i/sum n, slab) pairs, in client order."""
        mflat = mask.flat if mask is not None else None
        if not self.ctx.is_distributed:
            ops.weighted_aggregate(out.flat, [s.flat for _, s in local], [w for w, _ in local], mflat)
            return out
        if local:
            ops.weighted_aggregate(self.partial, [s.flat for _, s in local], [w for w, _ in local], None)
        else:
            self.partial.zero_()
        if self._fused is not None:
            self._fused.reduce_mask_broadcast(out.flat, mflat)
        else:
            dist.all_reduce(self.partial, op=dist.ReduceOp.SUM)
            if mflat is not None:
                n = mflat.numel()
                self.partial[:n].mul_(mflat)
            out.flat.copy_(self.partial)
        return out
