"""Cross-rank model aggregation.

Every rank first reduces its *own* sampled clients into a local weighted partial (one flat
kernel, ``ops.weighted_aggregate``); the cross-rank sum + SalientGrads mask multiply + write-back
of the new global model into every rank's slab is then either

* ``fused``  one sm_100a kernel per rank over NVLink symmetric memory (``csrc/sym_agg.cu``):
             the sample-count weighting of the rank's own clients, reduce-scatter by P2P loads / bulk-async
             copies / in-switch multimem reduction of the peers' partials, mask multiply, all-gather by
             stores into every peer's global slab -- no NCCL launch, no extra passes; or
* ``nccl``   the baseline the fused kernel is measured against: ``all_reduce`` of the partial
             followed by an elementwise mask multiply (what a thin wrapper around the reference's
             ``_aggregate`` would do, ``sailentgrads_api.py:212-227``).
"""
from __future__ import annotations

import os
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
        auto = mode == "auto"
        if auto:
            mode = "fused" if (ctx.is_distributed and self.device.type == "cuda") else "nccl"
        if not ctx.is_distributed:
            mode = "local"       # one rank: the weighted-aggregate kernel writes the global slab, nothing crosses a link
        self.mode = mode
        self.partial = torch.zeros(layout.total_numel, dtype=torch.float32, device=self.device)
        self._fused = None
        if ctx.is_distributed and mode == "fused":
            self._fused = self._make_fused(ctx, layout.total_numel, strict=not auto)
            if self._fused is None:
                self.mode = "nccl"
            else:
                self.partial = self._fused.local_partial   # the staging buffer peers read from

    def _make_fused(self, ctx: DistContext, numel: int, strict: bool):
        """The symmetric-memory aggregator, or -- when it was only chosen by ``auto`` and any rank cannot set it up (no P2P /
        NVLink between the GPUs, several nodes, more ranks than the kernel supports) -- None on EVERY rank, so the job runs on
        the NCCL path instead of failing to start.  An explicit ``fused`` request still raises."""
        fused, err = None, None
        try:
            from .symmetric import SymmetricAggregator
            fused = SymmetricAggregator(ctx, numel, self.device)
        except Exception as e:           # noqa: BLE001 -- any set-up failure demotes `auto` to NCCL
            if strict:
                raise
            err = e
        ok = torch.tensor([0.0 if fused is None else 1.0], device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if ok.item() < 1.0:
            if strict:
                raise RuntimeError("fused aggregation could not be set up on every rank")
            import logging
            logging.getLogger(__name__).warning("fused symmetric-memory aggregation unavailable (%s); using the NCCL path", err or "a peer failed")
            return None
        return fused

    def adopt_global(self, slab: ParamSlab) -> None:
        """Let the global model live where the fused kernel's all-gather writes it (no write-back pass); a no-op on the other paths."""
        if self._fused is not None and os.environ.get("NIDT_AGG_ADOPT", "1") != "0":
            self._fused.adopt(slab)

    def aggregate(self, out: ParamSlab, local: Sequence[Tuple[float, ParamSlab]], mask: Optional[ParamSlab] = None) -> ParamSlab:
        """``out = mask * sum_{all ranks} sum_{local} weight * slab``; ``local`` holds this rank's
        (already normalised weight n_i/sum n, slab) pairs, in client order."""
        mflat = mask.flat if mask is not None else None
        if not self.ctx.is_distributed:
            ops.weighted_aggregate(out.flat, [s.flat for _, s in local], [w for w, _ in local], mflat)
            return out
        if self._fused is not None:
            # the sample-count weighting of this rank's clients happens inside the fused kernel, in client order
            self._fused.aggregate(out.flat, [(w, s.flat) for w, s in local], mflat)
            return out
        if local:
            ops.weighted_aggregate(self.partial, [s.flat for _, s in local], [w for w, _ in local], None)
        else:
            self.partial.zero_()
        dist.all_reduce(self.partial, op=dist.ReduceOp.SUM)
        if mflat is not None:
            n = mflat.numel()
            self.partial[:n].mul_(mflat)
        out.flat.copy_(self.partial)
        return out
