"""Compute engines: what runs one local step / eval batch on a model bound to a ParamSlab.

``TorchEngine``  module autograd over slab-bound parameters.  This is the CPU path and the
                 scaffolding path for architectures without a native plan (2-D image models);
                 the optimizer / mask / saliency / aggregation still run as flat slab ops.
``B200Engine``   (``ops/engine_b200.py``) the sm_100a path for plan-based 3-D models: hand-written
                 conv / BN / ReLU / MaxPool / head kernels launched directly on the slab.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops
from ..runtime.slab import ParamSlab, SlabLayout


@dataclass
class StepHParams:
    lr: float
    momentum: float = 0.0
    wd: float = 0.0
    clip_norm: float = 10.0


class LazyMean:
    """A mean loss that is still on the device: ``sum`` (0-d tensor, the per-step mean losses added up) over ``steps``.  Lets the
    engines hand out epoch losses without any kernel launch; :meth:`SlabTrainer.flush_losses` reads many of them in one copy."""
    __slots__ = ("sum", "steps")

    def __init__(self, total: torch.Tensor, steps: int):
        self.sum, self.steps = total, max(1, int(steps))

    def item(self) -> float:
        return float(self.sum.item()) / self.steps


class EvalCounters:
    """Device-side accumulators (correct, loss_sum, total): no per-batch ``.item()`` syncs
    (the reference does three per batch, ``my_model_trainer.py:270-273``)."""

    def __init__(self, device, buf: Optional[torch.Tensor] = None):
        """``buf``: a 3-element float64 view to accumulate into (several counters can share one tensor, read back in one copy)."""
        self.buf = torch.zeros(3, dtype=torch.float64, device=device) if buf is None else buf

    def reset(self):
        self.buf.zero_()

    @staticmethod
    def as_result(c, l, t):
        return {"test_correct": c, "test_loss": l, "test_total": t, "test_acc": (c / t if t else 0.0)}

    def result(self):
        from ..runtime import xfer
        c, l, t = xfer.to_host(self.buf).tolist()
        return {"test_correct": c, "test_loss": l, "test_total": t, "test_acc": (c / t if t else 0.0)}


class TorchEngine:
    def __init__(self, model: nn.Module, device, loss_kind: str = "bce", volumetric: bool = True, input_scale: float = 1.0 / 255.0,
                 amp: bool = False):
        self.device = torch.device(device)
        self.model = model.to(self.device)
        self.layout = SlabLayout.from_module(self.model)
        self.loss_kind = loss_kind
        self.volumetric = volumetric
        self.input_scale = input_scale
        self.grad = ParamSlab(self.layout, device=self.device, params_only=True)
        self.slab: Optional[ParamSlab] = None
        self.loss_acc = torch.zeros((), dtype=torch.float32, device=self.device)
        self.loss_steps = 0
        self.native = False
        self.amp = bool(amp) and self.device.type == "cuda"     # bf16 autocast (library kernels) for models without a native plan

    # -- parameter binding ---------------------------------------------------------------
    def new_slab(self) -> ParamSlab:
        """A fresh slab initialised from the module's current values (unbound)."""
        return ParamSlab.from_module(self.model, device=self.device, bind=False)

    def attach(self, slab: ParamSlab) -> None:
        if self.slab is not None and self.slab is not slab:
            self.slab.sync_int_buffers_from(self.model)
        self.slab = slab
        slab.bind_module(self.model, grad=self.grad)
        slab.sync_int_buffers_to(self.model)

    def detach_sync(self) -> None:
        if self.slab is not None:
            self.slab.sync_int_buffers_from(self.model)

    # -- helpers ---------------------------------------------------------------------------
    def _prep(self, x: torch.Tensor) -> torch.Tensor:
        x = x.to(self.device, non_blocking=True)
        if self.volumetric:
            if x.dtype == torch.uint8:
                x = x.to(torch.float32) * self.input_scale
            if x.dim() == 4:
                x = x.unsqueeze(1)
        return x

    def _loss(self, logits, y):
        if isinstance(logits, (list, tuple)):
            logits = logits[0]
        if self.loss_kind == "bce":   # BCEWithLogitsLoss()(logits[B,1], y[:,None]) == mean over B
            return F.binary_cross_entropy_with_logits(logits.flatten(), y.to(self.device).float().flatten())
        return F.cross_entropy(logits, y.to(self.device).long())

    # -- one local step ----------------------------------------------------------------------
    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        """Forward + backward; gradients land in ``self.grad.flat`` (zeroed first)."""
        self.model.train(train_mode)
        self.grad.flat.zero_()
        for name, p in self.model.named_parameters():   # autograd may have replaced .grad; re-pin cheaply
            if p.grad is None or p.grad.data_ptr() != self.grad.view(name).data_ptr():
                p.grad = self.grad.view(name)
        with torch.autocast("cuda", dtype=torch.bfloat16, enabled=self.amp):
            out = self.model(self._prep(x))
        loss = self._loss(out.float() if torch.is_tensor(out) else [o.float() for o in out], y)
        loss.backward()
        return loss.detach()

    def train_step(self, x, y, hp: StepHParams, mask: Optional[torch.Tensor], mom: Optional[torch.Tensor], first_step: bool,
                   grad_mask: Optional[torch.Tensor] = None) -> None:
        loss = self.compute_grads(x, y, train_mode=True)
        self.loss_acc += loss
        self.loss_steps += 1
        if grad_mask is not None:      # Sub-FedAvg masks the gradient before clipping (subavg/my_model_trainer.py:66-68)
            self.grad.flat.mul_(grad_mask)
        ops.masked_sgd_step(self.slab.params_flat, self.grad.flat, mom, mask, hp.lr, hp.momentum, hp.wd, hp.clip_norm, first_step)

    def pop_mean_loss(self, sync: bool = True):
        """Mean loss since the last call; ``sync=False`` returns a 0-d device tensor (no host synchronisation)."""
        mean = self.loss_acc / max(1, self.loss_steps)
        self.loss_acc = torch.zeros_like(self.loss_acc)
        self.loss_steps = 0
        return float(mean.item()) if sync else mean

    # -- evaluation -----------------------------------------------------------------------------
    @torch.no_grad()
    def eval_batch(self, x, y, counters: EvalCounters, double_sigmoid: bool = True) -> None:
        self.model.eval()
        logits = self.model(self._prep(x))
        if isinstance(logits, (list, tuple)):
            logits = logits[0]
        y = y.to(self.device)
        if self.loss_kind == "bce":
            correct, loss = ops.reference.bce_eval_counters(logits, y, double_sigmoid)
            # the reference accumulates mean-loss * batch (``my_model_trainer.py:271``) == summed loss
        else:
            loss = F.cross_entropy(logits, y.long(), reduction="sum")
            correct = (logits.argmax(1) == y.long()).float().sum()
        counters.buf += torch.stack([correct.double(), loss.double(), torch.tensor(float(y.shape[0]), dtype=torch.float64, device=self.device)])


class SpatialTorchEngine(TorchEngine):
    """:class:`TorchEngine` whose forward / backward run depth-slab parallel over ALL ranks of the default process group
    (:class:`~..parallel.spatial.SpatialParallelModel`): every rank executes the same FL programme on the same clients, holds a slab
    of each activation, and ends every step with identical gradients (feature-extractor shares are summed over ranks), so the
    replicas of the model never diverge.  For one client whose activations do not fit a single device."""

    def __init__(self, model: nn.Module, device, **kw):
        super().__init__(model, device, **kw)
        from ..parallel.spatial import SpatialParallelModel
        self.spm = SpatialParallelModel(self.model)

    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        self.model.train(train_mode)
        self.grad.flat.zero_()
        for name, p in self.model.named_parameters():
            if p.grad is None or p.grad.data_ptr() != self.grad.view(name).data_ptr():
                p.grad = self.grad.view(name)
        out = self.spm(self._prep(x))
        loss = self._loss(out.float() if torch.is_tensor(out) else [o.float() for o in out], y)
        loss.backward()
        self.spm.reduce_gradients()
        return loss.detach()

    @torch.no_grad()
    def eval_batch(self, x, y, counters: EvalCounters, double_sigmoid: bool = True) -> None:
        self.model.eval()
        logits = self.spm(self._prep(x))
        if isinstance(logits, (list, tuple)):
            logits = logits[0]
        y = y.to(self.device)
        if self.loss_kind == "bce":
            correct, loss = ops.reference.bce_eval_counters(logits, y, double_sigmoid)
        else:
            loss = F.cross_entropy(logits, y.long(), reduction="sum")
            correct = (logits.argmax(1) == y.long()).float().sum()
        counters.buf += torch.stack([correct.double(), loss.double(), torch.tensor(float(y.shape[0]), dtype=torch.float64, device=self.device)])


class DataParallelTorchEngine(TorchEngine):
    """:class:`TorchEngine` whose training step covers ``1 / world`` of every mini-batch: every rank executes the same FL programme
    on the same clients (``ReplicatedContext``), takes its contiguous share of each batch, BatchNorm layers use the statistics of
    the WHOLE batch (:mod:`..parallel.syncbn`: one small all-reduce per layer forward and backward -- the reference's
    SynchronizedBatchNorm, ``batchnorm_utils.py:150-237``, as one process per device), and the gradients are summed over the ranks,
    so the replicas of the model never diverge.  The loss is the sum over a rank's samples divided by the WHOLE batch size, which puts
    every rank's backward pass in the same scale whatever the shares are (ragged and empty shares included).  The PyTorch-op
    counterpart of :class:`~..ops.engine_dp.DataParallelB200Engine`; with gloo it runs on CPUs (``--data_parallel --device cpu``)."""

    def __init__(self, model: nn.Module, device, group=None, **kw):
        import torch.distributed as dist
        from ..parallel.syncbn import convert_sync_batchnorm
        super().__init__(convert_sync_batchnorm(model, group), device, **kw)
        self._dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        from ..ops.engine_dp import share
        B = int(x.shape[0])
        lo, hi = share(B, self.world, self.rank)
        self.model.train(train_mode)
        self.grad.flat.zero_()
        for name, p in self.model.named_parameters():
            if p.grad is None or p.grad.data_ptr() != self.grad.view(name).data_ptr():
                p.grad = self.grad.view(name)
        out = self.model(self._prep(x[lo:hi]))
        if isinstance(out, (list, tuple)):
            out = out[0]
        yl = y[lo:hi].to(self.device)
        if self.loss_kind == "bce":
            loss = F.binary_cross_entropy_with_logits(out.float().flatten(), yl.float().flatten(), reduction="sum") / B
        else:
            loss = F.cross_entropy(out.float(), yl.long(), reduction="sum") / B
        loss.backward()
        self._dist.all_reduce(self.grad.flat, group=self.group)
        total = loss.detach().clone()
        self._dist.all_reduce(total, group=self.group)
        return total

