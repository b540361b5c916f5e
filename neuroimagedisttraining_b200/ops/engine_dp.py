"""Data-parallel training of ONE client on the sm_100a engine: the ranks of a group split every mini-batch, BatchNorm uses the
statistics of the whole batch, every rank applies the same update.

What crosses devices per step (all of it tiny next to the activations, which never move):

* forward, per BatchNorm layer: the per-channel (sum, sum of squares) of the conv output -- 2*C numbers -- summed over the ranks
  before the finalisation kernel turns them into scale / shift / running statistics;
* backward, per BatchNorm layer: the two per-channel reductions of the BatchNorm backward (sum dY, sum dY*xhat) -- 2*C numbers;
* the gradient slab (one all-reduce) and the mean loss.

With unequal shares (16 samples on 3 ranks = 6 + 5 + 5) each rank's loss is a mean over ITS samples, so its whole backward pass
runs in a scale that is B / B_r times the full batch's; everything downstream of the loss is linear in dY, hence: partial sums are
multiplied by f = B_r / B before the cross-rank sum and divided by it afterwards, gradients are multiplied by f before theirs.
The result equals full-batch training on one device up to summation order (``tests/test_native_dp_gpu.py``: 2 and 3 ranks against
one engine, ``tests/dist_worker_dp_gpu.py``: the same over NCCL).

This is the native counterpart of the reference's only collective call site, SynchronizedBatchNorm
(fedml_api/model/cv/batchnorm_utils.py:150-237: sum / sum-of-squares reduce + broadcast inside DataParallel replicas), here
one process per GPU and an all-reduce; the kernels are the single-device ones, run one launch at a time instead of from the step
executor (the BatchNorm backward is split into its reduce / finalise / apply launches, ``nidt_bn_relu_pool_bwd_phases``).
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import List, Optional, Tuple

import torch

from . import flat as flat_ops
from . import native
from .engine_b200 import B200Engine, _Block


class DistGroup:
    """The ranks that train a client together, over ``torch.distributed`` (NCCL on GPUs)."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self._dist, self.group = dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def all_reduce(self, t: torch.Tensor) -> None:
        self._dist.all_reduce(t, group=self.group)


class ThreadGroup:
    """``world`` engines on ONE device, one Python thread each, summing through host barriers: lets a single-GPU test run the
    multi-rank arithmetic (all threads launch on the same stream, so host order is device order)."""

    def __init__(self, world: int):
        self.world = world
        self._barrier = threading.Barrier(world)
        self._slots: List[Optional[torch.Tensor]] = [None] * world

    def member(self, rank: int) -> "_ThreadMember":
        return _ThreadMember(self, rank)


class _ThreadMember:
    def __init__(self, g: ThreadGroup, rank: int):
        self.g, self.rank, self.world = g, rank, g.world

    def all_reduce(self, t: torch.Tensor) -> None:
        g = self.g
        g._slots[self.rank] = t.clone()
        g._barrier.wait()
        total = g._slots[0].clone()
        for other in g._slots[1:]:
            total += other
        g._barrier.wait()
        t.copy_(total)


def share(B: int, world: int, rank: int) -> Tuple[int, int]:
    """Rows [lo, hi) of a batch of B that ``rank`` takes: contiguous, sizes differ by at most one."""
    base, extra = divmod(B, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


class DataParallelB200Engine(B200Engine):
    """:class:`B200Engine` whose training step covers ``1 / world`` of the batch; see the module docstring.  Evaluation batches are
    not split (every rank holds the model and evaluates what it is given)."""

    def __init__(self, model, device, dp, max_batch: int = 16, **kw):
        super().__init__(model, device, max_batch=max_batch, **kw)
        self.dp = dp
        self.alg_bwd0 = self.fused_bwd0 = False      # block 0 backward through dY (the variants without it are single-device paths)
        self.use_graphs = False          # the step is issued launch by launch (cross-rank sums sit between the launches)
        self._dp_f = 1.0                 # this rank's share of the current batch, B_r / B
        self._dp_B = 0                   # whole-batch size of the current step (0 = not inside a split step)
        self._seed_extra = 104729 * dp.rank      # ranks draw different dropout masks for their rows
        for b in self.blocks:
            Cc = b.spec.cout
            b.dp_sums = torch.zeros(2 * 2 * Cc, dtype=torch.float32, device=self.device)       # [hi | lo][2][C]
            b.dp_c12_local = torch.zeros(2 * Cc, dtype=torch.float32, device=self.device)
        self._rows = C.c_int(0)

    # ------------------------------------------------------------------------- cross-rank sums
    def _global_rows(self, partials: torch.Tensor, rows: int, Cc: int, out: torch.Tensor, weight: float) -> int:
        """Sum ``rows`` rows of [2][C] partial sums on this device (fp64), weight, add over the ranks, unweight; the result is
        written as two fp32 rows (value = hi + lo) so the finalisation kernels, which add rows in fp64, lose nothing."""
        tot = partials[:rows * 2 * Cc].view(rows, 2 * Cc).sum(0, dtype=torch.float64)
        if weight != 1.0:
            tot *= weight
        self.dp.all_reduce(tot)
        if weight != 1.0:
            tot /= weight
        hi = tot.float()
        out[:2 * Cc] = hi
        out[2 * Cc:] = (tot - hi.double()).float()
        return 2

    def _bn_batch_sums(self, b: _Block, tiles: int, count: float):
        if not self._dp_B:
            return super()._bn_batch_sums(b, tiles, count)
        Cc = b.spec.cout
        rows = self._global_rows(b.stats, tiles, Cc, b.dp_sums, 1.0)
        return b.dp_sums.data_ptr(), rows, count / self._dp_f          # elements per channel in the whole batch

    def _bn_backward(self, b: _Block, B: int) -> None:
        if not self._dp_B:
            return super()._bn_backward(b, B)
        lib, st, s = self.lib, self._stream(), b.spec
        D, H, W = b.c_dhw
        Cc = s.cout
        args = (b.y.data_ptr(), b.ymax.data_ptr() if b.ymax is not None else None, b.da.data_ptr(), b.dy.data_ptr(), B, D, H, W, Cc,
                s.pool_k, s.pool_s, b.scale.data_ptr(), b.shift.data_ptr(), b.mean.data_ptr(), b.invstd.data_ptr(),
                b.bwd_partials.data_ptr(), b.c12.data_ptr(), None, None, int(s.relu))
        self._chk(lib.nidt_bn_relu_pool_bwd_phases(*args, 1, C.byref(self._rows), st), "bn_bwd_reduce")
        rows = self._rows.value
        count = float(B * D * H * W)
        # BN weight / bias gradients: this rank's contribution (summed with the other gradients later)
        self._chk(lib.nidt_bn_bwd_finalize(b.bwd_partials.data_ptr(), rows, Cc, count, b.dp_c12_local.data_ptr(),
                                           self._p(self.grad, s.bn + ".weight"), self._p(self.grad, s.bn + ".bias"), st), "bn_bwd_finalize")
        # the two batch means of the BatchNorm backward over the WHOLE batch, in this rank's loss scale
        rows = self._global_rows(b.bwd_partials, rows, Cc, b.dp_sums, self._dp_f)
        self._chk(lib.nidt_bn_bwd_finalize(b.dp_sums.data_ptr(), rows, Cc, count / self._dp_f, b.c12.data_ptr(), None, None, st),
                  "bn_bwd_finalize")
        self._chk(lib.nidt_bn_relu_pool_bwd_phases(*args, 4, None, st), "bn_bwd_apply")
        native.count_launches(4)

    # ---------------------------------------------------------------------------- public API
    def _split_step(self, x: torch.Tensor, y: torch.Tensor) -> None:
        """Forward + backward of this rank's rows; afterwards ``grad`` and the loss accumulator hold whole-batch values."""
        x = self._check_x(x)
        B = int(x.shape[0])
        dp = self.dp
        y = y.to(self.device, torch.float32).contiguous()
        if B < dp.world:               # fewer samples than ranks: everybody computes the whole (tiny) batch, nothing to exchange
            self._dp_B, self._dp_f = 0, 1.0
            extra, self._seed_extra = self._seed_extra, 0          # ... with the same dropout masks
            try:
                Bl = self._forward(x, train=True)
                self._head(Bl, y, train=True)
                self._backward(Bl)
            finally:
                self._seed_extra = extra
            self.step += 1
            return
        lo, hi = share(B, dp.world, dp.rank)
        self._dp_B, self._dp_f = B, (hi - lo) / B
        try:
            before = self.loss_acc.clone()
            Bl = self._forward(x[lo:hi], train=True)
            self._head(Bl, y[lo:hi], train=True)
            self._backward(Bl)
            g = self.grad.flat
            if self._dp_f != 1.0:
                g.mul_(self._dp_f)
            dp.all_reduce(g)
            part = (self.loss_acc - before) * self._dp_f
            dp.all_reduce(part)
            self.loss_acc.copy_(before + part)
        finally:
            self._dp_B = 0
        self.step += 1

    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        if not train_mode:
            raise NotImplementedError("eval-mode gradient screening runs on the TorchEngine")
        self._split_step(x, y)
        return self.loss_acc

    def saliency_step(self, x, y, scores, scale) -> bool:
        return False                     # scores come from the all-reduced gradient slab (the caller's flat pass)

    def train_step(self, x, y, hp, mask, mom, first_step: bool, grad_mask: Optional[torch.Tensor] = None) -> None:
        self._split_step(x, y)
        self.loss_steps += 1
        if grad_mask is not None:
            self.grad.flat.mul_(grad_mask)
        flat_ops.masked_sgd_step(self.slab.params_flat, self.grad.flat, mom, mask, hp.lr, hp.momentum, hp.wd, hp.clip_norm, first_step)
        self._packed_dirty = True
