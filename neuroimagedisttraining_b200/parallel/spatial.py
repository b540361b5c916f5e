"""Spatial (depth-slab) decomposition of ONE client's volumes across ranks, with halo exchange.

The reference has no sequence / context parallelism (CNNs on a fixed 121x145x121 grid, batch 16 on one GPU); the
analogue for volumetric inputs is domain decomposition: each rank holds a contiguous slab of the depth axis of every activation,
and a layer with kernel k / stride s / padding p first pulls the input planes its output slab needs but a neighbour owns (the
halo: k - s planes per side for the usual shapes), then runs the ordinary local convolution or pooling with no depth padding.
This lets a single client use several GPUs when its activations do not fit one (larger batches, the 3-D ResNet at high
resolution), at the price of one small plane exchange per layer -- on B200 these are P2P copies of a few MB over NVLink.

Pieces:
  * :class:`DepthPartition`  -- who owns which planes of a depth axis (balanced, contiguous).
  * :func:`reslab`           -- autograd-aware redistribution "I own [i0,i1), I need [a,b)" through point-to-point sends; planes
                                outside the volume are zeros (= the layer's padding); the backward pass routes every plane's
                                gradient back to its owner and sums what several consumers needed.
  * :func:`slab_conv3d`, :func:`slab_max_pool3d`, :func:`slab_avg_pool3d` -- the layer wrappers.
  * :class:`SlabParallelFeatures` -- runs an ``nn.Sequential`` of Conv3d / BatchNorm3d / ReLU / MaxPool3d / AvgPool3d slab-parallel
                                (BatchNorm statistics span the ranks through :mod:`.syncbn`), gathers the final feature map.
  * :func:`allreduce_gradients` -- every rank sees only its slab's share of dW; sum them.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist
import torch.nn as nn
import torch.nn.functional as F

from .syncbn import SynchronizedBatchNorm, _SyncBNFn


def _triple(v) -> Tuple[int, int, int]:
    return (v, v, v) if isinstance(v, int) else tuple(v)


class DepthPartition:
    """Planes ``[bounds[r], bounds[r+1])`` of a depth axis of length ``depth`` belong to rank ``r``."""

    def __init__(self, depth: int, world: int, bounds: Optional[Sequence[int]] = None):
        self.depth, self.world = int(depth), int(world)
        if bounds is None:
            base, rem = divmod(self.depth, self.world)
            bounds = [0]
            for r in range(self.world):
                bounds.append(bounds[-1] + base + (1 if r < rem else 0))
        self.bounds = [int(b) for b in bounds]
        assert len(self.bounds) == self.world + 1 and self.bounds[0] == 0 and self.bounds[-1] == self.depth

    def own(self, rank: int) -> Tuple[int, int]:
        return self.bounds[rank], self.bounds[rank + 1]

    def needed_input(self, rank: int, k: int, s: int, p: int) -> Tuple[int, int]:
        """Input planes (may run outside ``[0, Din)`` = padding) that produce this rank's output planes."""
        o0, o1 = self.own(rank)
        if o1 <= o0:
            return 0, 0
        return o0 * s - p, (o1 - 1) * s - p + k

    def __repr__(self):
        return f"DepthPartition(depth={self.depth}, bounds={self.bounds})"


def out_depth(din: int, k: int, s: int, p: int, ceil_mode: bool = False) -> int:
    assert not ceil_mode, "ceil_mode pooling is not supported slab-parallel"
    return (din + 2 * p - k) // s + 1


# ---------------------------------------------------------------------------------------- plane redistribution
def _exchange(x: torch.Tensor, have: Sequence[Tuple[int, int]], want: Sequence[Tuple[int, int]], rank: int, depth: int, group,
              accumulate: bool) -> torch.Tensor:
    """Every rank r holds planes ``have[r]`` (clipped to the volume when they describe a padded range) and must end up with planes
    ``want[r]``.  ``accumulate=False``: copy (forward; a plane has one owner).  ``accumulate=True``: sum (backward; a plane's
    gradient arrives from every rank that consumed it).  Depth is dim 2 of ``x``."""
    h0, h1 = have[rank]
    w0, w1 = want[rank]
    shape = list(x.shape)
    shape[2] = max(w1 - w0, 0)
    out = x.new_zeros(shape)
    ops, recvs = [], []
    world = len(have)
    for peer in range(world):
        # what I hold that the peer wants
        ph0, ph1 = want[peer]
        s0, s1 = max(h0, ph0, 0), min(h1, ph1, depth)
        if s1 > s0:
            chunk = x[:, :, s0 - h0:s1 - h0]
            if peer == rank:
                if accumulate:
                    out[:, :, s0 - w0:s1 - w0] += chunk
                else:
                    out[:, :, s0 - w0:s1 - w0] = chunk
            else:
                ops.append(dist.P2POp(dist.isend, chunk.contiguous(), peer if group is None else dist.get_global_rank(group, peer), group))
        # what the peer holds that I want
        qh0, qh1 = have[peer]
        r0, r1 = max(qh0, w0, 0), min(qh1, w1, depth)
        if r1 > r0 and peer != rank:
            buf_shape = list(x.shape)
            buf_shape[2] = r1 - r0
            buf = x.new_empty(buf_shape)
            ops.append(dist.P2POp(dist.irecv, buf, peer if group is None else dist.get_global_rank(group, peer), group))
            recvs.append((r0, r1, buf))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    for r0, r1, buf in recvs:
        if accumulate:
            out[:, :, r0 - w0:r1 - w0] += buf
        else:
            out[:, :, r0 - w0:r1 - w0] = buf
    return out


class _Reslab(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, own, need, rank, depth, group):
        ctx.own, ctx.need, ctx.rank, ctx.depth, ctx.group = own, need, rank, depth, group
        return _exchange(x, own, need, rank, depth, group, accumulate=False)

    @staticmethod
    def backward(ctx, g):
        # reverse routing: I hold gradients for the planes I needed, their owners want them back (summed over consumers)
        return _exchange(g.contiguous(), ctx.need, ctx.own, ctx.rank, ctx.depth, ctx.group, accumulate=True), None, None, None, None, None


def reslab(x: torch.Tensor, part: DepthPartition, need: Sequence[Tuple[int, int]], rank: int, group=None) -> torch.Tensor:
    """``x`` = this rank's slab under ``part`` -> planes ``need[rank]`` (zeros outside the volume)."""
    own = [part.own(r) for r in range(part.world)]
    if part.world == 1:
        a, b = need[0]
        lo, hi = max(0, -a), max(0, b - part.depth)
        body = x[:, :, max(a, 0):min(b, part.depth)]
        return F.pad(body, (0, 0, 0, 0, lo, hi)) if (lo or hi) else body
    return _Reslab.apply(x, tuple(own), tuple(tuple(n) for n in need), rank, part.depth, group)


# ---------------------------------------------------------------------------------------------- layer wrappers
def _windowed(x, part: DepthPartition, rank: int, k: int, s: int, p: int, group) -> Tuple[torch.Tensor, DepthPartition]:
    out_part = DepthPartition(out_depth(part.depth, k, s, p), part.world)
    need = [out_part.needed_input(r, k, s, p) for r in range(part.world)]
    return reslab(x, part, need, rank, group), out_part


def slab_conv3d(x, part: DepthPartition, rank: int, weight, bias=None, stride=1, padding=0, group=None):
    """``F.conv3d`` on a depth slab: halo planes come from the neighbours, depth padding is materialised as zero planes."""
    (kd, _, _), (sd, sh, sw), (pd, ph, pw) = weight.shape[2:], _triple(stride), _triple(padding)
    xin, out_part = _windowed(x, part, rank, kd, sd, pd, group)
    if xin.shape[2] < kd:                      # this rank owns no output plane
        n, _, _, h, w = xin.shape
        ho, wo = (h + 2 * ph - weight.shape[3]) // sh + 1, (w + 2 * pw - weight.shape[4]) // sw + 1
        # keep the graph connected: this rank must still take part in the backward exchanges and the gradient all-reduce
        return xin.new_zeros((n, weight.shape[0], 0, ho, wo)) + 0 * weight.sum() + 0 * xin.sum(), out_part
    return F.conv3d(xin, weight, bias, (sd, sh, sw), (0, ph, pw)), out_part


def _slab_pool(fn, x, part, rank, kernel_size, stride, padding, group, pad_value):
    (kd, kh, kw), (sd, sh, sw), (pd, ph, pw) = _triple(kernel_size), _triple(stride if stride is not None else kernel_size), _triple(padding)
    xin, out_part = _windowed(x, part, rank, kd, sd, pd, group)
    if pd and pad_value != 0.0:
        # planes outside the volume must not win a max: overwrite the zero padding reslab produced
        a, b = out_part.needed_input(rank, kd, sd, pd)
        lo, hi = max(0, -a), max(0, b - part.depth)
        if lo or hi:
            xin = xin.clone()
            if lo:
                xin[:, :, :lo] = pad_value
            if hi:
                xin[:, :, xin.shape[2] - hi:] = pad_value
    if xin.shape[2] < kd:
        n, c, _, h, w = xin.shape
        return xin.new_zeros((n, c, 0, (h + 2 * ph - kh) // sh + 1, (w + 2 * pw - kw) // sw + 1)) + 0 * xin.sum(), out_part
    if ph or pw:
        xin = F.pad(xin, (pw, pw, ph, ph, 0, 0), value=pad_value)
    return fn(xin, (kd, kh, kw), (sd, sh, sw), 0), out_part


def slab_max_pool3d(x, part, rank, kernel_size, stride=None, padding=0, group=None):
    return _slab_pool(F.max_pool3d, x, part, rank, kernel_size, stride, padding, group, float("-inf"))


def slab_avg_pool3d(x, part, rank, kernel_size, stride=None, padding=0, group=None):
    """``count_include_pad=True`` semantics (torch's default): padded planes contribute zeros."""
    return _slab_pool(F.avg_pool3d, x, part, rank, kernel_size, stride, padding, group, 0.0)


class _GatherDepth(torch.autograd.Function):
    """All ranks end up with the whole depth axis; backward hands every rank the gradient of its own slab, summed over ranks
    (each rank's copy of the gathered tensor may feed a different loss term; with identical heads the caller divides by world)."""

    @staticmethod
    def forward(ctx, x, part, rank, group):
        ctx.part, ctx.rank, ctx.group = part, rank, group
        full = [(0, part.depth)] * part.world
        own = [part.own(r) for r in range(part.world)]
        return _exchange(x, own, full, rank, part.depth, group, accumulate=False)

    @staticmethod
    def backward(ctx, g):
        part = ctx.part
        full = [(0, part.depth)] * part.world
        own = [part.own(r) for r in range(part.world)]
        return _exchange(g.contiguous(), full, own, ctx.rank, part.depth, ctx.group, accumulate=True), None, None, None


def gather_depth(x, part: DepthPartition, rank: int, group=None) -> torch.Tensor:
    if part.world == 1:
        return x
    return _GatherDepth.apply(x, part, rank, group)


def scatter_depth(x_full: torch.Tensor, world: int, rank: int) -> Tuple[torch.Tensor, DepthPartition]:
    """This rank's slab of a full volume batch ``[N, C, D, H, W]`` (the loader side of the decomposition)."""
    part = DepthPartition(x_full.shape[2], world)
    a, b = part.own(rank)
    return x_full[:, :, a:b].contiguous(), part


# ------------------------------------------------------------------------------------------------- module runner
class SlabParallelFeatures(nn.Module):
    """Runs a feature extractor (``nn.Sequential`` of Conv3d / BatchNorm3d / ReLU / MaxPool3d / AvgPool3d / Dropout / Identity)
    on depth slabs.  The wrapped module's parameters are used as they are, so ``state_dict`` keys do not change and the same
    module still runs un-split.  ``forward`` takes this rank's input slab and returns the gathered feature map, identical on
    every rank; the loss computed from it on every rank is the same, hence ``gather_depth``'s summed backward is divided by
    the world size to keep gradients equal to the single-process ones."""

    def __init__(self, features: nn.Sequential, group=None):
        super().__init__()
        self.features = features
        self.group = group

    def _world_rank(self):
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(self.group), dist.get_rank(self.group)
        return 1, 0

    def forward(self, x_slab: torch.Tensor, part: DepthPartition) -> torch.Tensor:
        world, rank = self._world_rank()
        assert part.world == world
        x, part = self._run(self.features, x_slab, part, rank)
        out = gather_depth(x, part, rank, self.group)
        if world > 1 and out.requires_grad:
            out.register_hook(lambda g: g / world)
        return out

    def _run(self, m: nn.Module, x: torch.Tensor, part: DepthPartition, rank: int) -> Tuple[torch.Tensor, DepthPartition]:
        """One module on a slab -> (output slab, its partition).  Containers recurse; residual blocks (``conv1/bn1/conv2/bn2[/conv3/bn3]``
        + optional ``downsample``, the 3-D ResNet's BasicBlock / Bottleneck) run both branches on slabs and add them."""
        if isinstance(m, nn.Sequential):
            for child in m:
                x, part = self._run(child, x, part, rank)
            return x, part
        if all(hasattr(m, a) for a in ("conv1", "bn1", "conv2", "bn2", "downsample")):
            idt, ipart = (x, part) if m.downsample is None else self._run(m.downsample, x, part, rank)
            y, p = self._run(m.conv1, x, part, rank)
            y = F.relu(self._batchnorm(m.bn1, y))
            y, p = self._run(m.conv2, y, p, rank)
            y = self._batchnorm(m.bn2, y)
            if hasattr(m, "conv3"):
                y, p = self._run(m.conv3, F.relu(y), p, rank)
                y = self._batchnorm(m.bn3, y)
            assert p.bounds == ipart.bounds, "the two branches of a residual block must end on the same slabs"
            return F.relu(y + idt), p
        if isinstance(m, nn.Conv3d):
            assert m.dilation == (1, 1, 1) and m.groups == 1 and m.padding_mode == "zeros", "plain Conv3d only"
            return slab_conv3d(x, part, rank, m.weight, m.bias, m.stride, m.padding, self.group)
        if isinstance(m, nn.modules.batchnorm._BatchNorm):
            return self._batchnorm(m, x), part
        if isinstance(m, nn.MaxPool3d):
            assert not m.ceil_mode and _triple(m.dilation) == (1, 1, 1)
            return slab_max_pool3d(x, part, rank, m.kernel_size, m.stride, m.padding, self.group)
        if isinstance(m, nn.AvgPool3d):
            assert not m.ceil_mode and m.count_include_pad
            return slab_avg_pool3d(x, part, rank, m.kernel_size, m.stride, m.padding, self.group)
        if isinstance(m, (nn.ReLU, nn.Dropout, nn.Identity, nn.LeakyReLU, nn.Sigmoid, nn.Tanh)):
            return (F.relu(x) if isinstance(m, nn.ReLU) else m(x)), part
        raise NotImplementedError(f"{type(m).__name__} has no slab-parallel form")

    def _batchnorm(self, m, x):
        if not m.training:
            return F.batch_norm(x, m.running_mean, m.running_var, m.weight, m.bias, False, 0.0, m.eps)
        if isinstance(m, SynchronizedBatchNorm):
            return m(x)
        y, mean, var_unbiased = _SyncBNFn.apply(x, m.weight, m.bias, m.eps, self.group)
        with torch.no_grad():
            m.num_batches_tracked += 1
            mom = m.momentum if m.momentum is not None else 1.0 / float(m.num_batches_tracked)
            m.running_mean.mul_(1 - mom).add_(mom * mean)
            m.running_var.mul_(1 - mom).add_(mom * var_unbiased)
        return y


def allreduce_gradients(params, group=None) -> None:
    """Sum the per-slab shares of every parameter gradient (one flat all-reduce)."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return
    grads = []
    for p in params:
        if not p.requires_grad:
            continue
        if p.grad is None:          # a rank whose slab produced no output plane for this layer still takes part in the sum
            p.grad = torch.zeros_like(p)
        grads.append(p.grad)
    if not grads:
        return
    flat = torch.cat([g.reshape(-1) for g in grads])
    dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    o = 0
    for g in grads:
        g.copy_(flat[o:o + g.numel()].view_as(g))
        o += g.numel()


def halo_bytes(layers: Sequence[Tuple[int, int, int, int]], in_shape: Tuple[int, int, int], batch: int, world: int, elem_bytes: int = 2) -> List[int]:
    """Bytes a middle rank receives per layer for ``layers`` = (channels_in, k, s, p) along a chain starting at ``in_shape`` -- the
    planning number for deciding whether a client should be split (compare with the layer's compute time)."""
    d, h, w = in_shape
    out = []
    for cin, k, s, p in layers:
        part_in = DepthPartition(d, world)
        do = out_depth(d, k, s, p)
        part_out = DepthPartition(do, world)
        r = world // 2
        a, b = part_out.needed_input(r, k, s, p)
        i0, i1 = part_in.own(r)
        extra = max(0, i0 - max(a, 0)) + max(0, min(b, d) - i1)
        out.append(extra * h * w * cin * batch * elem_bytes)
        d, h, w = do, out_depth(h, k, s, p), out_depth(w, k, s, p)
    return out


class SpatialParallelModel(nn.Module):
    """A ``features`` + head model (the AlexNet3D family, Tiny3DCNN) or the 3-D ResNet trained by ALL ranks on the SAME batch, each rank holding a depth
    slab of the activations.  Parameters are replicated; after ``loss.backward()`` call :meth:`reduce_gradients` and step the
    optimizer on every rank -- the replicas stay identical because every rank then applies the same gradients (with dropout in
    the head, seed torch identically on all ranks so that they draw the same masks).

        spm = SpatialParallelModel(model)              # model.features: nn.Sequential, head: model.classifier / model.regressor
        loss = criterion(spm(x_full_batch), y); loss.backward(); spm.reduce_gradients(); optimizer.step()
    """

    def __init__(self, model: nn.Module, group=None):
        super().__init__()
        self.model = model
        self.group = group
        if hasattr(model, "features"):
            features = model.features
            head = getattr(model, "classifier", None) or getattr(model, "regressor", None)
        elif all(hasattr(model, a) for a in ("conv1", "bn1", "maxpool", "layer1", "layer2", "layer3", "avgpool", "fc", "fc2")):
            # the 3-D ResNet (models.volumetric.ResNet3D_l3): stem + three stages + average pool, then fc -> fc2
            features = nn.Sequential(model.conv1, model.bn1, model.relu, model.maxpool, model.layer1, model.layer2, model.layer3, model.avgpool)
            head = nn.Sequential(model.fc, model.fc2)
        else:
            features = head = None
        if features is None or head is None:
            raise ValueError("expected `features` + `classifier` / `regressor`, or the 3-D ResNet layout")
        self._features = features
        self.slab_features = SlabParallelFeatures(features, group)
        self.head = head

    def forward(self, x_full: torch.Tensor) -> torch.Tensor:
        world, rank = self.slab_features._world_rank()
        xs, part = scatter_depth(x_full, world, rank)
        f = self.slab_features(xs, part)
        return self.head(f.flatten(1))

    def reduce_gradients(self) -> None:
        """Feature-extractor gradients are per-slab shares -> summed over ranks.  The head saw the full feature map on every rank, so
        its gradients are already complete and identical."""
        allreduce_gradients(list(self._features.parameters()), self.group)
