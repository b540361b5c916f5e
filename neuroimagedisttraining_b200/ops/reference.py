"""Plain-PyTorch fp32 restatements of every custom op.

These are (a) the CPU path, so the whole framework runs and is testable without a GPU,
and (b) the numerics oracle the ``-m gpu`` tests compare each sm_100a kernel against.
They are *not* the GPU product path: on a CUDA tensor ``ops.*`` always dispatches to the
native library and raises if it is missing (``ops/native.py``).
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch

CLIP_EPS = 1e-6  # torch.nn.utils.clip_grad_norm_ uses max_norm / (total_norm + 1e-6)


# ----------------------------------------------------------------------------- optimizer
def masked_sgd_step(w: torch.Tensor, g: torch.Tensor, mom: Optional[torch.Tensor], mask: Optional[torch.Tensor],
                    lr: float, momentum: float, wd: float, clip_norm: float, first_step: bool) -> torch.Tensor:
    """clip_grad_norm_(clip_norm) -> SGD(momentum, weight_decay) -> ``w *= mask``; in place on ``w``/``mom``.

    Reference: ``sailentgrads/my_model_trainer.py:209,224-231`` (K8+K9+K10 of SURVEY §2.5).
    Returns the total gradient norm (0-d tensor).  ``clip_norm <= 0`` disables clipping."""
    total = torch.linalg.vector_norm(g.double()).float()
    d = g
    if clip_norm > 0:
        coef = torch.clamp(clip_norm / (total + CLIP_EPS), max=1.0)
        d = g * coef
    if wd != 0:
        d = d + wd * w
    if momentum != 0:
        assert mom is not None
        if first_step:
            mom.copy_(d)
        else:
            mom.mul_(momentum).add_(d)
        d = mom
    w.add_(d, alpha=-lr)
    if mask is not None:
        w.mul_(mask)
    return total


def prox_sgd_pull(w: torch.Tensor, w_ref: torch.Tensor, lr: float, lamda: float) -> None:
    """Ditto's proximal pull ``w -= lr * lamda * (w - w_global)`` (``ditto/my_model_trainer.py:63-64``)."""
    w.add_(w - w_ref, alpha=-lr * lamda)


# --------------------------------------------------------------------------- aggregation
def weighted_aggregate(out: torch.Tensor, xs: Sequence[torch.Tensor], weights: Sequence[float],
                       mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = sum_i weights[i] * xs[i]`` accumulated in list order with separate multiply and
    add (the reference's ``w_global[k] = x0*w0; w_global[k] += xi*wi`` loop,
    ``sailentgrads_api.py:212-227`` / ``fedavg_api.py:102-117``), then an optional mask multiply."""
    assert len(xs) == len(weights) and len(xs) > 0
    n = out.numel()
    acc = xs[0][:n] * weights[0]
    for x, wt in zip(xs[1:], weights[1:]):
        acc = acc + x[:n] * wt
    if mask is not None:
        m = mask.numel()
        acc[:m] = acc[:m] * mask
    out.copy_(acc)
    return out


def mask_weighted_aggregate(out: torch.Tensor, xs: Sequence[torch.Tensor], masks: Sequence[torch.Tensor]) -> torch.Tensor:
    """Mask-count-weighted mean: ``out = sum_i x_i*m_i / max(sum_i m_i, 1)`` over the masked region,
    plain mean elsewhere (DisPFL ``_aggregate_func`` ``dispfl_api.py:222-240``, Sub-FedAvg ``subavg_api.py:123-139``)."""
    n = out.numel()
    m = masks[0].numel()
    cnt = torch.zeros(m, dtype=torch.float32, device=out.device)
    acc = torch.zeros(n, dtype=torch.float32, device=out.device)
    for x, mk in zip(xs, masks):
        cnt += mk
        acc[:m] += x[:m] * mk
        acc[m:] += x[m:n]
    acc[:m] = torch.where(cnt > 0, acc[:m] / cnt.clamp(min=1.0), torch.zeros_like(cnt))
    acc[m:] /= len(xs)
    out.copy_(acc)
    return out


# ------------------------------------------------------------------------------ saliency
def saliency_accumulate(scores: torch.Tensor, g: torch.Tensor, w: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    """``scores += scale * |g * w|`` -- SNIP's ``|dL/dm|`` for a multiplicative mask at m=1
    (``snip.py:9-16,60-71``: forward uses ``w*m`` so ``dL/dm = dL/dw_eff * w``)."""
    scores.add_((g * w).abs(), alpha=scale)
    return scores


def global_topk_mask(scores: torch.Tensor, segments: Sequence[Tuple[int, int]], keep_ratio: float,
                     mask_out: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """One shared mask from a global top-k over the prunable segments of a score slab.

    Semantics of ``snip.get_mask_from_grads`` (``snip.py:80-116``): concatenate the scored tensors,
    divide by their sum, ``k = int(N*keep_ratio)``, threshold = k-th largest, keep ``s/sum >= thr``
    (ties kept); every non-scored parameter gets ones.  ``segments`` are (offset, numel) pairs.
    Returns (mask_out, threshold)."""
    parts = [scores[o:o + n] for o, n in segments]
    allv = torch.cat(parts)
    norm = allv.sum()
    allv = allv / norm
    k = int(allv.numel() * keep_ratio)
    if k <= 0:
        thr = torch.tensor(float("inf"), device=scores.device)
    else:
        thr = torch.topk(allv, k, sorted=True)[0][-1]
    mask_out.fill_(1.0)
    for o, n in segments:
        mask_out[o:o + n] = ((scores[o:o + n] / norm) >= thr).float()
    return mask_out, thr


def zero_padding_(flat: torch.Tensor, layout) -> torch.Tensor:
    """Re-zero the alignment gaps of a slab-shaped tensor (masks are 0 there)."""
    prev = 0
    for e in layout.entries.values():
        if e.offset >= flat.numel():
            break
        if e.offset > prev:
            flat[prev:e.offset] = 0
        prev = e.offset + e.numel
    if prev < flat.numel():
        flat[prev:] = 0
    return flat


# ----------------------------------------------------------------------------- statistics
def count_nonzero(flat: torch.Tensor) -> torch.Tensor:
    return torch.count_nonzero(flat)


def hamming_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Number of differing mask entries (``DisPFL/slim_util.py:14-19``)."""
    return ((a != 0) ^ (b != 0)).sum()


def squared_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``sum (a-b)^2`` (``model_difference``, ``DisPFL/slim_util.py:10-12``)."""
    return torch.sum((a - b) ** 2)


# ------------------------------------------------------------------ DisPFL prune / regrow
def fire_mask(w: torch.Tensor, mask: torch.Tensor, drop_ratio: float) -> Tuple[torch.Tensor, int]:
    """Per-tensor magnitude prune: drop the ``ceil(drop_ratio * nnz(mask))`` smallest-|w| active
    weights (``DisPFL/client.py:71-82``).  Returns (new mask, number removed)."""
    num_nonzero = int(mask.sum().item())
    num_remove = math.ceil(drop_ratio * num_nonzero)
    flat_w = torch.where(mask.flatten() > 0, w.flatten().abs(), torch.full_like(w.flatten(), 100000.0))
    _, idx = torch.sort(flat_w)
    new_mask = mask.clone().flatten()
    new_mask[idx[:num_remove]] = 0
    return new_mask.view_as(mask), num_remove


def regrow_mask(mask: torch.Tensor, num_remove: int, grad: Optional[torch.Tensor] = None,
                generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Regrow ``num_remove`` inactive weights: largest |grad| (``dis_gradient_check`` off) or
    uniformly at random (``DisPFL/client.py:86-99``)."""
    new_mask = mask.clone().flatten()
    if num_remove <= 0:
        return new_mask.view_as(mask)
    if grad is not None:
        score = torch.where(new_mask == 0, grad.flatten().abs(), torch.full_like(new_mask, -100000.0))
        _, idx = torch.sort(score, descending=True)
        new_mask[idx[:num_remove]] = 1
    else:
        zero_idx = torch.nonzero(new_mask == 0).flatten()
        perm = torch.randperm(zero_idx.numel(), generator=generator, device="cpu").to(zero_idx.device)
        new_mask[zero_idx[perm[:num_remove]]] = 1
    return new_mask.view_as(mask)


# ------------------------------------------------------------------------------ eval head
def bce_eval_counters(logits: torch.Tensor, target: torch.Tensor, double_sigmoid: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """(correct count, summed loss) for one eval batch.  ``double_sigmoid`` reproduces the
    reference's ``BCEWithLogits(sigmoid(logit))`` (``my_model_trainer.py:263-264``)."""
    p = torch.sigmoid(logits.float()).flatten()
    t = target.float().flatten()
    z = p if double_sigmoid else logits.float().flatten()
    loss = torch.nn.functional.binary_cross_entropy_with_logits(z, t, reduction="sum")
    correct = ((p >= 0.5).float() == t).float().sum()
    return correct, loss


def conv1_wgrad_from_pooled(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], gamma: torch.Tensor, beta: torch.Tensor,
                            da: torch.Tensor, eps: float = 1e-5) -> Tuple[torch.Tensor, torch.Tensor]:
    """PyTorch statement of ``csrc/conv_in_alg.cu``: the weight gradient of ``Conv3d(1, C, 5, stride 2) -> BatchNorm3d (batch
    statistics) -> ReLU -> MaxPool3d(3, 3)`` from the POOLED gradient ``da`` [N, C, Dp, Hp, Wp] without ever forming dY:

        dW[c, t] = s_c * ( S[c, t] - c1_c * sx[t] - c2_c * invstd_c * ((W G)[c, t] + (b_c - mu_c) * sx[t]) )

    ``G = P^T P`` and ``sx = P^T 1`` over the patch matrix ``P`` [positions, 125] of the input (they depend on ``x`` only -- the kernels
    cache them per sample), ``S`` sums the winning position's patch of every pooling window times its ``da``, ``c1 = mean(dz)``,
    ``c2 = mean(dz * yhat)`` are the two BatchNorm-backward means over ALL positions, ``s = gamma * invstd``.  Returns ``(dW, dgamma)``;
    the conv bias gradient is identically zero.  The forward is recomputed here (this is an oracle, not a fast path).
    ref: autograd through salient_models.py:147-150."""
    N, _, D, H, W_ = x.shape
    C = w.shape[0]
    patches = x[:, 0].unfold(1, 5, 2).unfold(2, 5, 2).unfold(3, 5, 2)                 # [N, Do, Ho, Wo, 5, 5, 5]
    Do, Ho, Wo = patches.shape[1:4]
    P = patches.reshape(N * Do * Ho * Wo, 125)
    y = P @ w.reshape(C, 125).t()
    if b is not None:
        y = y + b
    mu, var = y.mean(0), y.var(0, unbiased=False)
    invstd = (var + eps).rsqrt()
    yhat = (y - mu) * invstd
    z = torch.relu(yhat * gamma + beta).reshape(N, Do, Ho, Wo, C).permute(0, 4, 1, 2, 3)
    pooled, idx = torch.nn.functional.max_pool3d(z, 3, 3, return_indices=True)       # idx: flat position inside (Do, Ho, Wo)
    gate = (pooled > 0).to(da.dtype)
    n_i = torch.arange(N, device=x.device).view(N, 1, 1, 1, 1).expand_as(idx)
    pos = (n_i * (Do * Ho * Wo) + idx).reshape(-1)                                     # row of P / y of every window's winner
    ch = torch.arange(C, device=x.device).view(1, C, 1, 1, 1).expand_as(idx).reshape(-1)
    g = (da * gate).reshape(-1)
    count = float(P.shape[0])
    S = torch.zeros(C, 125, dtype=x.dtype, device=x.device)
    S.index_add_(0, ch, g.unsqueeze(1) * P[pos])                                       # one patch per (window, channel)
    c1 = torch.zeros(C, dtype=x.dtype, device=x.device).index_add_(0, ch, g) / count
    c2 = torch.zeros(C, dtype=x.dtype, device=x.device).index_add_(0, ch, g * yhat[pos, ch]) / count
    G, sx = P.t() @ P, P.sum(0)
    bb = b if b is not None else torch.zeros(C, dtype=x.dtype, device=x.device)
    inner = w.reshape(C, 125) @ G + (bb - mu).unsqueeze(1) * sx.unsqueeze(0)
    dW = (gamma * invstd).unsqueeze(1) * (S - c1.unsqueeze(1) * sx.unsqueeze(0) - (c2 * invstd).unsqueeze(1) * inner)
    return dW.reshape(C, 1, 5, 5, 5), c2 * count
