"""Flat-slab ops (one launch over a whole model instead of per-tensor Python loops)."""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from . import native, reference

_scratch: Dict[Tuple[str, int], torch.Tensor] = {}


def _get_scratch(kind: str, device: torch.device, nbytes: int) -> torch.Tensor:
    key = (kind, device.index if device.index is not None else torch.cuda.current_device())
    t = _scratch.get(key)
    if t is None or t.numel() < nbytes:
        t = torch.zeros(nbytes, dtype=torch.uint8, device=device)
        _scratch[key] = t
    return t


def _norm_scratch(device) -> torch.Tensor:
    lib = native.load()
    return _get_scratch("norm", device, 4 * lib.nidt_norm_scratch_floats())


def prunable_flags(layout, device=None) -> torch.Tensor:
    """uint8 per element of the trainable region: 0 padding, 1 real (never pruned), 2 prunable weight."""
    f = torch.zeros(layout.param_numel, dtype=torch.uint8)
    for name in layout.param_names:
        e = layout[name]
        f[e.offset:e.offset + e.numel] = 2 if e.prunable else 1
    return f.to(device) if device is not None else f


def prunable_segments(layout) -> List[Tuple[int, int]]:
    return [(layout[n].offset, layout[n].numel) for n in layout.prunable_names]


# ------------------------------------------------------------------------------ optimizer
def masked_sgd_step(w: torch.Tensor, g: torch.Tensor, mom: Optional[torch.Tensor], mask: Optional[torch.Tensor],
                    lr: float, momentum: float = 0.0, wd: float = 0.0, clip_norm: float = 10.0,
                    first_step: bool = False, norm_out: Optional[torch.Tensor] = None) -> Optional[torch.Tensor]:
    """Fused clip-by-global-norm + SGD(momentum, weight decay) + mask multiply over a flat slab.
    Replaces K8+K9+K10 (``sailentgrads/my_model_trainer.py:209,224-231``): 3 x 24 small launches -> 2."""
    if not w.is_cuda:
        t = reference.masked_sgd_step(w, g, mom, mask, lr, momentum, wd, clip_norm, first_step)
        if norm_out is not None:
            norm_out.copy_(t)
        return norm_out
    lib = native.load()
    n = w.numel()
    if mask is not None:
        assert mask.numel() >= n
    native.check(lib.nidt_masked_sgd_step(native.ptr(w), native.ptr(g), native.ptr(mom) if momentum != 0 else None,
                                          native.ptr(mask), n, lr, momentum, wd, clip_norm, int(first_step),
                                          native.ptr(_norm_scratch(w.device)), native.ptr(norm_out),
                                          native.stream_ptr(w.device)), "nidt_masked_sgd_step")
    native.count_launches(2)
    return norm_out


def grad_norm(g: torch.Tensor) -> torch.Tensor:
    if not g.is_cuda:
        return torch.linalg.vector_norm(g)
    lib = native.load()
    out = torch.empty((), dtype=torch.float32, device=g.device)
    native.check(lib.nidt_grad_norm(native.ptr(g), g.numel(), native.ptr(_norm_scratch(g.device)), native.ptr(out),
                                    native.stream_ptr(g.device)), "nidt_grad_norm")
    native.count_launches(2)
    return out


def prox_pull(w: torch.Tensor, w_ref: torch.Tensor, lr: float, lamda: float) -> None:
    if not w.is_cuda:
        return reference.prox_sgd_pull(w, w_ref, lr, lamda)
    lib = native.load()
    native.check(lib.nidt_prox_pull(native.ptr(w), native.ptr(w_ref), w.numel(), lr * lamda, native.stream_ptr(w.device)), "nidt_prox_pull")
    native.count_launches(1)


# ---------------------------------------------------------------------------- aggregation
def weighted_aggregate(out: torch.Tensor, xs: Sequence[torch.Tensor], weights: Sequence[float],
                       mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``out = sum_i weights[i]*xs[i]`` (client order, separate mul/add => bit-identical to the
    reference loop) with an optional fused mask multiply over the trainable prefix."""
    if not out.is_cuda:
        return reference.weighted_aggregate(out, xs, weights, mask)
    lib = native.load()
    n = out.numel()
    k = len(xs)
    assert k == len(weights) and k > 0
    srcs = (C.c_void_p * k)(*[native.ptr(x) for x in xs])
    wts = (C.c_float * k)(*[float(w) for w in weights])
    n_mask = 0 if mask is None else min(mask.numel(), n)
    native.check(lib.nidt_weighted_aggregate(native.ptr(out), srcs, wts, k, native.ptr(mask), n_mask, n,
                                             native.stream_ptr(out.device)), "nidt_weighted_aggregate")
    native.count_launches((k + lib.nidt_max_agg_sources() - 1) // lib.nidt_max_agg_sources())
    return out


# ------------------------------------------------------------------------------- saliency
def saliency_accumulate(scores: torch.Tensor, g: torch.Tensor, w: torch.Tensor, scale: float = 1.0) -> torch.Tensor:
    if not scores.is_cuda:
        return reference.saliency_accumulate(scores, g, w, scale)
    lib = native.load()
    native.check(lib.nidt_saliency_accumulate(native.ptr(scores), native.ptr(g), native.ptr(w), scores.numel(), scale,
                                              native.stream_ptr(scores.device)), "nidt_saliency_accumulate")
    native.count_launches(1)
    return scores


def global_topk_mask(scores: torch.Tensor, layout, keep_ratio: float, mask_out: torch.Tensor,
                     flags: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, dict]:
    """Global top-k over the prunable weights -> one shared {0,1} mask (``snip.py:80-116``).
    On GPU: one cooperative kernel (sum, 4 radix-select passes, mask emission)."""
    n_prunable = layout.prunable_numel()
    k = int(n_prunable * keep_ratio)
    if not scores.is_cuda:
        _, thr = reference.global_topk_mask(scores, prunable_segments(layout), keep_ratio, mask_out)
        reference.zero_padding_(mask_out, layout)
        return mask_out, {"threshold": float(thr), "k": k}
    lib = native.load()
    if flags is None:
        flags = prunable_flags(layout, scores.device)
    scratch = _get_scratch("topk", scores.device, lib.nidt_topk_scratch_bytes())
    n = layout.param_numel
    native.check(lib.nidt_global_topk_mask(native.ptr(scores), native.ptr(flags), native.ptr(mask_out), n, k,
                                           native.ptr(scratch), native.stream_ptr(scores.device)), "nidt_global_topk_mask")
    native.count_launches(1)
    thr, norm, kept = C.c_float(), C.c_float(), C.c_uint()
    native.check(lib.nidt_topk_read_result(native.ptr(scratch), C.byref(thr), C.byref(norm), C.byref(kept),
                                           native.stream_ptr(scores.device)), "nidt_topk_read_result")
    return mask_out, {"threshold": thr.value, "norm": norm.value, "kept": kept.value, "k": k}


# ----------------------------------------------------------------------------- statistics
def count_nonzero(flat: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Device-side nnz (no host sync): replaces the per-tensor ``count_nonzero().item()`` loops
    (``my_model_trainer.py:144-158``, ``model_trainer.py:49-53``).  ``out`` (int64, 0-d) is added to."""
    if not flat.is_cuda:
        r = torch.count_nonzero(flat)
        if out is not None:
            out.add_(r)
            return out
        return r
    lib = native.load()
    if out is None:
        out = torch.zeros((), dtype=torch.int64, device=flat.device)
    native.check(lib.nidt_count_nonzero(native.ptr(flat), flat.numel(), native.ptr(out), native.stream_ptr(flat.device)), "nidt_count_nonzero")
    native.count_launches(1)
    return out


def hamming_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    if not a.is_cuda:
        return reference.hamming_distance(a, b)
    lib = native.load()
    out = torch.zeros((), dtype=torch.int64, device=a.device)
    native.check(lib.nidt_hamming(native.ptr(a), native.ptr(b), a.numel(), native.ptr(out), native.stream_ptr(a.device)), "nidt_hamming")
    native.count_launches(1)
    return out


def squared_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    if not a.is_cuda:
        return reference.squared_distance(a, b)
    lib = native.load()
    out = torch.empty((), dtype=torch.float32, device=a.device)
    native.check(lib.nidt_sq_dist(native.ptr(a), native.ptr(b), a.numel(), native.ptr(_norm_scratch(a.device)), native.ptr(out),
                                  native.stream_ptr(a.device)), "nidt_sq_dist")
    native.count_launches(2)
    return out


def gather_rows_u8(src: torch.Tensor, idx: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``src[idx]`` for a contiguous uint8 tensor ``[N, ...]`` and an int64 device index vector: one native launch that moves
    16 bytes per thread regardless of the (odd) row length; ``torch.index_select`` on CPU."""
    if not src.is_cuda:
        return src.index_select(0, idx)
    lib = native.load()
    rows = int(idx.numel())
    row_bytes = int(src[0].numel())
    out = torch.empty((rows,) + tuple(src.shape[1:]), dtype=torch.uint8, device=src.device) if out is None else out
    lib.nidt_gather_rows_u8.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p]
    lib.nidt_gather_rows_u8.restype = C.c_int
    native.check(lib.nidt_gather_rows_u8(native.ptr(src), native.ptr(idx), native.ptr(out), rows, row_bytes, native.stream_ptr(src.device)),
                 "nidt_gather_rows_u8")
    native.count_launches(1)
    return out


def mask_select_k(values: torch.Tensor, mask: torch.Tensor, k: int, mode: str) -> torch.Tensor:
    """DisPFL prune / regrow on one tensor, in place on ``mask`` (fp32 0/1, contiguous): ``mode="fire"`` clears the ``k`` active
    entries with the smallest ``|values|``, ``mode="regrow"`` sets the ``k`` inactive entries with the largest ``|values|``.
    CUDA: one cooperative radix-select kernel; CPU: the sort-based restatement."""
    assert mode in ("fire", "regrow") and mask.is_contiguous() and values.numel() == mask.numel()
    k = int(k)
    if not mask.is_cuda:
        flat, v = mask.view(-1), values.reshape(-1).abs()
        if k > 0:
            if mode == "fire":
                score = torch.where(flat > 0, v, torch.full_like(v, float("inf")))
                idx = torch.sort(score, stable=True)[1][:min(k, int((flat > 0).sum()))]
                flat[idx] = 0
            else:
                score = torch.where(flat == 0, v, torch.full_like(v, -1.0))
                idx = torch.sort(score, descending=True, stable=True)[1][:min(k, int((flat == 0).sum()))]
                flat[idx] = 1
        return mask
    lib = native.load()
    lib.nidt_mask_select_scratch_bytes.restype = C.c_int
    lib.nidt_mask_select_k.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int, C.c_void_p, C.c_void_p]
    lib.nidt_mask_select_k.restype = C.c_int
    scratch = _get_scratch("mask_select", mask.device, lib.nidt_mask_select_scratch_bytes())
    v = values.contiguous().float()
    native.check(lib.nidt_mask_select_k(native.ptr(v), native.ptr(mask), mask.numel(), k, 0 if mode == "fire" else 1, native.ptr(scratch),
                                        native.stream_ptr(mask.device)), "nidt_mask_select_k")
    native.count_launches(1)
    return mask
