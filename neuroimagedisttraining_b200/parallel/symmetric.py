"""NVLink symmetric memory for the fused aggregation kernels (``csrc/sym_agg.cu``).

``torch.distributed._symmetric_memory`` is used only to allocate identical buffers on every rank and
exchange their peer-mapped addresses (and the NVLS multicast address where the platform provides
one); the weighting + reduce-scatter + mask + all-gather itself is our kernel -- no NCCL call on the data path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from ..ops import native
from .context import DistContext

_declared = False
MODES = {"p2p": 0, "multimem": 1, "bulk": 2}


def _declare(lib) -> None:
    global _declared
    if _declared:
        return
    vp, i32, i64, u32, u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32, C.c_uint64
    for name in ("nidt_sym_max_ranks", "nidt_sym_flag_words", "nidt_sym_max_local", "nidt_sym_hist_words", "nidt_sym_bulk_max_ranks",
                 "nidt_sym_sal_scratch_bytes"):
        getattr(lib, name).argtypes = []
        getattr(lib, name).restype = C.c_int
    P = C.POINTER(vp)
    lib.nidt_sym_aggregate.argtypes = [P, P, P, vp, vp, i32, i32, P, C.POINTER(C.c_float), i32, vp, i64, i64, vp, u32, i32, i32, i32, vp]
    lib.nidt_sym_aggregate.restype = C.c_int
    lib.nidt_sym_saliency_topk_mask.argtypes = [P, P, P, vp, vp, P, P, i32, i32, i64, vp, vp, u64, vp, vp, u32, i32, vp]
    lib.nidt_sym_saliency_topk_mask.restype = C.c_int
    lib.nidt_sym_saliency_read_result.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_uint), vp]
    lib.nidt_sym_saliency_read_result.restype = C.c_int
    _declared = True


def _symm_alloc(n: int, dtype, device, group):
    """(tensor, handle, per-rank pointer array) of one symmetric allocation, zero-filled."""
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(n, dtype=dtype, device=device)
    t.zero_()
    h = symm.rendezvous(t, group=group)
    return t, h, [int(p) for p in h.buffer_ptrs]


class SymmetricAggregator:
    """Owns ``[partial | result]`` and a flag pad in symmetric memory.  :meth:`aggregate` launches the fused W1 kernel
    (local weighting + reduce-scatter + mask + all-gather); :meth:`saliency_topk_mask` the fused W2 kernel on buffers of its own.
    ``local_partial`` is where a caller that pre-reduces its clients writes the rank's contribution; ``local_result`` is where
    the new global model lands (:meth:`adopt` makes a slab live there, so no write-back pass is needed)."""

    def __init__(self, ctx: DistContext, numel: int, device, use_multimem: Optional[bool] = None):
        import torch.distributed._symmetric_memory as symm
        self.ctx = ctx
        self.device = torch.device(device)
        self.lib = native.load()
        _declare(self.lib)
        if ctx.world_size > self.lib.nidt_sym_max_ranks():
            raise ValueError(f"fused aggregation supports up to {self.lib.nidt_sym_max_ranks()} ranks")
        self.numel = int(numel)
        self.pad = (self.numel + 1023) // 1024 * 1024     # keep `result` 4 KiB aligned
        self.group = dist.group.WORLD
        try:
            symm.enable_symm_mem_for_group(self.group.group_name)
        except Exception:
            pass
        W = ctx.world_size
        self.buf, self.hdl, base = _symm_alloc(2 * self.pad, torch.float32, self.device, self.group)
        self.flags, self.fhdl, fptrs = _symm_alloc(self.lib.nidt_sym_flag_words(), torch.int32, self.device, self.group)
        torch.cuda.synchronize(self.device)
        dist.barrier(device_ids=[ctx.local_rank])
        self._partial = (C.c_void_p * W)(*base)
        self._result = (C.c_void_p * W)(*[b + 4 * self.pad for b in base])
        self._flags = (C.c_void_p * W)(*fptrs)
        mc = 0
        try:
            mc = int(self.hdl.multicast_ptr)
        except Exception:
            mc = 0
        self.mc_partial = mc or None
        self.mc_result = (mc + 4 * self.pad) if mc else None
        # NVLS multimem (in-switch reduction) vs P2P loads/stores: chosen by size unless forced (use_multimem=, NIDT_AGG_MULTIMEM=0/1
        # or NIDT_AGG_MODE=p2p|multimem|bulk).  Measured at 2/4/8 GPUs (profiles/agg_sweep_*): multimem wins up to 256 MB; above,
        # the P2P path staged through shared memory with bulk-async copies.
        env = os.environ.get("NIDT_AGG_MULTIMEM")
        if use_multimem is None and env in ("0", "1"):
            use_multimem = (env == "1")
        self._mm_forced = use_multimem
        self.has_multicast = bool(mc)
        self.multimem_max_bytes = int(os.environ.get("NIDT_AGG_MULTIMEM_MAX_BYTES", str(256 << 20)))
        self.bulk_min_bytes = int(os.environ.get("NIDT_AGG_BULK_MIN_BYTES", str(32 << 20)))
        self._align = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.local_partial = self.buf[:self.numel]
        self.local_result = self.buf[self.pad:self.pad + self.numel]
        self.epoch = 0
        self._adopted = False
        self._alt: Optional["SymmetricAggregator"] = None
        self._sal = None

    # ------------------------------------------------------------------ path selection
    def use_multimem_for(self, numel: int) -> bool:
        if not self.has_multicast:
            return False
        if self._mm_forced is not None:
            return bool(self._mm_forced)
        return 4 * int(numel) <= self.multimem_max_bytes

    @property
    def use_multimem(self) -> bool:
        return self.use_multimem_for(self.numel)

    @use_multimem.setter
    def use_multimem(self, force: Optional[bool]) -> None:
        """True / False forces the path for every size; None returns to the choice by size."""
        self._mm_forced = force

    def pick(self, numel: int) -> Tuple[str, float, int]:
        """(mode, share of each chunk sent through the in-switch path, CTAs or 0) for a slab of ``numel`` floats -- the fastest
        variant of the clock-stamped sweeps at 2, 4 and 8 GPUs (profiles/agg_sweep_n*_v3.jsonl): up to 4 ranks the shared-memory
        P2P path wins at every size (P2P moves 2(P-1)/P of the slab per direction, little more than the switch path's 1x); from
        5 ranks on, NVLS up to 256 MB and above that the shared-memory P2P path with half of every chunk through the switch."""
        env = os.environ.get("NIDT_AGG_MODE")
        bulk_ok = self.ctx.world_size <= self.lib.nidt_sym_bulk_max_ranks()
        if env in MODES and (env != "multimem" or self.has_multicast) and (env != "bulk" or bulk_ok):
            return env, 0.0, 0
        p2p = "bulk" if bulk_ok else "p2p"
        forced = self._mm_forced
        if not self.has_multicast or forced is False or (forced is None and self.ctx.world_size <= 4):
            return p2p, 0.0, 0
        if forced is True or 4 * int(numel) <= self.multimem_max_bytes:
            return "multimem", 0.0, (148 if 4 * int(numel) >= (64 << 20) else 0)
        return p2p, 0.5, 0

    def pick_mode(self, numel: int) -> str:
        mode, frac, _ = self.pick(numel)
        return mode + (f"+mm{frac:g}" if frac else "")

    def align_ranks(self) -> None:
        """Stream-ordered line-up of the ranks before a fused launch: a one-element NCCL all-reduce on the launching stream.
        The fused kernels spin on flags their peers write; ranks reach an aggregation with arbitrary skew (uneven client counts,
        a rank with no sampled client this round), and NCCL waits without a deadline where our device-side barrier is a bounded
        hang detector (``NIDT_SYM_TIMEOUT_S``).  After this call every rank's stream has passed the same point, so the kernels
        start within one collective's skew of each other.  No host synchronisation."""
        dist.all_reduce(self._align)

    # ------------------------------------------------------------------ W1
    def adopt(self, slab) -> None:
        """Make ``slab`` live in this aggregator's result buffer: an aggregation into it then needs no write-back pass (the
        all-gather of the fused kernel IS the write of the global model).  Aggregations into any other tensor go through a second
        set of buffers, so the adopted slab is never used as scratch."""
        assert slab.flat.numel() == self.numel
        self.local_result.copy_(slab.flat)
        slab.flat = self.local_result
        self._adopted = True

    def _alt_agg(self) -> "SymmetricAggregator":
        if self._alt is None:       # collective: every rank gets here in the same call
            self._alt = SymmetricAggregator(self.ctx, self.numel, self.device, self._mm_forced)
        return self._alt

    def aggregate(self, out: torch.Tensor, local: Sequence[Tuple[float, torch.Tensor]], mask: Optional[torch.Tensor], align: bool = True,
                  mode: Optional[str] = None, mm_frac: Optional[float] = None, ctas: int = 0, use_partial: bool = False) -> torch.Tensor:
        """``out = mask * sum_ranks sum_local weight * slab`` on every rank, one kernel launch per rank.  ``local``: this rank's
        (weight, flat tensor) pairs in client order; ``use_partial=True`` takes ``local_partial`` as the rank's finished
        contribution instead.  ``mode`` / ``mm_frac`` / ``ctas`` override the choice by size (sweeps)."""
        n = self.numel
        assert out.numel() == n and out.is_cuda
        in_place = out.data_ptr() == self.local_result.data_ptr()
        if self._adopted and not in_place:
            return self._alt_agg().aggregate(out, local, mask, align, mode, mm_frac, ctas, use_partial)
        local = list(local)
        k = len(local)
        if not use_partial:
            if k == 0:
                self.local_partial.zero_()
            elif k > self.lib.nidt_sym_max_local():
                from .. import ops
                ops.weighted_aggregate(self.local_partial, [t for _, t in local], [w for w, _ in local], None)
                k = 0
        else:
            k = 0
        srcs = (C.c_void_p * max(k, 1))(*[native.ptr(t) for _, t in local[:k]]) if k else None
        wts = (C.c_float * max(k, 1))(*[float(w) for w, _ in local[:k]]) if k else None
        if align:
            self.align_ranks()
        self.epoch += 1
        auto_mode, auto_frac, auto_ctas = self.pick(n)
        if mode is None:
            mode, ctas = auto_mode, (ctas or auto_ctas)
            frac = auto_frac if mm_frac is None else float(mm_frac)
        else:
            frac = 0.0 if mm_frac is None else float(mm_frac)
        if "NIDT_AGG_MM_FRAC" in os.environ and mm_frac is None:
            frac = float(os.environ["NIDT_AGG_MM_FRAC"])
        if not self.has_multicast:
            frac = 0.0
        n_mask = 0 if mask is None else min(mask.numel(), n) // 4 * 4
        native.check(self.lib.nidt_sym_aggregate(
            self._partial, self._result, self._flags, self.mc_partial, self.mc_result, self.ctx.rank, self.ctx.world_size,
            srcs, wts, k, native.ptr(mask), n_mask, n, None if in_place else native.ptr(out), self.epoch, MODES[mode],
            int(round(256 * frac)), int(ctas or int(os.environ.get("NIDT_AGG_CTAS", "0"))), native.stream_ptr(self.device)), "nidt_sym_aggregate")
        native.count_launches(1)
        return out

    def reduce_mask_broadcast(self, out: torch.Tensor, mask: Optional[torch.Tensor], align: bool = True, **kw) -> torch.Tensor:
        """``out = mask * sum_ranks local_partial`` (the rank's contribution was written to ``local_partial`` by the caller)."""
        return self.aggregate(out, [], mask, align=align, use_partial=True, **kw)

    # ------------------------------------------------------------------ W2
    def _sal_buffers(self, n: int):
        if self._sal is None or self._sal["n"] != n:     # collective
            W = self.ctx.world_size
            pad = (n + 1023) // 1024 * 1024
            buf, hdl, base = _symm_alloc(2 * pad, torch.float32, self.device, self.group)
            flags, fh, fptrs = _symm_alloc(self.lib.nidt_sym_flag_words(), torch.int32, self.device, self.group)
            hist, hh, hptrs = _symm_alloc(self.lib.nidt_sym_hist_words(), torch.int32, self.device, self.group)
            mask, mh, mptrs = _symm_alloc(pad, torch.float32, self.device, self.group)
            torch.cuda.synchronize(self.device)
            dist.barrier(device_ids=[self.ctx.local_rank])
            mc = 0
            try:
                mc = int(hdl.multicast_ptr)
            except Exception:
                mc = 0
            self._sal = {"n": n, "pad": pad, "keep": (buf, hdl, flags, fh, hist, hh, mask, mh), "partial_t": buf[:n], "mask_t": mask[:n],
                         "partial": (C.c_void_p * W)(*base), "result": (C.c_void_p * W)(*[b + 4 * pad for b in base]),
                         "flags": (C.c_void_p * W)(*fptrs), "hist": (C.c_void_p * W)(*hptrs), "mask": (C.c_void_p * W)(*mptrs),
                         "mc_partial": mc or None, "mc_result": (mc + 4 * pad) if mc else None, "epoch": 0,
                         "scratch": torch.zeros(self.lib.nidt_sym_sal_scratch_bytes(), dtype=torch.uint8, device=self.device)}
        return self._sal

    def saliency_topk_mask(self, scores: torch.Tensor, layout, keep_ratio: float, mask_out: torch.Tensor, flags: torch.Tensor,
                           read: bool = True, align: bool = True):
        """Fused W2: ``scores`` (this rank's partial) is reduce-scattered over NVLink, every rank histograms its shard, the
        shard histograms are exchanged and summed, the global k-th largest score gives the threshold and every rank emits its
        shard of the mask into all ranks' buffers -- one launch per rank.  ``scores`` receives the reduced values.
        Returns ``(mask_out, info)``."""
        n = layout.param_numel
        k = int(layout.prunable_numel() * keep_ratio)
        s = self._sal_buffers(n)
        s["partial_t"].copy_(scores[:n])
        if align:
            self.align_ranks()
        s["epoch"] += 1
        mm = bool(s["mc_partial"]) and self.pick(n)[0] == "multimem"
        native.check(self.lib.nidt_sym_saliency_topk_mask(
            s["partial"], s["result"], s["flags"], s["mc_partial"], s["mc_result"], s["hist"], s["mask"], self.ctx.rank, self.ctx.world_size, n,
            native.ptr(flags), native.ptr(mask_out), k, native.ptr(s["scratch"]), native.ptr(scores), s["epoch"], int(mm),
            native.stream_ptr(self.device)), "nidt_sym_saliency_topk_mask")
        native.count_launches(1)
        if not read:            # timing loops: no host read-back
            return mask_out, {"k": k}
        thr, norm, kept = C.c_float(), C.c_float(), C.c_uint()
        native.check(self.lib.nidt_sym_saliency_read_result(native.ptr(s["scratch"]), C.byref(thr), C.byref(norm), C.byref(kept),
                                                            native.stream_ptr(self.device)), "nidt_sym_saliency_read_result")
        return mask_out, {"threshold": thr.value, "norm": norm.value, "kept": kept.value, "k": k}
