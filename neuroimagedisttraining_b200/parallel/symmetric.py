"""NVLink symmetric memory for the fused aggregation kernel (``csrc/sym_agg.cu``).

``torch.distributed._symmetric_memory`` is used only to allocate identical buffers on every rank and
exchange their peer-mapped addresses (and the NVLS multicast address where the platform provides
one); the reduce-scatter + mask + all-gather itself is our kernel -- no NCCL call on this path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
import torch.distributed as dist

from ..ops import native
from .context import DistContext

_declared = False


def _declare(lib) -> None:
    global _declared
    if _declared:
        return
    vp, i32, i64, u32 = C.c_void_p, C.c_int, C.c_int64, C.c_uint32
    lib.nidt_sym_max_ranks.argtypes = []
    lib.nidt_sym_max_ranks.restype = C.c_int
    lib.nidt_sym_flag_words.argtypes = []
    lib.nidt_sym_flag_words.restype = C.c_int
    lib.nidt_sym_reduce_mask_broadcast.argtypes = [C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), vp, vp, i32, i32, vp, i64, i64, vp, u32, i32, vp]
    lib.nidt_sym_reduce_mask_broadcast.restype = C.c_int
    lib.nidt_sym_saliency_topk_mask.argtypes = [C.POINTER(vp), C.POINTER(vp), C.POINTER(vp), vp, vp, i32, i32, i64, vp, vp, C.c_uint64, vp, vp, u32, i32, vp]
    lib.nidt_sym_saliency_topk_mask.restype = C.c_int
    lib.nidt_topk_scratch_bytes.restype = C.c_int
    _declared = True


class SymmetricAggregator:
    """Owns ``[partial | result]`` and a flag pad in symmetric memory; ``reduce_mask_broadcast`` launches
    the fused kernel.  ``local_partial`` is where the rank writes its own weighted partial sum."""

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
        group = dist.group.WORLD
        try:
            symm.enable_symm_mem_for_group(group.group_name)
        except Exception:
            pass
        self.buf = symm.empty(2 * self.pad, dtype=torch.float32, device=self.device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group=group)
        self.flags = symm.empty(self.lib.nidt_sym_flag_words(), dtype=torch.int32, device=self.device)
        self.flags.zero_()
        self.fhdl = symm.rendezvous(self.flags, group=group)
        torch.cuda.synchronize(self.device)
        dist.barrier(device_ids=[ctx.local_rank])
        W = ctx.world_size
        base = [int(p) for p in self.hdl.buffer_ptrs]
        self._partial = (C.c_void_p * W)(*base)
        self._result = (C.c_void_p * W)(*[b + 4 * self.pad for b in base])
        self._flags = (C.c_void_p * W)(*[int(p) for p in self.fhdl.buffer_ptrs])
        mc = 0
        try:
            mc = int(self.hdl.multicast_ptr)
        except Exception:
            mc = 0
        self.mc_partial = mc or None
        self.mc_result = (mc + 4 * self.pad) if mc else None
        # NVLS multimem (in-switch reduction) vs plain P2P loads/stores: chosen by size unless forced (use_multimem= or
        # NIDT_AGG_MULTIMEM=0/1).  Measured at 2/4/8 GPUs (profiles/agg_sweep_*): multimem wins up to 256 MB, P2P above.
        env = os.environ.get("NIDT_AGG_MULTIMEM")
        if use_multimem is None and env in ("0", "1"):
            use_multimem = (env == "1")
        self._mm_forced = use_multimem
        self.has_multicast = bool(mc)
        self.multimem_max_bytes = int(os.environ.get("NIDT_AGG_MULTIMEM_MAX_BYTES", str(256 << 20)))
        self._align = torch.zeros(1, dtype=torch.float32, device=self.device)
        self.local_partial = self.buf[:self.numel]
        self.local_result = self.buf[self.pad:self.pad + self.numel]
        self.epoch = 0

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

    def align_ranks(self) -> None:
        """Stream-ordered line-up of the ranks before a fused launch: a one-element NCCL all-reduce on the launching stream.
        The fused kernels spin on flags their peers write; ranks reach an aggregation with arbitrary skew (uneven client counts,
        a rank with no sampled client this round), and NCCL waits without a deadline where our device-side barrier is a bounded
        hang detector (``NIDT_SYM_TIMEOUT_S``).  After this call every rank's stream has passed the same point, so the kernels
        start within one collective's skew of each other.  No host synchronisation."""
        dist.all_reduce(self._align)

    def reduce_mask_broadcast(self, out: torch.Tensor, mask: Optional[torch.Tensor], align: bool = True) -> torch.Tensor:
        """``out = mask * sum_ranks partial`` on every rank; one kernel launch per rank (``align``: see :meth:`align_ranks`;
        benchmarks that time the kernel alone line the ranks up themselves)."""
        n = self.numel
        assert out.numel() == n and out.is_cuda
        if align:
            self.align_ranks()
        self.epoch += 1
        n_mask = 0 if mask is None else min(mask.numel(), n)
        native.check(self.lib.nidt_sym_reduce_mask_broadcast(
            self._partial, self._result, self._flags, self.mc_partial, self.mc_result, self.ctx.rank, self.ctx.world_size,
            native.ptr(mask), n_mask, n, native.ptr(out), self.epoch, int(self.use_multimem), native.stream_ptr(self.device)),
            "nidt_sym_reduce_mask_broadcast")
        native.count_launches(1)
        return out

    def saliency_topk_mask(self, scores: torch.Tensor, layout, keep_ratio: float, mask_out: torch.Tensor, flags: torch.Tensor):
        """Fused W2: ``scores`` (this rank's partial) is all-reduced over NVLink, the global k-th largest
        normalised score is found and ``mask_out`` is emitted, in one launch per rank.  ``scores`` receives
        the reduced values.  Returns ``(mask_out, info)``."""
        n = layout.param_numel
        k = int(layout.prunable_numel() * keep_ratio)
        self.local_partial[:n].copy_(scores[:n])
        self.align_ranks()
        if not hasattr(self, "_topk_scratch"):
            self._topk_scratch = torch.zeros(self.lib.nidt_topk_scratch_bytes(), dtype=torch.uint8, device=self.device)
        self.epoch += 1
        native.check(self.lib.nidt_sym_saliency_topk_mask(
            self._partial, self._result, self._flags, self.mc_partial, self.mc_result, self.ctx.rank, self.ctx.world_size, n,
            native.ptr(flags), native.ptr(mask_out), k, native.ptr(self._topk_scratch), native.ptr(scores), self.epoch,
            int(self.use_multimem_for(n)), native.stream_ptr(self.device)), "nidt_sym_saliency_topk_mask")
        native.count_launches(1)
        thr, norm, kept = C.c_float(), C.c_float(), C.c_uint()
        native.check(self.lib.nidt_topk_read_result(native.ptr(self._topk_scratch), C.byref(thr), C.byref(norm), C.byref(kept),
                                                    native.stream_ptr(self.device)), "nidt_topk_read_result")
        return mask_out, {"threshold": thr.value, "norm": norm.value, "kept": kept.value, "k": k}
