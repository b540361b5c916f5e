"""sm_100a compute engine for plan-based 3-D CNNs (AlexNet3D family).

Walks ``model.plan`` / ``model.head_spec`` and launches the hand-written kernels directly on the
parameter slab -- no autograd, no cuDNN/cuBLAS, no intermediate fp32 tensors:

  conv_in (Cin=1, tcgen05, im2col built in smem from the 8-bit volume) | conv3d_fprop_tc (tcgen05+TMA)
  -> bn_finalize -> bn_relu_pool_fwd  ...  head_fwd_bwd  ...  bn_relu_pool_bwd -> conv3d_wgrad_tc
  -> conv3d_fprop_tc as dgrad ... -> masked_sgd_step

Activations are channels-last (NDHWC) bf16; master weights, BN statistics, gradients and the
optimizer state are fp32 in the slab.  Interface-compatible with :class:`fl.engine.TorchEngine`.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Tuple

import torch
import torch.nn as nn

from ..models.volumetric import ConvBlockSpec, HeadSpec, trace_shapes
from ..runtime.slab import ParamSlab, SlabLayout
from . import flat as flat_ops
from . import native

_declared = False


def _declare(lib) -> None:
    global _declared
    if _declared:
        return
    i32, i64, f32, f64, vp, u64 = C.c_int, C.c_int64, C.c_float, C.c_double, C.c_void_p, C.c_uint64
    sigs = {
        "nidt_bn_bwd_grid": [],
        "nidt_bn_finalize": [vp, i32, i32, f64, vp, vp, vp, vp, vp, f32, f32, vp, vp, vp, vp, vp],
        "nidt_bn_eval_prepare": [i32, vp, vp, vp, vp, f32, vp, vp, vp],
        "nidt_bn_relu_pool_fwd": [vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp, i32, vp],
        "nidt_bn_relu_pool_bwd": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp],
        "nidt_bn_relu_pool_bwd_phases": [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, vp, vp],
        "nidt_bn_bwd_finalize": [vp, i32, i32, f64, vp, vp, vp, vp],
        "nidt_head_fwd_bwd": [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, u64, i32, i32, vp],
        "nidt_conv3d_fprop_tc": [vp, vp, vp, vp, vp] + [i32] * 17 + [vp],
        "nidt_conv3d_m_tiles": [i32] * 8,
        "nidt_conv3d_wgrad_tc": [vp, vp, vp, vp] + [i32] * 17 + [vp],
        "nidt_pack_conv_weights": [vp, i32, i32, i32, vp, vp, vp],
        "nidt_conv_in_grid": [],
        "nidt_conv_in_fprop": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp],
        "nidt_conv_in_wgrad": [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    lib.nidt_conv_in_s2d_supported.argtypes = [i32] * 6
    lib.nidt_conv_in_s2d_supported.restype = i32
    lib.nidt_conv_in_s2d_wgrad_scratch_floats.argtypes = []
    lib.nidt_conv_in_s2d_wgrad_scratch_floats.restype = i64
    lib.nidt_conv_in_s2d_fprop.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    lib.nidt_conv_in_s2d_fprop.restype = i32
    lib.nidt_conv_in_s2d_wgrad.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    lib.nidt_conv_in_s2d_wgrad.restype = i32
    lib.nidt_conv3d_wgrad_scratch_floats.argtypes = [i32, i32, i32, i32]
    lib.nidt_conv3d_wgrad_scratch_floats.restype = i64
    lib.nidt_pack_conv_weights_multi.argtypes = [i32, vp, vp, vp, vp, vp, vp, vp]
    lib.nidt_pack_conv_weights_multi.restype = i32
    lib.nidt_plan_pack_weights.argtypes = [vp, vp, vp]
    lib.nidt_plan_pack_weights.restype = i32
    lib.nidt_plan_train_step.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, u64, i32, vp, vp]
    lib.nidt_plan_train_step.restype = i32
    lib.nidt_plan_eval.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp]
    lib.nidt_plan_eval.restype = i32
    lib.nidt_plan_train_step_graph.argtypes = [vp, vp, vp, i32, vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, u64, i32, vp, vp, vp, vp]
    lib.nidt_plan_train_step_graph.restype = i32
    lib.nidt_plan_eval_graph.argtypes = [vp, vp, vp, i32, vp, vp, i32, vp, vp, vp]
    lib.nidt_plan_eval_graph.restype = i32
    lib.nidt_bn_relu_pool3_fwd_amax.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, vp]
    lib.nidt_bn_relu_pool3_fwd_amax.restype = i32
    lib.nidt_bn_relu_pool3_bwd_stats.argtypes = [vp, vp, i32, i32, i32, i32, i32, vp, vp, vp, vp, vp, vp, vp, vp, i32, vp]
    lib.nidt_bn_relu_pool3_bwd_stats.restype = i32
    lib.nidt_conv_in_s2d_wgrad_fused.argtypes = [vp] * 11 + [i32, i32, i32, i32, f32, vp, vp, f32, vp]
    lib.nidt_conv_in_s2d_wgrad_fused.restype = i32
    lib.nidt_plan_saliency_step.argtypes = [vp, vp, vp, i32, vp, vp, vp, f32, i64, u64, vp, vp]
    lib.nidt_plan_saliency_step.restype = i32
    lib.nidt_plan_graph_clear.argtypes = [vp]
    lib.nidt_plan_graph_clear.restype = i32
    lib.nidt_plan_graph_count.argtypes = []
    lib.nidt_plan_graph_count.restype = i32
    for name in ("nidt_conv_in_gram_floats", "nidt_conv_in_gram_max_samples", "nidt_conv_in_alg_grid"):
        getattr(lib, name).argtypes = []
        getattr(lib, name).restype = i32
    lib.nidt_conv_in_gram_scratch_floats.argtypes = [i32, i32]
    lib.nidt_conv_in_gram_scratch_floats.restype = i64
    lib.nidt_conv_in_alg_scratch_floats.argtypes = []
    lib.nidt_conv_in_alg_scratch_floats.restype = i64
    lib.nidt_conv_in_gram_samples.argtypes = [vp, vp, vp, i32, vp, vp, i32, i32, i32, f32, vp]
    lib.nidt_conv_in_gram_samples.restype = i32
    lib.nidt_conv_in_gram_sum.argtypes = [vp, vp, i32, vp, vp]
    lib.nidt_conv_in_gram_sum.restype = i32
    lib.nidt_conv_in_alg_wgrad.argtypes = [vp] * 13 + [i32, i32, i32, i32, f32, vp, vp, f32, vp]
    lib.nidt_conv_in_alg_wgrad.restype = i32
    lib.nidt_conv_in_s2d_fprop_poolw.argtypes = [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, vp]
    lib.nidt_conv_in_s2d_fprop_poolw.restype = i32
    lib.nidt_bn_relu_pool3w_fwd.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, i32, vp]
    lib.nidt_bn_relu_pool3w_fwd.restype = i32
    for name, argtypes in {
        "nidt_ew_bf16": [vp, vp, vp, i64, i32, vp],
        "nidt_maxpool3d_fwd": [vp, vp, vp] + [i32] * 8 + [vp],
        "nidt_maxpool3d_bwd": [vp, vp, vp] + [i32] * 8 + [vp],
        "nidt_upsample_zero": [vp, vp] + [i32] * 9 + [vp],
        "nidt_avgpool3d_fwd": [vp, vp] + [i32] * 6 + [vp],
        "nidt_avgpool3d_bwd": [vp, vp] + [i32] * 6 + [vp],
        "nidt_linear_head_fwd_bwd": [vp] * 16 + [i32] * 5 + [vp],
        "nidt_conv3d_halo_stats_rows": [],
        "nidt_conv3d_halo_supported": [i32] * 8,
        "nidt_conv3d_halo_fprop": [vp, vp, vp, vp, vp] + [i32] * 12 + [vp],
    }.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    _declared = True


import functools


class _BlockC(C.Structure):
    """Mirror of ``nidt_block_t`` (include/nidt_b200.h)."""
    _fields_ = ([(n, C.c_int) for n in ("cin", "cout", "k", "stride", "pad", "pool_k", "pool_s", "relu", "has_bias",
                                        "in_d", "in_h", "in_w", "c_d", "c_h", "c_w")]
                + [(n, C.c_int64) for n in ("off_w", "off_b", "off_gamma", "off_beta", "off_rm", "off_rv", "off_nbt")]
                + [(n, C.c_void_p) for n in ("y", "dy", "a", "da", "ymax", "scale", "shift", "mean", "invstd", "c12", "bwd_partials", "stats",
                                             "wf", "wd", "wg_scratch")]
                + [(n, C.c_int) for n in ("stats_tiles", "s2d", "block_n", "dgrad_block_n", "splits")]
                + [("box", C.c_int * 4), ("dbox", C.c_int * 4), ("amax", C.c_void_p), ("gram", C.c_void_p), ("alg_scratch", C.c_void_p),
                   ("yw", C.c_void_p), ("aw", C.c_void_p), ("bwd0_mode", C.c_int), ("poolw_train", C.c_int)])


class _PlanC(C.Structure):
    """Mirror of ``nidt_plan_t``."""
    _fields_ = [("n_blocks", C.c_int), ("max_batch", C.c_int), ("blocks", _BlockC * 8),
                ("off_fc1_w", C.c_int64), ("off_fc1_b", C.c_int64), ("off_fc2_w", C.c_int64), ("off_fc2_b", C.c_int64),
                ("head_s", C.c_int), ("head_c", C.c_int), ("head_hidden", C.c_int), ("p_drop", C.c_float), ("input_scale", C.c_float),
                ("loss_acc", C.c_void_p), ("eval_logits", C.c_void_p), ("norm_scratch", C.c_void_p)]


@functools.lru_cache(maxsize=None)
def choose_box(W: int, H: int, D: int, N: int, max_rows: int = 128) -> Tuple[int, int, int, int]:
    """Box of output positions (bw,bh,bd,bn), bw*bh*bd*bn <= 128, maximising tensor-core row use:
    valid positions / (tiles * 128).  Ties prefer wider (more contiguous) boxes."""
    best, best_key = (1, 1, 1, 1), (-1.0, 0)
    total = W * H * D * N
    for bw in range(1, min(W, max_rows) + 1):
        tw = -(-W // bw)
        for bh in range(1, min(H, max_rows // bw) + 1):
            th = -(-H // bh)
            for bd in range(1, min(D, max_rows // (bw * bh)) + 1):
                td = -(-D // bd)
                bn = min(N, max_rows // (bw * bh * bd))
                if bn < 1:
                    continue
                tn = -(-N // bn)
                util = total / (tw * th * td * tn * 128.0)
                key = (round(util, 4), bw * 1000 + bh)
                if key > best_key:
                    best, best_key = (bw, bh, bd, bn), key
    return best


class GramCache:
    """Per-sample Gram matrices of the first layer's input patches (``csrc/conv_in_alg.cu``), on the device, keyed by the sample's
    id in the volume store.  They depend on the volume only, so a sample pays for its matrix (0.33 ms) the first time it is trained
    on and never again; 64 KiB per sample.  ``NIDT_GRAM_CACHE_MB`` bounds the buffer (default 8192; when full it starts over)."""

    def __init__(self, lib, device, in_shape, scale: float, max_batch: int):
        self.lib, self.device, self.in_shape, self.scale = lib, device, tuple(in_shape), float(scale)
        self.floats = lib.nidt_conv_in_gram_floats()
        self.max_list = lib.nidt_conv_in_gram_max_samples()
        self.limit = max(64, (int(os.environ.get("NIDT_GRAM_CACHE_MB", "8192")) << 20) // (4 * self.floats))
        self.slots: dict = {}
        self.cap = 0
        self.buf: Optional[torch.Tensor] = None
        self.scratch: Optional[torch.Tensor] = None
        self.scratch_n = 0
        self.built = 0                                     # matrices computed so far (cold misses)
        self.key = None                                    # which volume store the ids refer to (a new store empties the cache)
        self.batch = torch.zeros(self.floats, dtype=torch.float64, device=device)       # the current batch's sum: a stable address

    def _grow(self, need: int) -> None:
        cap = min(self.limit, max(256, 2 * self.cap, need))
        if need > cap:                                     # over the limit: start over (the matrices are recomputed when next seen)
            self.slots.clear()
            cap = min(self.limit, max(256, need))
            self.buf, self.cap = None, 0
        new = torch.empty((cap, self.floats), dtype=torch.float32, device=self.device)
        if self.buf is not None and self.slots:
            new[:self.cap] = self.buf
        self.buf, self.cap = new, cap

    def use_store(self, key) -> None:
        """The ids that follow index the store identified by ``key``; matrices cached for another store are dropped."""
        if key != self.key:
            self.slots.clear()
            self.key = key

    def sum_for(self, ids, x: torch.Tensor, stream: int) -> bool:
        """``self.batch`` <- sum of the Gram matrices of the samples ``ids`` (row i of the u8 batch ``x`` is sample ids[i]); computes
        the ones not seen before.  False when the batch cannot be served (too many samples for one launch)."""
        n = len(ids)
        if n == 0 or n > self.max_list or n != x.shape[0]:
            return False
        ids = [int(i) for i in ids]
        miss = [k for k, i in enumerate(ids) if i not in self.slots]
        if miss:
            if len(self.slots) + len(miss) > self.cap:
                self._grow(len(self.slots) + len(miss))
                miss = [k for k, i in enumerate(ids) if i not in self.slots]
            first = {}
            for k in miss:                                 # a sample may occur twice in a batch
                first.setdefault(ids[k], k)
            rows = list(first.values())
            base = len(self.slots)
            for j, k in enumerate(rows):
                self.slots[ids[k]] = base + j
            if self.scratch_n < len(rows):
                self.scratch_n = max(len(rows), min(self.max_list, x.shape[0]))
                self.scratch = torch.empty(self.lib.nidt_conv_in_gram_scratch_floats(self.scratch_n, self.in_shape[0]), dtype=torch.float32,
                                           device=self.device)
            r = (C.c_int * len(rows))(*rows)
            s = (C.c_int * len(rows))(*[self.slots[ids[k]] for k in rows])
            native.check(self.lib.nidt_conv_in_gram_samples(x.data_ptr(), r, s, len(rows), self.buf.data_ptr(), self.scratch.data_ptr(),
                                                            *self.in_shape, self.scale, stream), "nidt_conv_in_gram_samples")
            native.count_launches(2)
            self.built += len(rows)
        s = (C.c_int * n)(*[self.slots[i] for i in ids])
        native.check(self.lib.nidt_conv_in_gram_sum(self.buf.data_ptr(), s, n, self.batch.data_ptr(), stream), "nidt_conv_in_gram_sum")
        native.count_launches(1)
        return True


class _Block:
    pass


class B200Engine:
    native = True

    def __init__(self, model: nn.Module, device, max_batch: int = 16, input_scale: float = 1.0 / 255.0, seed: int = 1234,
                 wgrad_ctas: int = 296, use_s2d: bool = True):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("B200Engine needs a CUDA device")
        if not hasattr(model, "plan") or not hasattr(model, "head_spec"):
            raise ImportError("model has no native plan")
        self.lib = native.load()
        _declare(self.lib)
        self.model = model
        self.plan: List[ConvBlockSpec] = list(model.plan)
        self.head: HeadSpec = model.head_spec
        self.in_shape = tuple(model.in_shape)
        self.layout = SlabLayout.from_module(model)
        self.input_scale = float(input_scale)
        self.max_batch = int(max_batch)
        self.seed = seed
        self.step = 0
        p0 = self.plan[0]
        if not (p0.cin == 1 and p0.cout == 64 and p0.k in (3, 5)):
            raise ImportError("first block must be a 1->64 channel conv with k in (3,5)")
        for b in self.plan[1:]:
            if b.cin % 64 or b.stride != 1 or b.cout % 8:
                raise ImportError(f"block {b.conv}: native path needs Cin%64==0, stride 1")
        if self.head.out_features != 1:
            raise ImportError("native head supports one logit")
        self.grad = ParamSlab(self.layout, device=self.device, params_only=True)
        self.slab: Optional[ParamSlab] = None
        # one slot per (client, epoch) mean loss: popping a loss moves to the next zeroed slot instead of launching a divide + a fill
        self._loss_ring = torch.zeros(1024, dtype=torch.float32, device=self.device)
        self._loss_pos = 0
        self.loss_acc = self._loss_ring[0]
        self.loss_steps = 0
        self._packed_dirty = True
        self.profile = None
        self._tag = ""
        self.use_s2d = use_s2d
        self.s2d_wgrad = True
        self._seed_extra = 0
        self._mode = 0                    # block-0 backward variant of the step being issued by the Python launch loop (see _bwd0_mode)
        self.batch_ids = None             # store ids of the samples of the NEXT training batch (set by the trainer, consumed by the step)
        self.batch_ids_key = None         # identity of the store those ids index (per-sample caches are per store)
        self._alloc(wgrad_ctas)
        # native step executor (csrc/executor.cu): one C call per step instead of one per kernel; NIDT_PY_EXEC=1 keeps the
        # Python launch loop (it is also what start_profile() switches to, for per-kernel timings)
        self.native_exec = os.environ.get("NIDT_PY_EXEC", "0") != "1"
        self._plan = self._build_plan()
        self._launches = C.c_int(0)
        # CUDA-graph replays of the native step (csrc/executor.cu): one graph per (slab, batch buffers); NIDT_GRAPHS=0 turns it off.
        # Needs stable batch buffers, so trainers make their stagers gather every batch into one buffer (``use_graphs``).
        self.use_graphs = self.native_exec and os.environ.get("NIDT_GRAPHS", "1") != "0"
        self._dyn = torch.zeros(64, dtype=torch.uint8, device=self.device)      # nidt_dyn_t (48 bytes)

    def __del__(self):
        try:
            self.lib.nidt_plan_graph_clear(C.byref(self._plan))
        except Exception:
            pass

    # ----------------------------------------------------------------------------- buffers
    def _alloc(self, wgrad_ctas: int) -> None:
        dev, B = self.device, self.max_batch
        bf = torch.bfloat16
        shapes = trace_shapes(self.plan, self.in_shape)
        self.blocks: List[_Block] = []
        lib = self.lib
        in_dhw = self.in_shape
        bn_grid = lib.nidt_bn_bwd_grid()
        for i, (spec, (cdhw, adhw)) in enumerate(zip(self.plan, shapes)):
            b = _Block()
            b.spec, b.in_dhw, b.c_dhw, b.a_dhw = spec, in_dhw, cdhw, adhw
            D, H, W = cdhw
            b.y = torch.empty((B, D, H, W, spec.cout), dtype=bf, device=dev)
            b.dy = torch.empty_like(b.y)
            b.a = torch.empty((B,) + adhw + (spec.cout,), dtype=bf, device=dev) if spec.pool_k else torch.empty_like(b.y)
            b.da = torch.empty_like(b.a)
            b.ymax = torch.empty_like(b.a) if (spec.pool_k == 3 and spec.pool_s == 3) else None
            C_ = spec.cout
            b.scale, b.shift, b.mean, b.invstd = (torch.zeros(C_, dtype=torch.float32, device=dev) for _ in range(4))
            b.c12 = torch.zeros(2 * C_, dtype=torch.float32, device=dev)
            b.bwd_partials = torch.zeros(bn_grid * 2 * C_, dtype=torch.float32, device=dev)
            taps = spec.k ** 3
            if i == 0:
                b.stats = torch.zeros(lib.nidt_conv_in_grid() * 2 * C_, dtype=torch.float32, device=dev)
                b.stats_tiles = lib.nidt_conv_in_grid()
                iD, iH, iW = in_dhw
                b.s2d = bool(lib.nidt_conv_in_s2d_supported(iD, iH, iW, spec.k, spec.stride, spec.pad)) and self.use_s2d
                # NIDT_FUSED_BWD0=1: winning pool taps for the fused BatchNorm-backward + weight-gradient kernel of this block
                # (nidt_conv_in_s2d_wgrad_fused: dY never written to HBM).  Bit-identical to the separate kernels, but MEASURED SLOWER
                # at batch 16 (0.62 + 0.03 ms against 0.247 + 0.235 ms, profiles/profile_step_fused_bwd0.txt): the in-place y -> dY
                # transform competes with the 8 staging warps for the same issue slots, which this kernel is already bound by.  Off.
                # NIDT_ALG_BWD0 (default on): the backward of this block without any pass over y / dY -- weight gradient from one input
                # patch per pooling window + the batch's cached patch Gram matrix (csrc/conv_in_alg.cu); used for the steps whose
                # sample ids the caller supplies (``batch_ids``), the dY-based kernels otherwise.
                self.fused_bwd0 = os.environ.get("NIDT_FUSED_BWD0", "0") == "1"
                # (pooled width <= 19: the sparse pass walks one strip of up to 19 windows per unit; wider rows are split into two strips,
                # a path no measured configuration has exercised, so such shapes keep the dY-based kernels)
                self.alg_bwd0 = (os.environ.get("NIDT_ALG_BWD0", "1") != "0" and b.s2d and b.ymax is not None and spec.cout == 64 and spec.k == 5
                                 and W >= 3 and (W - 3) // 3 + 1 <= 19)
                want_amax = b.s2d and b.ymax is not None and spec.cout == 64 and (self.fused_bwd0 or self.alg_bwd0)
                b.amax = torch.empty(b.a.shape, dtype=torch.uint8, device=dev) if want_amax else None
                self.gram = GramCache(lib, dev, in_dhw, self.input_scale, B) if self.alg_bwd0 else None
                b.alg_scratch = (torch.empty(lib.nidt_conv_in_alg_scratch_floats(), dtype=torch.float32, device=dev) if self.alg_bwd0 else None)
                # NIDT_POOLW: when nobody reads the full-resolution conv output afterwards, the conv epilogue reduces every 3 output
                # columns to the pooling winner and the pooling kernel finishes from a third of the data (nidt_conv_in_s2d_fprop_poolw +
                # nidt_bn_relu_pool3w_fwd); bit-identical results.  "eval" (default): evaluation batches only -- MEASURED at batch 16:
                # evaluation 0.501 -> 0.456 ms, but a training step 1.10 -> 1.14 ms (with the BatchNorm statistics loop in the same
                # epilogue warps the extra work costs the conv more than the smaller pooling pass saves); "1": training steps with the
                # backward from pooled tensors too; "0": off.
                pw_env = os.environ.get("NIDT_POOLW", "eval")
                self.poolw = pw_env in ("1", "eval") and b.s2d and b.ymax is not None and spec.cout == 64 and spec.k == 5 and W >= 3
                self.poolw_train = self.poolw and pw_env == "1"
                if self.poolw:
                    Wp = (W - 3) // 3 + 1
                    b.yw = torch.empty((B, D, H, Wp, 64), dtype=bf, device=dev)
                    b.aw = torch.empty((B, D, H, Wp, 8), dtype=torch.int16, device=dev)
                else:
                    b.yw = b.aw = None
                n_scr = max(lib.nidt_conv_in_grid() * 128 * 64, lib.nidt_conv_in_s2d_wgrad_scratch_floats())
                b.wg_scratch = torch.zeros(n_scr, dtype=torch.float32, device=dev)
            else:
                b.box = choose_box(W, H, D, B)
                b.m_tiles = lib.nidt_conv3d_m_tiles(B, D, H, W, *b.box)
                b.block_n = 128 if (spec.cout % 128 == 0 and b.m_tiles * (spec.cout // 128) >= 148) else 64
                b.stats = torch.zeros(b.m_tiles * 2 * C_, dtype=torch.float32, device=dev)
                b.stats_tiles = b.m_tiles
                b.wf = torch.empty((spec.cout, taps, spec.cin), dtype=bf, device=dev)
                b.wd = torch.empty((spec.cin, taps, spec.cout), dtype=bf, device=dev)
                iD, iH, iW = in_dhw
                b.dbox = choose_box(iW, iH, iD, B)
                b.dgrad_block_n = 128 if (spec.cin % 128 == 0 and lib.nidt_conv3d_m_tiles(B, iD, iH, iW, *b.dbox) * (spec.cin // 128) >= 148) else 64
                work = taps * (-(-spec.cout // 128)) * (spec.cin // 64)
                b.splits = max(1, min(b.m_tiles, wgrad_ctas // work))
                b.wg_scratch = torch.zeros(lib.nidt_conv3d_wgrad_scratch_floats(spec.cout, spec.cin, spec.k, b.splits),
                                           dtype=torch.float32, device=dev)
            self.blocks.append(b)
            in_dhw = adhw
        d, h, w = shapes[-1][1]
        self.head_S = d * h * w
        self.head_C = self.plan[-1].cout
        assert self.head_S * self.head_C == self.head.in_features
        self.eval_logits = torch.zeros(B, dtype=torch.float32, device=dev)

    def _build_plan(self) -> _PlanC:
        lay = self.layout
        off = lambda name: int(lay[name].offset) if name in lay else -1
        ptr = lambda t: (t.data_ptr() if t is not None else None)
        p = _PlanC()
        p.n_blocks, p.max_batch = len(self.blocks), self.max_batch
        assert len(self.blocks) <= 8
        for i, b in enumerate(self.blocks):
            s, c = b.spec, p.blocks[i]
            c.cin, c.cout, c.k, c.stride, c.pad = s.cin, s.cout, s.k, s.stride, s.pad
            c.pool_k, c.pool_s, c.relu, c.has_bias = int(s.pool_k or 0), int(s.pool_s or 0), int(s.relu), int(bool(s.bias))
            (c.in_d, c.in_h, c.in_w), (c.c_d, c.c_h, c.c_w) = b.in_dhw, b.c_dhw
            c.off_w, c.off_b = off(s.conv + ".weight"), (off(s.conv + ".bias") if s.bias else -1)
            c.off_gamma, c.off_beta = off(s.bn + ".weight"), off(s.bn + ".bias")
            c.off_rm, c.off_rv, c.off_nbt = off(s.bn + ".running_mean"), off(s.bn + ".running_var"), off(s.bn + ".num_batches_tracked")
            c.y, c.dy, c.a, c.da, c.ymax = ptr(b.y), ptr(b.dy), ptr(b.a), ptr(b.da), ptr(b.ymax)
            c.scale, c.shift, c.mean, c.invstd = ptr(b.scale), ptr(b.shift), ptr(b.mean), ptr(b.invstd)
            c.c12, c.bwd_partials, c.stats, c.wg_scratch = ptr(b.c12), ptr(b.bwd_partials), ptr(b.stats), ptr(b.wg_scratch)
            c.stats_tiles = int(b.stats_tiles)
            if i == 0:
                c.s2d = int(b.s2d and self.s2d_wgrad)
                c.amax = ptr(getattr(b, "amax", None)) if (b.s2d and self.s2d_wgrad) else None
                if self.alg_bwd0 and c.amax:
                    c.gram, c.alg_scratch = self.gram.batch.data_ptr(), b.alg_scratch.data_ptr()
                if self.poolw and c.s2d:
                    c.yw, c.aw, c.poolw_train = b.yw.data_ptr(), b.aw.data_ptr(), int(self.poolw_train)
                c.bwd0_mode = 1 if (self.fused_bwd0 and c.amax) else 0
            else:
                c.wf, c.wd = ptr(b.wf), ptr(b.wd)
                c.block_n, c.dgrad_block_n, c.splits = b.block_n, b.dgrad_block_n, b.splits
                for j in range(4):
                    c.box[j], c.dbox[j] = b.box[j], b.dbox[j]
        h = self.head
        p.off_fc1_w, p.off_fc1_b, p.off_fc2_w, p.off_fc2_b = off(h.fc1 + ".weight"), off(h.fc1 + ".bias"), off(h.fc2 + ".weight"), off(h.fc2 + ".bias")
        p.head_s, p.head_c, p.head_hidden, p.p_drop, p.input_scale = self.head_S, self.head_C, h.hidden, float(h.p_drop), self.input_scale
        p.eval_logits = self.eval_logits.data_ptr()
        # own grad-norm partials: several engines (client lanes) may run on different streams at once
        self._norm_scratch_t = torch.zeros(self.lib.nidt_norm_scratch_floats(), dtype=torch.float32, device=self.device)
        p.norm_scratch = self._norm_scratch_t.data_ptr()
        return p

    def _native_ready(self) -> bool:
        """Native executor usable for this call; packs the weights of a newly attached slab."""
        if not self.native_exec or self.profile is not None:
            return False
        self._plan.loss_acc = self.loss_acc.data_ptr()      # pop_mean_loss() replaces the accumulator tensor
        if self._packed_dirty:
            native.check(self.lib.nidt_plan_pack_weights(C.byref(self._plan), self.slab.flat.data_ptr(), self._stream()), "nidt_plan_pack_weights")
            native.count_launches(1)
            self._packed_dirty = False
        return True

    # ------------------------------------------------------------------------- slab plumbing
    def new_slab(self) -> ParamSlab:
        return ParamSlab.from_module(self.model, device=self.device, bind=False)

    def attach(self, slab: ParamSlab) -> None:
        if slab is not self.slab:
            self._packed_dirty = True
        self.slab = slab

    def detach_sync(self) -> None:
        return None

    def _p(self, slab: ParamSlab, name: str) -> int:
        return slab.flat.data_ptr() + 4 * self.layout[name].offset

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _chk(self, rc: int, what: str) -> None:
        native.check(rc, what)
        if self.profile is not None:      # per-launch CUDA-event timing (scripts/profile_step.py)
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.profile.append((what + self._tag, ev))

    def start_profile(self) -> None:
        self.profile = []
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.profile.append(("<start>", ev))

    def stop_profile(self):
        """-> list of (kernel tag, ms) in launch order (time between consecutive completion events)."""
        torch.cuda.synchronize(self.device)
        out = [(n, self.profile[i - 1][1].elapsed_time(e)) for i, (n, e) in enumerate(self.profile) if i > 0]
        self.profile = None
        return out

    # ------------------------------------------------------------------------------ kernels
    def _pack_weights(self) -> None:
        lib, st = self.lib, self._stream()
        for b in self.blocks[1:]:
            s = b.spec
            self._chk(lib.nidt_pack_conv_weights(self._p(self.slab, s.conv + ".weight"), s.cout, s.cin, s.k ** 3,
                                                 b.wf.data_ptr(), b.wd.data_ptr(), st), "pack_conv_weights")
            native.count_launches(1)
        self._packed_dirty = False

    def _forward(self, x: torch.Tensor, train: bool) -> int:
        """Runs the conv stack; returns the batch size.  ``x``: uint8 [B, D, H, W] on the device."""
        assert x.dtype == torch.uint8 and x.is_cuda and x.dim() == 4, "B200Engine consumes 8-bit volumes [B,D,H,W]"
        x = x.contiguous()
        B = x.shape[0]
        assert B <= self.max_batch and tuple(x.shape[1:]) == self.in_shape
        lib, st, slab = self.lib, self._stream(), self.slab
        if self._packed_dirty:
            self._pack_weights()
        self._x = x
        prev = None
        for i, b in enumerate(self.blocks):
            s = b.spec
            self._tag = f"[{i}]"
            D, H, W = b.c_dhw
            stats = b.stats.data_ptr() if train else None
            bias = self._p(slab, s.conv + ".bias") if s.bias else None
            poolw = (i == 0 and getattr(self, "poolw", False) and b.s2d and self.s2d_wgrad
                     and (not train or (self.poolw_train and self._mode == 2 and getattr(b, "amax", None) is not None)))
            if i == 0:
                iD, iH, iW = b.in_dhw
                if poolw:     # nobody reads the full-resolution y afterwards: reduced along W in the conv epilogue
                    self._chk(lib.nidt_conv_in_s2d_fprop_poolw(x.data_ptr(), self._p(slab, s.conv + ".weight"), bias, self._p(slab, s.bn + ".weight"),
                                                               b.yw.data_ptr(), b.aw.data_ptr(), stats, B, iD, iH, iW, self.input_scale, st),
                              "conv_in_s2d_fprop_poolw")
                elif b.s2d:   # 5^3 / stride 2: space-to-depth lines feed tcgen05 directly, no im2col
                    self._chk(lib.nidt_conv_in_s2d_fprop(x.data_ptr(), self._p(slab, s.conv + ".weight"), bias, b.y.data_ptr(), stats,
                                                         B, iD, iH, iW, self.input_scale, st), "conv_in_s2d_fprop")
                else:
                    self._chk(lib.nidt_conv_in_fprop(x.data_ptr(), self._p(slab, s.conv + ".weight"), bias, b.y.data_ptr(), stats,
                                                     B, iD, iH, iW, s.k, s.stride, s.pad, self.input_scale, st), "conv_in_fprop")
            else:
                iD, iH, iW = b.in_dhw
                bw, bh, bd, bn = b.box if B == self.max_batch else choose_box(W, H, D, B)
                b.cur_box = (bw, bh, bd, bn)
                self._chk(lib.nidt_conv3d_fprop_tc(prev.data_ptr(), b.wf.data_ptr(), bias, b.y.data_ptr(), stats, B, iD, iH, iW, s.cin,
                                                   D, H, W, s.cout, s.k, s.stride, s.pad, bw, bh, bd, bn, b.block_n, st), "conv3d_fprop_tc")
            native.count_launches(1)
            g, be = self._p(slab, s.bn + ".weight"), self._p(slab, s.bn + ".bias")
            rm, rv = self._p(slab, s.bn + ".running_mean"), self._p(slab, s.bn + ".running_var")
            if train:
                tiles = b.stats_tiles if (i == 0 or B == self.max_batch) else lib.nidt_conv3d_m_tiles(B, D, H, W, *b.cur_box)
                stats_ptr, tiles, count = self._bn_batch_sums(b, tiles, float(B * D * H * W))
                self._chk(lib.nidt_bn_finalize(stats_ptr, tiles, s.cout, count, g, be, rm, rv,
                                               self._p(slab, s.bn + ".num_batches_tracked"), 0.1, 1e-5, b.scale.data_ptr(), b.shift.data_ptr(),
                                               b.mean.data_ptr(), b.invstd.data_ptr(), st), "bn_finalize")
            else:
                self._chk(lib.nidt_bn_eval_prepare(s.cout, g, be, rm, rv, 1e-5, b.scale.data_ptr(), b.shift.data_ptr(), st), "bn_eval_prepare")
            if poolw:
                self._chk(lib.nidt_bn_relu_pool3w_fwd(b.yw.data_ptr(), b.aw.data_ptr(), b.a.data_ptr(), b.ymax.data_ptr() if train else None,
                                                      b.amax.data_ptr() if train else None, B, D, H, W, s.cout, b.scale.data_ptr(), b.shift.data_ptr(),
                                                      int(s.relu), st), "bn_relu_pool3w_fwd")
            elif train and i == 0 and self._mode != 0 and getattr(b, "amax", None) is not None and self.s2d_wgrad:
                self._chk(lib.nidt_bn_relu_pool3_fwd_amax(b.y.data_ptr(), b.a.data_ptr(), b.ymax.data_ptr(), b.amax.data_ptr(), B, D, H, W, s.cout,
                                                          b.scale.data_ptr(), b.shift.data_ptr(), int(s.relu), st), "bn_relu_pool_fwd")
            else:
                self._chk(lib.nidt_bn_relu_pool_fwd(b.y.data_ptr(), b.a.data_ptr(), b.ymax.data_ptr() if (train and b.ymax is not None) else None, B, D, H, W, s.cout, s.pool_k, s.pool_s,
                                                    b.scale.data_ptr(), b.shift.data_ptr(), int(s.relu), st), "bn_relu_pool_fwd")
            native.count_launches(2)
            prev = b.a
        return B

    def _head(self, B: int, y: torch.Tensor, train: bool, counters=None, double_sigmoid=True, want_grads=True) -> None:
        lib, st, slab, h = self.lib, self._stream(), self.slab, self.head
        self._tag = ""
        last = self.blocks[-1]
        gp = (lambda n: self._p(self.grad, n)) if (train and want_grads) else (lambda n: None)
        self._chk(lib.nidt_head_fwd_bwd(
            last.a.data_ptr(), y.data_ptr(), self._p(slab, h.fc1 + ".weight"), self._p(slab, h.fc1 + ".bias"),
            self._p(slab, h.fc2 + ".weight"), self._p(slab, h.fc2 + ".bias"),
            gp(h.fc1 + ".weight"), gp(h.fc1 + ".bias"), gp(h.fc2 + ".weight"), gp(h.fc2 + ".bias"),
            last.da.data_ptr() if train else None, self.loss_acc.data_ptr() if train else None,
            counters.buf.data_ptr() if counters is not None else None, self.eval_logits.data_ptr(),
            B, self.head_S, self.head_C, h.hidden, float(h.p_drop), self.seed + 7919 * self.step + self._seed_extra, int(train),
            int(double_sigmoid), st),
            "head_fwd_bwd")
        native.count_launches(1)

    def _backward(self, B: int) -> None:
        lib, st, slab = self.lib, self._stream(), self.slab
        for i in range(len(self.blocks) - 1, -1, -1):
            b = self.blocks[i]
            s = b.spec
            self._tag = f"[{i}]"
            D, H, W = b.c_dhw
            if i == 0 and self._mode == 2:
                # BatchNorm-backward reductions over the pooled tensors, then the weight gradient from one input patch per pooling
                # window and the batch's cached patch Gram matrix: no pass over y or dY (csrc/conv_in_alg.cu)
                iD, iH, iW = b.in_dhw
                self._chk(lib.nidt_bn_relu_pool3_bwd_stats(b.ymax.data_ptr(), b.da.data_ptr(), B, D, H, W, s.cout, b.scale.data_ptr(), b.shift.data_ptr(),
                                                           b.mean.data_ptr(), b.invstd.data_ptr(), b.bwd_partials.data_ptr(), b.c12.data_ptr(),
                                                           self._p(self.grad, s.bn + ".weight"), self._p(self.grad, s.bn + ".bias"), int(s.relu), st),
                          "bn_pool3_bwd_stats")
                self._chk(lib.nidt_conv_in_alg_wgrad(self._x.data_ptr(), b.amax.data_ptr(), b.da.data_ptr(), self._p(slab, s.conv + ".weight"),
                                                     self._p(slab, s.conv + ".bias") if s.bias else None, b.scale.data_ptr(), b.mean.data_ptr(),
                                                     b.invstd.data_ptr(), b.c12.data_ptr(), self.gram.batch.data_ptr(),
                                                     self._p(self.grad, s.conv + ".weight"), self._p(self.grad, s.conv + ".bias") if s.bias else None,
                                                     b.alg_scratch.data_ptr(), B, iD, iH, iW, self.input_scale, None, None, 0.0, st),
                          "conv_in_alg_wgrad")
                native.count_launches(4)
                continue
            if i == 0 and self._mode == 1 and getattr(b, "amax", None) is not None and b.s2d and self.s2d_wgrad:
                # BatchNorm-backward reductions, then dY formed inside the weight-gradient kernel (never written to HBM)
                iD, iH, iW = b.in_dhw
                self._chk(lib.nidt_bn_relu_pool3_bwd_stats(b.ymax.data_ptr(), b.da.data_ptr(), B, D, H, W, s.cout, b.scale.data_ptr(), b.shift.data_ptr(),
                                                           b.mean.data_ptr(), b.invstd.data_ptr(), b.bwd_partials.data_ptr(), b.c12.data_ptr(),
                                                           self._p(self.grad, s.bn + ".weight"), self._p(self.grad, s.bn + ".bias"), int(s.relu), st),
                          "bn_pool3_bwd_stats")
                gb = self._p(self.grad, s.conv + ".bias") if s.bias else None
                self._chk(lib.nidt_conv_in_s2d_wgrad_fused(self._x.data_ptr(), b.y.data_ptr(), b.amax.data_ptr(), b.da.data_ptr(), b.scale.data_ptr(),
                                                           b.mean.data_ptr(), b.invstd.data_ptr(), b.c12.data_ptr(), self._p(self.grad, s.conv + ".weight"), gb,
                                                           b.wg_scratch.data_ptr(), B, iD, iH, iW, self.input_scale, None, None, 0.0, st),
                          "conv_in_s2d_wgrad_fused")
                native.count_launches(4)
                continue
            self._bn_backward(b, B)
            iD, iH, iW = b.in_dhw
            if i == 0:
                gb = self._p(self.grad, s.conv + ".bias") if s.bias else None
                if b.s2d and self.s2d_wgrad:
                    self._chk(lib.nidt_conv_in_s2d_wgrad(self._x.data_ptr(), b.dy.data_ptr(), self._p(self.grad, s.conv + ".weight"), gb,
                                                         b.wg_scratch.data_ptr(), B, iD, iH, iW, self.input_scale, st), "conv_in_s2d_wgrad")
                else:
                    self._chk(lib.nidt_conv_in_wgrad(self._x.data_ptr(), b.dy.data_ptr(), self._p(self.grad, s.conv + ".weight"), gb,
                                                     b.wg_scratch.data_ptr(), B, iD, iH, iW, s.k, s.stride, s.pad, self.input_scale, st), "conv_in_wgrad")
                native.count_launches(2)
            else:
                prev = self.blocks[i - 1]
                bw, bh, bd, bn = b.cur_box
                splits = b.splits if B == self.max_batch else max(1, min(b.splits, lib.nidt_conv3d_m_tiles(B, D, H, W, bw, bh, bd, bn)))
                self._chk(lib.nidt_conv3d_wgrad_tc(b.dy.data_ptr(), prev.a.data_ptr(), self._p(self.grad, s.conv + ".weight"),
                                                   b.wg_scratch.data_ptr(), B, iD, iH, iW, s.cin, D, H, W, s.cout, s.k, s.stride, s.pad,
                                                   bw, bh, bd, bn, splits, st), "conv3d_wgrad_tc")
                # dgrad = fprop of dY with transposed, tap-flipped weights and pad' = k-1-pad
                dbox = b.dbox if B == self.max_batch else choose_box(iW, iH, iD, B)
                self._chk(lib.nidt_conv3d_fprop_tc(b.dy.data_ptr(), b.wd.data_ptr(), None, prev.da.data_ptr(), None, B, D, H, W, s.cout,
                                                   iD, iH, iW, s.cin, s.k, 1, s.k - 1 - s.pad, *dbox, b.dgrad_block_n, st), "conv3d_dgrad_tc")
                native.count_launches(3)
            # Conv bias gradients: block 0 gets sum(dY) for free from the bias column of its wgrad GEMM.  For the
            # other blocks it is identically zero in exact arithmetic (BatchNorm subtracts the batch mean right
            # after the conv): the grad slab keeps 0 there, weight decay still applies in the update.

    def _bn_batch_sums(self, b: _Block, tiles: int, count: float):
        """(pointer, rows, elements per channel) of the [rows][2][C] partial sums the BatchNorm finalisation reads: the conv epilogue's
        per-tile sums of this device's batch (the data-parallel engine replaces them with the sums over all ranks)."""
        return b.stats.data_ptr(), tiles, count

    def _bn_backward(self, b: _Block, B: int) -> None:
        """dY of a block from d(pooled activation): BatchNorm + ReLU + MaxPool backward, BN weight / bias gradients into the slab."""
        s = b.spec
        D, H, W = b.c_dhw
        self._chk(self.lib.nidt_bn_relu_pool_bwd(b.y.data_ptr(), b.ymax.data_ptr() if b.ymax is not None else None, b.da.data_ptr(), b.dy.data_ptr(),
                                                 B, D, H, W, s.cout, s.pool_k, s.pool_s, b.scale.data_ptr(), b.shift.data_ptr(), b.mean.data_ptr(),
                                                 b.invstd.data_ptr(), b.bwd_partials.data_ptr(), b.c12.data_ptr(),
                                                 self._p(self.grad, s.bn + ".weight"), self._p(self.grad, s.bn + ".bias"), int(s.relu),
                                                 self._stream()), "bn_relu_pool_bwd")
        native.count_launches(3)

    # ---------------------------------------------------------------------------- public API
    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        """Forward + backward; gradients land in ``self.grad.flat``.  ``train_mode=False`` (DisPFL
        gradient screening) uses running statistics and no dropout in the forward pass."""
        if not train_mode:
            raise NotImplementedError("eval-mode gradient screening runs on the TorchEngine")
        y = y.to(self.device, torch.float32).contiguous()
        if self._native_ready():
            self._native_step(x, y, None, None, None, False, do_update=False)
            return self.loss_acc
        self._mode = self._bwd0_mode(self._check_x(x))
        B = self._forward(x, train=True)
        self._head(B, y, train=True)
        self._backward(B)
        self.step += 1
        return self.loss_acc

    def saliency_step(self, x, y, scores: torch.Tensor, scale: float) -> bool:
        """One SNIP evaluation with the scores produced by the kernels that produce the gradients: ``scores += scale*|g*w|`` for
        every conv weight leaves the split-K reduction of its wgrad, the classifier's segment is one flat pass (``snip.py:60-71``).
        Non-prunable parameters (biases, BatchNorm) are not scored -- the mask is 1 there whatever their score.  Returns False when
        this engine configuration has no fused path (the caller then scores the gradient slab in a separate pass)."""
        b0 = self.blocks[0]
        if not (self._native_ready() and b0.s2d and self.s2d_wgrad):
            return False
        lay, h = self.layout, self.head
        o = lay[h.fc1 + ".weight"].offset
        if any(lay[s.conv + ".weight"].offset >= o for s in self.plan) or lay[h.fc2 + ".weight"].offset < o:
            return False
        x = self._check_x(x)
        self._x = x
        self._plan.blocks[0].bwd0_mode = self._bwd0_mode(x)
        y = y.to(self.device, torch.float32).contiguous()
        native.check(self.lib.nidt_plan_saliency_step(
            C.byref(self._plan), x.data_ptr(), y.data_ptr(), int(x.shape[0]), self.slab.flat.data_ptr(), self.grad.flat.data_ptr(),
            native.ptr(scores), float(scale), lay.param_numel, self.seed + 7919 * self.step, C.byref(self._launches), self._stream()),
            "nidt_plan_saliency_step")
        native.count_launches(self._launches.value)
        self.step += 1
        return True

    def _check_x(self, x: torch.Tensor) -> torch.Tensor:
        assert x.dtype == torch.uint8 and x.is_cuda and x.dim() == 4, "B200Engine consumes 8-bit volumes [B,D,H,W]"
        x = x.contiguous()
        assert x.shape[0] <= self.max_batch and tuple(x.shape[1:]) == self.in_shape
        return x

    def _bwd0_mode(self, x: torch.Tensor) -> int:
        """Backward variant of block 0 for this batch: 2 (no pass over y / dY) when the sample ids are known, so that the cached Gram
        matrices can be summed; else 1 (NIDT_FUSED_BWD0) or 0."""
        ids, self.batch_ids = self.batch_ids, None
        base = 1 if (self.fused_bwd0 and self._plan.blocks[0].amax) else 0
        if ids is None or not self.alg_bwd0 or not self.s2d_wgrad or not self._plan.blocks[0].amax:
            return base
        self.gram.use_store(self.batch_ids_key)
        return 2 if self.gram.sum_for(ids, x, self._stream()) else base

    def _native_step(self, x, y, hp, mask, mom, first_step: bool, do_update: bool) -> None:
        x = self._check_x(x)
        self._x = x
        self._plan.blocks[0].bwd0_mode = self._bwd0_mode(x)
        n = self.layout.param_numel
        if self.use_graphs:
            native.check(self.lib.nidt_plan_train_step_graph(
                C.byref(self._plan), x.data_ptr(), y.data_ptr(), int(x.shape[0]), self.slab.flat.data_ptr(), self.grad.flat.data_ptr(),
                native.ptr(mom) if (do_update and hp.momentum != 0) else None, native.ptr(mask) if do_update else None, n,
                float(hp.lr) if do_update else 0.0, float(hp.momentum) if do_update else 0.0, float(hp.wd) if do_update else 0.0,
                float(hp.clip_norm) if do_update else 0.0, int(first_step), self.seed + 7919 * self.step, int(do_update),
                self.loss_acc.data_ptr(), self._dyn.data_ptr(), C.byref(self._launches), self._stream()), "nidt_plan_train_step_graph")
            native.count_launches(self._launches.value)
            self.step += 1
            return
        native.check(self.lib.nidt_plan_train_step(
            C.byref(self._plan), x.data_ptr(), y.data_ptr(), int(x.shape[0]), self.slab.flat.data_ptr(), self.grad.flat.data_ptr(),
            native.ptr(mom) if (do_update and hp.momentum != 0) else None, native.ptr(mask) if do_update else None, n,
            float(hp.lr) if do_update else 0.0, float(hp.momentum) if do_update else 0.0, float(hp.wd) if do_update else 0.0,
            float(hp.clip_norm) if do_update else 0.0, int(first_step), self.seed + 7919 * self.step, int(do_update),
            C.byref(self._launches), self._stream()), "nidt_plan_train_step")
        native.count_launches(self._launches.value)
        self.step += 1

    def train_step(self, x, y, hp, mask: Optional[torch.Tensor], mom: Optional[torch.Tensor], first_step: bool,
                   grad_mask: Optional[torch.Tensor] = None) -> None:
        if grad_mask is None and self._native_ready():
            if mask is not None:
                assert mask.numel() >= self.layout.param_numel
            self._native_step(x, y.to(self.device, torch.float32).contiguous(), hp, mask, mom, first_step, do_update=True)
            self.loss_steps += 1
            return
        self.compute_grads(x, y, True)
        self.loss_steps += 1
        if grad_mask is not None:
            self.grad.flat.mul_(grad_mask)
        flat_ops.masked_sgd_step(self.slab.params_flat, self.grad.flat, mom, mask, hp.lr, hp.momentum, hp.wd, hp.clip_norm, first_step)
        self._packed_dirty = True

    def pop_mean_loss(self, sync: bool = True):
        """Mean loss since the last call; ``sync=False`` returns a 0-d device tensor (no host synchronisation)."""
        from ..fl.engine import LazyMean
        out = LazyMean(self.loss_acc, self.loss_steps)
        self._loss_pos += 1
        if self._loss_pos >= self._loss_ring.numel():      # handles keep the old ring alive until they are read
            self._loss_ring = torch.zeros_like(self._loss_ring)
            self._loss_pos = 0
        self.loss_acc = self._loss_ring[self._loss_pos]
        self.loss_steps = 0
        return out.item() if sync else out

    def eval_batch(self, x, y, counters, double_sigmoid: bool = True) -> None:
        y = y.to(self.device, torch.float32).contiguous()
        if self._native_ready():
            x = self._check_x(x)
            if self.use_graphs and counters is not None:
                native.check(self.lib.nidt_plan_eval_graph(C.byref(self._plan), x.data_ptr(), y.data_ptr(), int(x.shape[0]), self.slab.flat.data_ptr(),
                                                           counters.buf.data_ptr(), int(double_sigmoid), self._dyn.data_ptr(),
                                                           C.byref(self._launches), self._stream()), "nidt_plan_eval_graph")
                native.count_launches(self._launches.value)
                return
            native.check(self.lib.nidt_plan_eval(C.byref(self._plan), x.data_ptr(), y.data_ptr(), int(x.shape[0]), self.slab.flat.data_ptr(),
                                                 counters.buf.data_ptr() if counters is not None else None, int(double_sigmoid),
                                                 C.byref(self._launches), self._stream()), "nidt_plan_eval")
            native.count_launches(self._launches.value)
            return
        B = self._forward(x, train=False)
        self._head(B, y, train=False, counters=counters, double_sigmoid=double_sigmoid)

    def logits(self, x) -> torch.Tensor:
        """Eval-mode logits of one batch (testing aid)."""
        B = self._forward(x, train=False)
        yz = torch.zeros(B, dtype=torch.float32, device=self.device)
        self._head(B, yz, train=False)
        return self.eval_logits[:B].clone()
