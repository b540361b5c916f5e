"""sm_100a compute engine for the 3-D ResNets (``models.volumetric.ResNet3D_l3`` with ``BasicBlock3D`` -- BASELINE config 4 -- or
``Bottleneck3D`` blocks).

Same contract as :class:`.engine_b200.B200Engine` (no autograd, no cuDNN/cuBLAS, parameters and gradients in flat slabs,
channels-last bf16 activations), for a residual topology:

  stem    conv_in (1->64, k3 s2 p3, tcgen05, im2col in smem from the 8-bit volume) -> BN+ReLU -> MaxPool(3,2,1)
  basic   conv3x3x3 (tcgen05+TMA; stride 2 = element-strided TMA boxes) -> BN+ReLU -> conv3x3x3 -> BN
          (+ 1x1x1 stride-2 conv -> BN on the shortcut) -> add + ReLU
  bottle  conv1x1x1 -> BN+ReLU -> conv3x3x3 (stride) -> BN+ReLU -> conv1x1x1 (x4 channels) -> BN (+ shortcut conv -> BN) -> add + ReLU;
          the 1x1x1 convolutions are the same tcgen05 implicit-GEMM kernel with one tap (a plain [positions x Cin] x [Cin x Cout] GEMM)
  head    AvgPool3d(3) -> Linear(F,512) -> Linear(512,1) -> BCEWithLogits (+ full backward)

Backward: ReLU gate -> BN backward -> wgrad (split-K tcgen05) and dgrad (the fprop kernel on dY with flipped weights; for
stride-2 convs on the zero-upsampled dY) -> gradient accumulation at the block input -> ... -> max-pool backward -> stem wgrad.
Reference model: ``ResNet_l3(BasicBlock | Bottleneck, layers)``, fedml_api/model/cv/salient_models.py:13-139 (its hard-coded
9216-feature FC does not fit the ABCD shape; here the FC is sized from the input, see models/volumetric.py).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional

import torch
import torch.nn as nn

from ..models.volumetric import BasicBlock3D, Bottleneck3D, ResNet3D_l3
from ..runtime.slab import ParamSlab, SlabLayout
from . import flat as flat_ops
from . import native
from .engine_b200 import _declare, choose_box


def _od(n, k, s, p):
    return (n + 2 * p - k) // s + 1


class _Conv:
    """One tcgen05 convolution + the BatchNorm that follows it, with every buffer it needs."""
    pass


class ResNet3DEngine:
    native = True

    def __init__(self, model: nn.Module, device, max_batch: int = 8, input_scale: float = 1.0 / 255.0, seed: int = 1234,
                 wgrad_ctas: int = 296):
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("ResNet3DEngine needs a CUDA device")
        if not isinstance(model, ResNet3D_l3):
            raise ImportError("model is not a ResNet3D_l3")
        for layer in (model.layer1, model.layer2, model.layer3):
            for blk in layer:
                if not isinstance(blk, (BasicBlock3D, Bottleneck3D)):
                    raise ImportError("native ResNet path supports BasicBlock3D / Bottleneck3D blocks")
        if model.num_classes != 1:
            raise ImportError("native head supports one logit")
        self.lib = native.load()
        _declare(self.lib)
        self.model = model
        self.in_shape = tuple(model.in_shape)
        self.layout = SlabLayout.from_module(model)
        self.input_scale = float(input_scale)
        self.max_batch = int(max_batch)
        self.seed, self.step = seed, 0
        self.grad = ParamSlab(self.layout, device=self.device, params_only=True)
        self.slab: Optional[ParamSlab] = None
        self.loss_acc = torch.zeros((), dtype=torch.float32, device=self.device)
        self.loss_steps = 0
        self._packed_dirty = True
        self.profile = None
        self._tag = ""
        self.wgrad_ctas = wgrad_ctas
        self._alloc()

    # ----------------------------------------------------------------------------- buffers
    def _new_conv(self, conv: str, bn: str, cin: int, cout: int, k: int, stride: int, pad: int, in_dhw) -> _Conv:
        dev, B, bf, lib = self.device, self.max_batch, torch.bfloat16, self.lib
        c = _Conv()
        c.conv, c.bn, c.cin, c.cout, c.k, c.stride, c.pad, c.in_dhw = conv, bn, cin, cout, k, stride, pad, tuple(in_dhw)
        c.out_dhw = tuple(_od(n, k, stride, pad) for n in in_dhw)
        D, H, W = c.out_dhw
        c.y = torch.empty((B, D, H, W, cout), dtype=bf, device=dev)
        c.dy = torch.empty_like(c.y)
        f32 = dict(dtype=torch.float32, device=dev)
        c.scale, c.shift, c.mean, c.invstd = (torch.zeros(cout, **f32) for _ in range(4))
        c.c12 = torch.zeros(2 * cout, **f32)
        c.bwd_partials = torch.zeros(lib.nidt_bn_bwd_grid() * 2 * cout, **f32)
        if cin == 1:        # stem
            c.stats_tiles = lib.nidt_conv_in_grid()
            c.stats = torch.zeros(c.stats_tiles * 2 * cout, **f32)
            c.wg_scratch = torch.zeros(lib.nidt_conv_in_grid() * 128 * 64, **f32)
            return c
        taps = k ** 3
        c.box = choose_box(W, H, D, B)
        c.m_tiles = lib.nidt_conv3d_m_tiles(B, D, H, W, *c.box)
        c.block_n = 128 if (cout % 128 == 0 and c.m_tiles * (cout // 128) >= 148) else 64
        c.stats_tiles = c.m_tiles
        c.stats = torch.zeros(c.m_tiles * 2 * cout, **f32)
        c.wf = torch.empty((cout, taps, cin), dtype=bf, device=dev)
        c.wd = torch.empty((cin, taps, cout), dtype=bf, device=dev)
        iD, iH, iW = c.in_dhw
        c.dbox = choose_box(iW, iH, iD, B)
        c.dgrad_block_n = 128 if (cin % 128 == 0 and lib.nidt_conv3d_m_tiles(B, iD, iH, iW, *c.dbox) * (cin // 128) >= 148) else 64
        work = taps * (-(-cout // 128)) * (cin // 64)
        c.splits = max(1, min(c.m_tiles, self.wgrad_ctas // work))
        c.wg_scratch = torch.zeros(lib.nidt_conv3d_wgrad_scratch_floats(cout, cin, k, c.splits), **f32)
        if stride > 1:      # zero-upsampled dY for the data gradient
            c.u_dhw = tuple(stride * (n - 1) + 1 for n in c.out_dhw)
            c.u = torch.zeros((B,) + c.u_dhw + (cout,), dtype=bf, device=dev)
        return c

    def _act(self, dhw, ch) -> torch.Tensor:
        return torch.empty((self.max_batch,) + tuple(dhw) + (ch,), dtype=torch.bfloat16, device=self.device)

    def _alloc(self) -> None:
        m = self.model
        dev = self.device
        self.stem = self._new_conv("conv1", "bn1", 1, 64, 3, 2, 3, self.in_shape)
        st = self.stem
        st.r = self._act(st.out_dhw, 64)           # relu(bn(conv))
        st.dr = torch.empty_like(st.r)
        self.pool_dhw = tuple(_od(n, 3, 2, 1) for n in st.out_dhw)
        st.p = self._act(self.pool_dhw, 64)        # max-pooled
        st.dp = torch.empty_like(st.p)
        st.idx = torch.empty(st.p.shape, dtype=torch.uint8, device=dev)
        self.blocks: List[object] = []
        dhw, ch = self.pool_dhw, 64
        for lname in ("layer1", "layer2", "layer3"):
            for bi, blk in enumerate(getattr(m, lname)):
                pre = f"{lname}.{bi}"
                b = _Conv()
                b.name = pre
                planes, stride = blk.conv1.out_channels, blk.stride
                if isinstance(blk, Bottleneck3D):
                    specs = [(planes, 1, 1, 0), (planes, 3, stride, 1), (planes * 4, 1, 1, 0)]
                else:
                    specs = [(planes, 3, stride, 1), (planes, 3, 1, 1)]
                b.main, cin, cdhw = [], ch, dhw                # the main branch: conv -> BN (-> ReLU except after the last one)
                for j, (co, k, s_, p_) in enumerate(specs):
                    c = self._new_conv(f"{pre}.conv{j + 1}", f"{pre}.bn{j + 1}", cin, co, k, s_, p_, cdhw)
                    b.main.append(c)
                    cin, cdhw = co, c.out_dhw
                out_ch = cin
                b.ds = None
                if blk.downsample is not None:
                    b.ds = self._new_conv(pre + ".downsample.0", pre + ".downsample.1", ch, out_ch, 1, stride, 0, dhw)
                    assert b.ds.out_dhw == cdhw
                    b.zd = self._act(b.ds.out_dhw, out_ch)
                else:
                    assert ch == out_ch and dhw == cdhw, "identity shortcut needs equal shapes"
                b.acts = [self._act(c.out_dhw, c.cout) for c in b.main[:-1]]      # relu(bn(conv_j)) for all but the last conv
                b.dacts = [torch.empty_like(a) for a in b.acts]
                b.z = self._act(cdhw, out_ch)                  # bn(last conv)
                b.out = self._act(cdhw, out_ch)
                b.dout = torch.empty_like(b.out)          # gradient w.r.t. the block output (written by the next block / head)
                b.dz = torch.empty_like(b.out)            # gated gradient shared by the main branch and the shortcut
                b.dtmp = self._act(dhw, ch) if b.ds is not None else None
                b.c1, b.c2, b.a1, b.da1, b.z2 = b.main[0], b.main[1], b.acts[0], b.dacts[0], b.z      # names the basic-block tests use
                self.blocks.append(b)
                dhw, ch = cdhw, out_ch
        self.last_dhw, self.last_ch = dhw, ch
        self.avg_k = 3
        S = 1
        for n in dhw:
            S *= n // self.avg_k
        self.F = S * ch
        assert self.F == m.fc.in_features, (self.F, m.fc.in_features)
        self.Hd = m.fc.out_features
        B = self.max_batch
        f32 = dict(dtype=torch.float32, device=dev)
        self.feat, self.dfeat = torch.zeros((B, self.F), **f32), torch.zeros((B, self.F), **f32)
        self.hidden, self.dhidden = torch.zeros(B * self.Hd, **f32), torch.zeros(B * self.Hd, **f32)
        self.eval_logits = torch.zeros(B, **f32)
        self.convs = [self.stem] + [c for b in self.blocks for c in (tuple(b.main) + ((b.ds,) if b.ds is not None else ()))]

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

    def _chk(self, rc: int, what: str, launches: int = 1) -> None:
        native.check(rc, what)
        native.count_launches(launches)
        if self.profile is not None:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record()
            self.profile.append((what + self._tag, ev))

    def start_profile(self) -> None:
        self.profile = []
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        self.profile.append(("<start>", ev))

    def stop_profile(self):
        torch.cuda.synchronize(self.device)
        out = [(n, self.profile[i - 1][1].elapsed_time(e)) for i, (n, e) in enumerate(self.profile) if i > 0]
        self.profile = None
        return out

    # ------------------------------------------------------------------------------ pieces
    def _pack_weights(self) -> None:
        """fp32 master weights -> bf16 GEMM operands for every tcgen05 conv, eight layers per launch."""
        lib, st = self.lib, self._stream()
        convs = self.convs[1:]
        for g0 in range(0, len(convs), 8):
            grp = convs[g0:g0 + 8]
            n = len(grp)
            w = (C.c_void_p * n)(*[self._p(self.slab, c.conv + ".weight") for c in grp])
            wf = (C.c_void_p * n)(*[c.wf.data_ptr() for c in grp])
            wd = (C.c_void_p * n)(*[c.wd.data_ptr() for c in grp])
            co = (C.c_int * n)(*[c.cout for c in grp])
            ci = (C.c_int * n)(*[c.cin for c in grp])
            tp = (C.c_int * n)(*[c.k ** 3 for c in grp])
            self._chk(lib.nidt_pack_conv_weights_multi(n, w, co, ci, tp, wf, wd, st), "pack_conv_weights")
        self._packed_dirty = False

    def _conv_bn(self, c: _Conv, x: torch.Tensor, B: int, train: bool) -> None:
        """c.y = conv(x); BN scale/shift from batch statistics (train) or running statistics (eval)."""
        lib, st, slab = self.lib, self._stream(), self.slab
        D, H, W = c.out_dhw
        iD, iH, iW = c.in_dhw
        stats = c.stats.data_ptr() if train else None
        self._tag = f"[{c.conv}]"
        if c.cin == 1:
            self._chk(lib.nidt_conv_in_fprop(x.data_ptr(), self._p(slab, c.conv + ".weight"), None, c.y.data_ptr(), stats, B, iD, iH, iW,
                                             c.k, c.stride, c.pad, self.input_scale, st), "conv_in_fprop")
            tiles = c.stats_tiles
        else:
            box = c.box if B == self.max_batch else choose_box(W, H, D, B)
            c.cur_box = box
            self._chk(lib.nidt_conv3d_fprop_tc(x.data_ptr(), c.wf.data_ptr(), None, c.y.data_ptr(), stats, B, iD, iH, iW, c.cin, D, H, W,
                                               c.cout, c.k, c.stride, c.pad, *box, c.block_n, st), "conv3d_fprop_tc")
            tiles = c.stats_tiles if B == self.max_batch else lib.nidt_conv3d_m_tiles(B, D, H, W, *box)
        g, be = self._p(slab, c.bn + ".weight"), self._p(slab, c.bn + ".bias")
        rm, rv = self._p(slab, c.bn + ".running_mean"), self._p(slab, c.bn + ".running_var")
        if train:
            self._chk(lib.nidt_bn_finalize(c.stats.data_ptr(), tiles, c.cout, float(B * D * H * W), g, be, rm, rv,
                                           self._p(slab, c.bn + ".num_batches_tracked"), 0.1, 1e-5, c.scale.data_ptr(), c.shift.data_ptr(),
                                           c.mean.data_ptr(), c.invstd.data_ptr(), st), "bn_finalize")
        else:
            self._chk(lib.nidt_bn_eval_prepare(c.cout, g, be, rm, rv, 1e-5, c.scale.data_ptr(), c.shift.data_ptr(), st), "bn_eval_prepare")

    def _bn_apply(self, c: _Conv, out: torch.Tensor, B: int, relu: bool) -> None:
        D, H, W = c.out_dhw
        self._chk(self.lib.nidt_bn_relu_pool_fwd(c.y.data_ptr(), out.data_ptr(), None, B, D, H, W, c.cout, 0, 0, c.scale.data_ptr(),
                                                 c.shift.data_ptr(), int(relu), self._stream()), "bn_act_fwd")

    def _bn_bwd(self, c: _Conv, da: torch.Tensor, B: int, relu: bool) -> None:
        """c.dy = d loss / d conv output, given da = d loss / d (BN [+ReLU] output); BN weight/bias gradients into the slab."""
        D, H, W = c.out_dhw
        self._tag = f"[{c.conv}]"
        self._chk(self.lib.nidt_bn_relu_pool_bwd(c.y.data_ptr(), None, da.data_ptr(), c.dy.data_ptr(), B, D, H, W, c.cout, 0, 0,
                                                 c.scale.data_ptr(), c.shift.data_ptr(), c.mean.data_ptr(), c.invstd.data_ptr(),
                                                 c.bwd_partials.data_ptr(), c.c12.data_ptr(), self._p(self.grad, c.bn + ".weight"),
                                                 self._p(self.grad, c.bn + ".bias"), int(relu), self._stream()), "bn_act_bwd", 3)

    def _conv_bwd(self, c: _Conv, x: torch.Tensor, dx: Optional[torch.Tensor], B: int) -> None:
        """Weight gradient into the slab and (dx given) the data gradient of a tcgen05 conv from c.dy."""
        lib, st = self.lib, self._stream()
        D, H, W = c.out_dhw
        iD, iH, iW = c.in_dhw
        box = c.cur_box
        splits = c.splits if B == self.max_batch else max(1, min(c.splits, lib.nidt_conv3d_m_tiles(B, D, H, W, *box)))
        self._chk(lib.nidt_conv3d_wgrad_tc(c.dy.data_ptr(), x.data_ptr(), self._p(self.grad, c.conv + ".weight"), c.wg_scratch.data_ptr(),
                                           B, iD, iH, iW, c.cin, D, H, W, c.cout, c.k, c.stride, c.pad, *box, splits, st), "conv3d_wgrad_tc", 2)
        if dx is None:
            return
        dbox = c.dbox if B == self.max_batch else choose_box(iW, iH, iD, B)
        src, sD, sH, sW = c.dy, D, H, W
        if c.stride > 1:
            sD, sH, sW = c.u_dhw
            self._chk(lib.nidt_upsample_zero(c.dy.data_ptr(), c.u.data_ptr(), B, D, H, W, c.cout, sD, sH, sW, c.stride, st), "upsample_zero")
            src = c.u
        self._chk(lib.nidt_conv3d_fprop_tc(src.data_ptr(), c.wd.data_ptr(), None, dx.data_ptr(), None, B, sD, sH, sW, c.cout, iD, iH, iW,
                                           c.cin, c.k, 1, c.k - 1 - c.pad, *dbox, c.dgrad_block_n, st), "conv3d_dgrad_tc")

    def _ew(self, a, b, out, B: int, mode: int, what: str) -> None:
        n = B * (a.numel() // self.max_batch)
        self._chk(self.lib.nidt_ew_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), n, mode, self._stream()), what)

    # ------------------------------------------------------------------------------ passes
    def _forward(self, x: torch.Tensor, train: bool) -> int:
        assert x.dtype == torch.uint8 and x.is_cuda and x.dim() == 4, "ResNet3DEngine consumes 8-bit volumes [B,D,H,W]"
        x = x.contiguous()
        B = x.shape[0]
        assert B <= self.max_batch and tuple(x.shape[1:]) == self.in_shape
        lib, st = self.lib, self._stream()
        if self._packed_dirty:
            self._pack_weights()
        self._x = x
        s = self.stem
        self._conv_bn(s, x, B, train)
        self._bn_apply(s, s.r, B, True)
        D, H, W = s.out_dhw
        self._chk(lib.nidt_maxpool3d_fwd(s.r.data_ptr(), s.p.data_ptr(), s.idx.data_ptr() if train else None, B, D, H, W, 64, 3, 2, 1, st),
                  "maxpool3d_fwd")
        cur = s.p
        for b in self.blocks:
            h = cur
            for j, c in enumerate(b.main):
                self._conv_bn(c, h, B, train)
                last = j == len(b.main) - 1
                h = b.z if last else b.acts[j]
                self._bn_apply(c, h, B, not last)
            idt = cur
            if b.ds is not None:
                self._conv_bn(b.ds, cur, B, train)
                self._bn_apply(b.ds, b.zd, B, False)
                idt = b.zd
            self._tag = f"[{b.name}]"
            self._ew(b.z, idt, b.out, B, 0, "add_relu_fwd")
            b.x_in = cur
            cur = b.out
        D, H, W = self.last_dhw
        self._tag = ""
        self._chk(lib.nidt_avgpool3d_fwd(cur.data_ptr(), self.feat.data_ptr(), B, D, H, W, self.last_ch, self.avg_k, st), "avgpool3d_fwd")
        return B

    def _head(self, B: int, y: torch.Tensor, train: bool, counters=None, double_sigmoid=True) -> None:
        slab = self.slab
        gp = (lambda n: self._p(self.grad, n)) if train else (lambda n: None)
        self._chk(self.lib.nidt_linear_head_fwd_bwd(
            self.feat.data_ptr(), y.data_ptr(), self._p(slab, "fc.weight"), self._p(slab, "fc.bias"), self._p(slab, "fc2.weight"),
            self._p(slab, "fc2.bias"), gp("fc.weight"), gp("fc.bias"), gp("fc2.weight"), gp("fc2.bias"),
            self.dfeat.data_ptr() if train else None, self.hidden.data_ptr(), self.dhidden.data_ptr(),
            self.loss_acc.data_ptr() if train else None, counters.buf.data_ptr() if counters is not None else None,
            self.eval_logits.data_ptr(), B, self.F, self.Hd, int(train), int(double_sigmoid), self._stream()), "linear_head", 4 if train else 2)

    def _backward(self, B: int) -> None:
        lib, st = self.lib, self._stream()
        D, H, W = self.last_dhw
        last = self.blocks[-1]
        self._tag = ""
        self._chk(lib.nidt_avgpool3d_bwd(self.dfeat.data_ptr(), last.dout.data_ptr(), B, D, H, W, self.last_ch, self.avg_k, st), "avgpool3d_bwd")
        for i in range(len(self.blocks) - 1, -1, -1):
            b = self.blocks[i]
            dx = self.blocks[i - 1].dout if i > 0 else self.stem.dp        # gradient w.r.t. this block's input
            self._tag = f"[{b.name}]"
            self._ew(b.dout, b.out, b.dz, B, 1, "relu_gate_bwd")
            for j in range(len(b.main) - 1, -1, -1):
                c = b.main[j]
                last = j == len(b.main) - 1
                self._bn_bwd(c, b.dz if last else b.dacts[j], B, not last)
                self._conv_bwd(c, b.acts[j - 1] if j > 0 else b.x_in, b.dacts[j - 1] if j > 0 else dx, B)
            self._tag = f"[{b.name}]"
            if b.ds is not None:
                self._bn_bwd(b.ds, b.dz, B, False)
                self._conv_bwd(b.ds, b.x_in, b.dtmp, B)
                self._ew(dx, b.dtmp, dx, B, 2, "grad_accumulate")
            else:
                self._ew(dx, b.dz, dx, B, 2, "grad_accumulate")
        s = self.stem
        D, H, W = s.out_dhw
        self._tag = "[stem]"
        self._chk(lib.nidt_maxpool3d_bwd(s.dp.data_ptr(), s.idx.data_ptr(), s.dr.data_ptr(), B, D, H, W, 64, 3, 2, 1, st), "maxpool3d_bwd")
        self._bn_bwd(s, s.dr, B, True)
        iD, iH, iW = s.in_dhw
        self._chk(lib.nidt_conv_in_wgrad(self._x.data_ptr(), s.dy.data_ptr(), self._p(self.grad, "conv1.weight"), None, s.wg_scratch.data_ptr(),
                                         B, iD, iH, iW, s.k, s.stride, s.pad, self.input_scale, st), "conv_in_wgrad", 2)

    # ---------------------------------------------------------------------------- public API
    def compute_grads(self, x, y, train_mode: bool = True) -> torch.Tensor:
        if not train_mode:
            raise NotImplementedError("eval-mode gradient screening runs on the TorchEngine")
        y = y.to(self.device, torch.float32).contiguous()
        B = self._forward(x, train=True)
        self._head(B, y, train=True)
        self._backward(B)
        self.step += 1
        return self.loss_acc

    def train_step(self, x, y, hp, mask: Optional[torch.Tensor], mom: Optional[torch.Tensor], first_step: bool,
                   grad_mask: Optional[torch.Tensor] = None) -> None:
        self.compute_grads(x, y, True)
        self.loss_steps += 1
        if grad_mask is not None:
            self.grad.flat.mul_(grad_mask)
        flat_ops.masked_sgd_step(self.slab.params_flat, self.grad.flat, mom, mask, hp.lr, hp.momentum, hp.wd, hp.clip_norm, first_step)
        self._packed_dirty = True

    def pop_mean_loss(self, sync: bool = True):
        mean = self.loss_acc / max(1, self.loss_steps)
        self.loss_acc = torch.zeros_like(self.loss_acc)
        self.loss_steps = 0
        return float(mean.item()) if sync else mean

    def eval_batch(self, x, y, counters, double_sigmoid: bool = True) -> None:
        y = y.to(self.device, torch.float32).contiguous()
        B = self._forward(x, train=False)
        self._head(B, y, train=False, counters=counters, double_sigmoid=double_sigmoid)

    def logits(self, x) -> torch.Tensor:
        B = self._forward(x, train=False)
        yz = torch.zeros(B, dtype=torch.float32, device=self.device)
        self._head(B, yz, train=False)
        return self.eval_logits[:B].clone()
