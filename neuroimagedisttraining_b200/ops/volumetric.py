"""Tensor-level wrappers of the volumetric sm_100a kernels (channels-last bf16 activations).
The training engine (:mod:`.engine_b200`) calls the C ABI directly on preallocated buffers; these
allocate their outputs and exist for tests, notebooks and one-off use."""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import native
from .engine_b200 import _declare, choose_box


def _lib():
    lib = native.load()
    _declare(lib)
    return lib


def _st(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _out_dim(n, k, s, p):
    return (n + 2 * p - k) // s + 1


def pack_conv_weights(w: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """fp32 [Cout,Cin,k,k,k] -> bf16 (wf [Cout,k^3,Cin] for fprop, wd [Cin,k^3,Cout] tap-flipped for dgrad)."""
    lib = _lib()
    Cout, Cin = w.shape[:2]
    taps = w.shape[2] * w.shape[3] * w.shape[4]
    w = w.contiguous().float()
    wf = torch.empty((Cout, taps, Cin), dtype=torch.bfloat16, device=w.device)
    wd = torch.empty((Cin, taps, Cout), dtype=torch.bfloat16, device=w.device)
    native.check(lib.nidt_pack_conv_weights(w.data_ptr(), Cout, Cin, taps, wf.data_ptr(), wd.data_ptr(), _st(w)), "pack")
    return wf, wd


def conv3d_fprop(x: torch.Tensor, wf: torch.Tensor, bias: Optional[torch.Tensor], k: int, stride: int = 1, pad: int = 0,
                 box=None, block_n: Optional[int] = None, stats: bool = False, out_dhw=None):
    """x: bf16 [N,D,H,W,Cin]; wf: bf16 [Cout,k^3,Cin] -> y bf16 [N,Do,Ho,Wo,Cout] (+ partials [m_tiles,2,Cout]).
    ``out_dhw`` overrides the output extent (used when the same kernel runs as dgrad)."""
    lib = _lib()
    N, D, H, W, Cin = x.shape
    Cout = wf.shape[0]
    Do, Ho, Wo = out_dhw if out_dhw is not None else (_out_dim(D, k, stride, pad), _out_dim(H, k, stride, pad), _out_dim(W, k, stride, pad))
    box = box or choose_box(Wo, Ho, Do, N)
    block_n = block_n or (128 if Cout % 128 == 0 else 64)
    y = torch.empty((N, Do, Ho, Wo, Cout), dtype=torch.bfloat16, device=x.device)
    mt = lib.nidt_conv3d_m_tiles(N, Do, Ho, Wo, *box)
    part = torch.zeros((mt, 2, Cout), dtype=torch.float32, device=x.device) if stats else None
    native.check(lib.nidt_conv3d_fprop_tc(x.data_ptr(), wf.data_ptr(), bias.data_ptr() if bias is not None else None, y.data_ptr(),
                                          part.data_ptr() if stats else None, N, D, H, W, Cin, Do, Ho, Wo, Cout, k, stride, pad,
                                          *box, block_n, _st(x)), "conv3d_fprop_tc")
    return (y, part) if stats else y


def upsample_zero(dy: torch.Tensor, s: int) -> torch.Tensor:
    """bf16 [N,D,H,W,C] -> [N,s(D-1)+1,...,C] with s-1 zeros between neighbours (data gradient of strided convs)."""
    lib = _lib()
    N, D, H, W, Cc = dy.shape
    Du, Hu, Wu = s * (D - 1) + 1, s * (H - 1) + 1, s * (W - 1) + 1
    u = torch.empty((N, Du, Hu, Wu, Cc), dtype=torch.bfloat16, device=dy.device)
    native.check(lib.nidt_upsample_zero(dy.data_ptr(), u.data_ptr(), N, D, H, W, Cc, Du, Hu, Wu, s, _st(dy)), "upsample_zero")
    return u


def conv3d_dgrad(dy: torch.Tensor, wd: torch.Tensor, k: int, pad: int, in_dhw, box=None, block_n=None, stride: int = 1) -> torch.Tensor:
    """dy: bf16 [N,Do,Ho,Wo,Cout]; wd: bf16 [Cin,k^3,Cout] -> dx bf16 [N,D,H,W,Cin].  stride > 1: the same stride-1
    kernel over the zero-upsampled dy."""
    if stride > 1:
        dy = upsample_zero(dy, stride)
    return conv3d_fprop(dy, wd, None, k, 1, k - 1 - pad, box=box, block_n=block_n, out_dhw=tuple(in_dhw))


def conv3d_wgrad(dy: torch.Tensor, x: torch.Tensor, k: int, stride: int = 1, pad: int = 0, box=None, splits: int = 1) -> torch.Tensor:
    """-> fp32 [Cout,Cin,k,k,k]"""
    lib = _lib()
    N, Do, Ho, Wo, Cout = dy.shape
    _, D, H, W, Cin = x.shape
    box = box or choose_box(Wo, Ho, Do, N)
    grad = torch.empty((Cout, Cin, k, k, k), dtype=torch.float32, device=x.device)
    scratch = torch.zeros(lib.nidt_conv3d_wgrad_scratch_floats(Cout, Cin, k, splits), dtype=torch.float32, device=x.device)
    native.check(lib.nidt_conv3d_wgrad_tc(dy.data_ptr(), x.data_ptr(), grad.data_ptr(), scratch.data_ptr(), N, D, H, W, Cin, Do, Ho, Wo, Cout,
                                          k, stride, pad, *box, splits, _st(x)), "conv3d_wgrad_tc")
    return grad


def conv_in_fprop(x_u8: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], k: int, stride: int, pad: int, scale: float,
                  stats: bool = False):
    """x_u8: uint8 [N,D,H,W]; w: fp32 [64,1,k,k,k] -> y bf16 [N,Do,Ho,Wo,64] (+ partials [grid,2,64])."""
    lib = _lib()
    N, D, H, W = x_u8.shape
    Do, Ho, Wo = _out_dim(D, k, stride, pad), _out_dim(H, k, stride, pad), _out_dim(W, k, stride, pad)
    y = torch.empty((N, Do, Ho, Wo, 64), dtype=torch.bfloat16, device=x_u8.device)
    part = torch.zeros((lib.nidt_conv_in_grid(), 2, 64), dtype=torch.float32, device=x_u8.device) if stats else None
    w = w.contiguous().float()
    native.check(lib.nidt_conv_in_fprop(x_u8.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None, y.data_ptr(),
                                        part.data_ptr() if stats else None, N, D, H, W, k, stride, pad, scale, _st(y)), "conv_in_fprop")
    return (y, part) if stats else y


def conv_in_wgrad(x_u8: torch.Tensor, dy: torch.Tensor, k: int, stride: int, pad: int, scale: float, with_bias: bool = False):
    """-> dW fp32 [64,1,k,k,k] (and dbias [64] = sum of dy, from the bias column of the same GEMM)."""
    lib = _lib()
    N, D, H, W = x_u8.shape
    grad = torch.empty((64, 1, k, k, k), dtype=torch.float32, device=dy.device)
    gb = torch.empty(64, dtype=torch.float32, device=dy.device) if with_bias else None
    scratch = torch.zeros(lib.nidt_conv_in_grid() * 128 * 64, dtype=torch.float32, device=dy.device)
    native.check(lib.nidt_conv_in_wgrad(x_u8.data_ptr(), dy.data_ptr(), grad.data_ptr(), gb.data_ptr() if with_bias else None,
                                        scratch.data_ptr(), N, D, H, W, k, stride, pad, scale, _st(dy)), "conv_in_wgrad")
    return (grad, gb) if with_bias else grad


def bn_finalize(partials: torch.Tensor, count: float, gamma, beta, running_mean=None, running_var=None, num_batches=None,
                momentum: float = 0.1, eps: float = 1e-5):
    lib = _lib()
    T, _, Cc = partials.shape
    dev = partials.device
    scale, shift, mean, invstd = (torch.empty(Cc, dtype=torch.float32, device=dev) for _ in range(4))
    p = lambda t: t.data_ptr() if t is not None else None
    native.check(lib.nidt_bn_finalize(partials.data_ptr(), T, Cc, float(count), gamma.data_ptr(), beta.data_ptr(), p(running_mean), p(running_var),
                                      p(num_batches), momentum, eps, scale.data_ptr(), shift.data_ptr(), mean.data_ptr(), invstd.data_ptr(),
                                      _st(partials)), "bn_finalize")
    return scale, shift, mean, invstd


def bn_relu_pool_fwd(y: torch.Tensor, scale, shift, pool_k: int = 0, pool_s: int = 0, relu: bool = True, want_ymax: bool = False):
    lib = _lib()
    N, D, H, W, Cc = y.shape
    if pool_k:
        od = ((D - pool_k) // pool_s + 1, (H - pool_k) // pool_s + 1, (W - pool_k) // pool_s + 1)
    else:
        od = (D, H, W)
    a = torch.empty((N,) + od + (Cc,), dtype=torch.bfloat16, device=y.device)
    ymax = torch.empty_like(a) if want_ymax else None
    native.check(lib.nidt_bn_relu_pool_fwd(y.data_ptr(), a.data_ptr(), ymax.data_ptr() if want_ymax else None, N, D, H, W, Cc, pool_k, pool_s,
                                           scale.data_ptr(), shift.data_ptr(), int(relu), _st(y)), "bn_relu_pool_fwd")
    return (a, ymax) if want_ymax else a


def bn_relu_pool_bwd(y, da, scale, shift, mean, invstd, pool_k: int = 0, pool_s: int = 0, relu: bool = True, ymax=None):
    """-> (dy bf16 like y, dgamma, dbeta)"""
    lib = _lib()
    N, D, H, W, Cc = y.shape
    dev = y.device
    dy = torch.empty_like(y)
    partials = torch.zeros(lib.nidt_bn_bwd_grid() * 2 * Cc, dtype=torch.float32, device=dev)
    c12 = torch.empty(2 * Cc, dtype=torch.float32, device=dev)
    dg, db = torch.empty(Cc, dtype=torch.float32, device=dev), torch.empty(Cc, dtype=torch.float32, device=dev)
    native.check(lib.nidt_bn_relu_pool_bwd(y.data_ptr(), ymax.data_ptr() if ymax is not None else None, da.data_ptr(), dy.data_ptr(), N, D, H, W, Cc, pool_k, pool_s, scale.data_ptr(),
                                           shift.data_ptr(), mean.data_ptr(), invstd.data_ptr(), partials.data_ptr(), c12.data_ptr(),
                                           dg.data_ptr(), db.data_ptr(), int(relu), _st(y)), "bn_relu_pool_bwd")
    return dy, dg, db


def conv_in_s2d_fprop(x_u8: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor], scale: float, stats: bool = False):
    """5x5x5 / stride-2 first layer through the space-to-depth kernel (no im2col).  Same contract as :func:`conv_in_fprop`."""
    lib = _lib()
    N, D, H, W = x_u8.shape
    Do, Ho, Wo = _out_dim(D, 5, 2, 0), _out_dim(H, 5, 2, 0), _out_dim(W, 5, 2, 0)
    y = torch.empty((N, Do, Ho, Wo, 64), dtype=torch.bfloat16, device=x_u8.device)
    part = torch.zeros((lib.nidt_conv_in_grid(), 2, 64), dtype=torch.float32, device=x_u8.device) if stats else None
    w = w.contiguous().float()
    native.check(lib.nidt_conv_in_s2d_fprop(x_u8.data_ptr(), w.data_ptr(), bias.data_ptr() if bias is not None else None, y.data_ptr(),
                                            part.data_ptr() if stats else None, N, D, H, W, scale, _st(y)), "conv_in_s2d_fprop")
    return (y, part) if stats else y


def conv_in_s2d_wgrad(x_u8: torch.Tensor, dy: torch.Tensor, scale: float) -> torch.Tensor:
    lib = _lib()
    N, D, H, W = x_u8.shape
    grad = torch.empty((64, 1, 5, 5, 5), dtype=torch.float32, device=dy.device)
    scratch = torch.zeros(lib.nidt_conv_in_s2d_wgrad_scratch_floats(), dtype=torch.float32, device=dy.device)
    native.check(lib.nidt_conv_in_s2d_wgrad(x_u8.data_ptr(), dy.data_ptr(), grad.data_ptr(), None, scratch.data_ptr(), N, D, H, W, scale, _st(dy)),
                 "conv_in_s2d_wgrad")
    return grad


# ------------------------------------------------------------------ residual-network glue (csrc/resnet_ops.cu)
def ew_bf16(a: torch.Tensor, b: torch.Tensor, mode: int, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """mode 0: relu(a+b); 1: where(b>0, a, 0); 2: a+b (bf16, numel % 8 == 0)."""
    lib = _lib()
    out = torch.empty_like(a) if out is None else out
    native.check(lib.nidt_ew_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), a.numel(), mode, _st(a)), "ew_bf16")
    return out


def maxpool3d_fwd(x: torch.Tensor, k: int, s: int, p: int):
    """bf16 [N,D,H,W,C] -> (out bf16, idx uint8) with overlapping windows / padding."""
    lib = _lib()
    N, D, H, W, Cc = x.shape
    od = (_out_dim(D, k, s, p), _out_dim(H, k, s, p), _out_dim(W, k, s, p))
    out = torch.empty((N,) + od + (Cc,), dtype=torch.bfloat16, device=x.device)
    idx = torch.empty((N,) + od + (Cc,), dtype=torch.uint8, device=x.device)
    native.check(lib.nidt_maxpool3d_fwd(x.data_ptr(), out.data_ptr(), idx.data_ptr(), N, D, H, W, Cc, k, s, p, _st(x)), "maxpool3d_fwd")
    return out, idx


def maxpool3d_bwd(dout: torch.Tensor, idx: torch.Tensor, in_dhw, k: int, s: int, p: int) -> torch.Tensor:
    lib = _lib()
    N, Cc = dout.shape[0], dout.shape[-1]
    D, H, W = in_dhw
    dx = torch.empty((N, D, H, W, Cc), dtype=torch.bfloat16, device=dout.device)
    native.check(lib.nidt_maxpool3d_bwd(dout.data_ptr(), idx.data_ptr(), dx.data_ptr(), N, D, H, W, Cc, k, s, p, _st(dout)), "maxpool3d_bwd")
    return dx


def avgpool3d_fwd(x: torch.Tensor, k: int) -> torch.Tensor:
    """bf16 [N,D,H,W,C] -> fp32 [N, C*S] in NCDHW flatten order."""
    lib = _lib()
    N, D, H, W, Cc = x.shape
    S = (D // k) * (H // k) * (W // k)
    feat = torch.empty((N, Cc * S), dtype=torch.float32, device=x.device)
    native.check(lib.nidt_avgpool3d_fwd(x.data_ptr(), feat.data_ptr(), N, D, H, W, Cc, k, _st(x)), "avgpool3d_fwd")
    return feat


def avgpool3d_bwd(dfeat: torch.Tensor, in_shape, k: int) -> torch.Tensor:
    lib = _lib()
    N, D, H, W, Cc = in_shape
    dx = torch.empty((N, D, H, W, Cc), dtype=torch.bfloat16, device=dfeat.device)
    native.check(lib.nidt_avgpool3d_bwd(dfeat.data_ptr(), dx.data_ptr(), N, D, H, W, Cc, k, _st(dfeat)), "avgpool3d_bwd")
    return dx


def linear_head_fwd_bwd(feat, y, W1, b1, W2, b2, train: bool = True, counters=None, double_sigmoid: bool = True):
    """-> dict(logits, loss (train), gW1, gb1, gW2, gb2, dfeat (train))"""
    lib = _lib()
    B, Fdim = feat.shape
    Hd = W1.shape[0]
    dev = feat.device
    f32 = dict(dtype=torch.float32, device=dev)
    hidden, dh = torch.empty(B * Hd, **f32), torch.empty(B * Hd, **f32)
    logits = torch.empty(B, **f32)
    loss = torch.zeros((), **f32)
    g = {}
    if train:
        g = {"gW1": torch.empty_like(W1), "gb1": torch.empty_like(b1), "gW2": torch.empty_like(W2), "gb2": torch.empty_like(b2),
             "dfeat": torch.empty_like(feat)}
    p = lambda n: g[n].data_ptr() if train else None
    native.check(lib.nidt_linear_head_fwd_bwd(feat.data_ptr(), y.data_ptr(), W1.data_ptr(), b1.data_ptr(), W2.data_ptr(), b2.data_ptr(),
                                              p("gW1"), p("gb1"), p("gW2"), p("gb2"), p("dfeat"), hidden.data_ptr(), dh.data_ptr(),
                                              loss.data_ptr() if train else None, counters.data_ptr() if counters is not None else None,
                                              logits.data_ptr(), B, Fdim, Hd, int(train), int(double_sigmoid), _st(feat)), "linear_head")
    g.update(logits=logits, loss=loss, hidden=hidden.view(B, Hd))
    return g


def conv3d_halo_fprop(x: torch.Tensor, wf: torch.Tensor, bias: Optional[torch.Tensor], pad: int, block_n: Optional[int] = None,
                      stats: bool = False, debug_flags: int = 0):
    """3x3x3 / stride-1 conv through the halo-resident kernel (csrc/conv_halo.cu).  Same contract as :func:`conv3d_fprop`;
    the partials have ``nidt_conv3d_halo_stats_rows()`` rows.  dgrad: pass ``wd`` and ``pad = 2 - pad``."""
    lib = _lib()
    N, D, H, W, Cin = x.shape
    Cout = wf.shape[0]
    Do, Ho, Wo = D + 2 * pad - 2, H + 2 * pad - 2, W + 2 * pad - 2
    block_n = block_n or (128 if Cout % 128 == 0 else 64)
    y = torch.empty((N, Do, Ho, Wo, Cout), dtype=torch.bfloat16, device=x.device)
    part = torch.empty((lib.nidt_conv3d_halo_stats_rows(), 2, Cout), dtype=torch.float32, device=x.device) if stats else None
    native.check(lib.nidt_conv3d_halo_fprop(x.data_ptr(), wf.data_ptr(), bias.data_ptr() if bias is not None else None, y.data_ptr(),
                                            part.data_ptr() if stats else None, N, D, H, W, Cin, Do, Ho, Wo, Cout, pad, block_n, debug_flags,
                                            _st(x)), "conv3d_halo_fprop")
    return (y, part) if stats else y
