"""``nidt_conv3d_wgrad_tc_static`` -- the weight-gradient entry point for callers that cannot allocate.

Some users of the library hold device pointers but cannot get more device memory where the call is
made: the body of a custom call inside another framework's executor (the executor owns the allocator),
or a service whose memory is partitioned once at start-up.  For them the split-K scratch of the conv
weight-gradient lives inside the library.  This file drives that entry point through the C ABI exactly
as such a caller would (ctypes, raw pointers, no scratch argument) and checks it against the ordinary
entry point, twice in a row with different shapes so the second call runs on whatever the first left in
the shared array.  It is collected last on purpose: if it ever breaks, ``pytest -x`` has already run
everything else.
"""
import ctypes as C

import pytest
import torch

from neuroimagedisttraining_b200.ops import native

ARGS = [C.c_void_p, C.c_void_p, C.c_void_p] + [C.c_int] * 17 + [C.c_void_p]


def _lib():
    lib = native.load()
    lib.nidt_conv3d_wgrad_tc_static.argtypes = ARGS
    lib.nidt_conv3d_wgrad_tc_static.restype = C.c_int
    lib.nidt_conv3d_wgrad_static_capacity_floats.argtypes = []
    lib.nidt_conv3d_wgrad_static_capacity_floats.restype = C.c_int64
    return lib


def test_static_wgrad_rejects_oversize_without_touching_the_gpu():
    """The capacity check comes before any CUDA call, so it also runs on a machine without a GPU."""
    lib = _lib()
    cap = lib.nidt_conv3d_wgrad_static_capacity_floats()
    assert cap == 4 << 20
    # 8 splits x 27 taps x 192 x 192 floats = 7.96 Mi > 4 Mi
    rc = lib.nidt_conv3d_wgrad_tc_static(None, None, None, 16, 5, 7, 5, 192, 5, 7, 5, 192, 3, 1, 1, 5, 7, 1, 3, 8, None)
    assert rc != 0
    msg = lib.nidt_last_error().decode()
    assert "built-in scratch" in msg and "7962624" in msg, msg


@pytest.mark.gpu
def test_static_wgrad_matches_regular_entry_point_twice():
    from neuroimagedisttraining_b200.ops import volumetric as V
    from neuroimagedisttraining_b200.ops.engine_b200 import choose_box
    lib = _lib()
    dev = "cuda"
    st = torch.cuda.current_stream().cuda_stream
    torch.manual_seed(11)
    # (N, Cin, D, H, W, Cout, k, pad, splits): conv2-like with 3 splits, then conv4-like with 1 split
    for (N, Cin, D, H, W, Cout, k, pad, splits) in [(2, 64, 19, 23, 19, 128, 3, 0, 3), (3, 192, 5, 7, 5, 192, 3, 1, 1)]:
        x = torch.randn(N, D, H, W, Cin, device=dev).bfloat16()
        Do, Ho, Wo = D + 2 * pad - k + 1, H + 2 * pad - k + 1, W + 2 * pad - k + 1
        dy = torch.randn(N, Do, Ho, Wo, Cout, device=dev).bfloat16()
        box = choose_box(Wo, Ho, Do, N)
        ref = V.conv3d_wgrad(dy, x, k, 1, pad, box=box, splits=splits)          # caller-provided scratch
        got = torch.full((Cout, Cin, k, k, k), float("nan"), dtype=torch.float32, device=dev)
        rc = lib.nidt_conv3d_wgrad_tc_static(dy.data_ptr(), x.data_ptr(), got.data_ptr(), N, D, H, W, Cin, Do, Ho, Wo, Cout,
                                             k, 1, pad, *box, splits, st)
        assert rc == 0, lib.nidt_last_error().decode()
        torch.cuda.synchronize()
        assert torch.isfinite(got).all()
        # both entry points run the same kernels in the same order, only the scratch address differs: bit-identical output
        # (anything less would mean partials left over from the previous call leaked into this one)
        assert torch.equal(got, ref), f"static vs regular wgrad: max err {(got - ref).abs().max().item():.4g}"
        free0 = torch.cuda.mem_get_info()[0]
        rc = lib.nidt_conv3d_wgrad_tc_static(dy.data_ptr(), x.data_ptr(), got.data_ptr(), N, D, H, W, Cin, Do, Ho, Wo, Cout,
                                             k, 1, pad, *box, splits, st)
        torch.cuda.synchronize()
        assert rc == 0 and torch.cuda.mem_get_info()[0] == free0, "the call changed the amount of free device memory"
        # and against PyTorch fp32, with the tolerance the other conv tests use
        xf = x.float().permute(0, 4, 1, 2, 3).contiguous()
        wz = torch.zeros(Cout, Cin, k, k, k, device=dev, requires_grad=True)
        torch.nn.functional.conv3d(xf, wz, None, 1, pad).backward(dy.float().permute(0, 4, 1, 2, 3).contiguous())
        s2 = wz.grad.abs().max().item() + 1e-12
        assert (got - wz.grad).abs().max().item() <= 1e-2 * s2
