"""ctypes binding of ``lib/libnidt_b200.so`` (C ABI declared in ``include/nidt_b200.h``).

The library is built in-tree by ``build_native.py`` and loaded lazily.  On a machine with a
GPU a missing library is a hard error -- there is no silent PyTorch fallback for CUDA tensors.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path
from typing import Optional

import torch

_LIB: Optional[C.CDLL] = None
_LIB_PATH = Path(__file__).resolve().parent.parent / "lib" / "libnidt_b200.so"

c_f32p = C.c_void_p
c_stream = C.c_void_p


class NativeError(RuntimeError):
    pass


def lib_path() -> Path:
    return Path(os.environ.get("NIDT_B200_LIB", str(_LIB_PATH)))


def available() -> bool:
    return lib_path().exists()


def _declare(lib: C.CDLL) -> None:
    i64, f32, i32, u64, vp = C.c_int64, C.c_float, C.c_int, C.c_uint64, C.c_void_p
    lib.nidt_last_error.restype = C.c_char_p
    sigs = {
        "nidt_abi_version": [],
        "nidt_norm_scratch_floats": [],
        "nidt_topk_scratch_bytes": [],
        "nidt_max_agg_sources": [],
        "nidt_masked_sgd_step": [vp, vp, vp, vp, i64, f32, f32, f32, f32, i32, vp, vp, vp],
        "nidt_grad_norm": [vp, i64, vp, vp, vp],
        "nidt_weighted_aggregate": [vp, C.POINTER(vp), C.POINTER(f32), i32, vp, i64, i64, vp],
        "nidt_saliency_accumulate": [vp, vp, vp, i64, f32, vp],
        "nidt_global_topk_mask": [vp, vp, vp, i64, u64, vp, vp],
        "nidt_topk_read_result": [vp, C.POINTER(f32), C.POINTER(f32), C.POINTER(C.c_uint), vp],
        "nidt_count_nonzero": [vp, i64, vp, vp],
        "nidt_hamming": [vp, vp, i64, vp, vp],
        "nidt_sq_dist": [vp, vp, i64, vp, vp, vp],
        "nidt_prox_pull": [vp, vp, i64, f32, vp],
    }
    for name, argtypes in sigs.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = C.c_int
    # optional symbol groups are declared by the modules that use them (engine, comm)


def load() -> C.CDLL:
    global _LIB
    if _LIB is None:
        p = lib_path()
        if not p.exists():
            raise NativeError(
                f"native library {p} is missing; run `python -m neuroimagedisttraining_b200.build_native` "
                "(or __graft_entry__.build()) -- CUDA tensors have no PyTorch fallback path")
        lib = C.CDLL(str(p), mode=C.RTLD_GLOBAL)
        _declare(lib)
        _LIB = lib
    return _LIB


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().nidt_last_error().decode(errors="replace")
        raise NativeError(f"{what or 'native call'} failed (rc={rc}): {msg}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "native ops need contiguous CUDA tensors"
    return t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


LAUNCH_COUNT = 0  # kernels of ours launched through this binding (bench.py's ``gpu_launches``)


def count_launches(n: int) -> None:
    global LAUNCH_COUNT
    LAUNCH_COUNT += n
