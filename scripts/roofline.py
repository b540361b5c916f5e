#!/usr/bin/env python
"""Roofline table of one AlexNet3D training step from a per-kernel profile (``scripts/profile_step.py`` output) and the measured
peaks of the box (``MEASURED_PEAKS.json``).  Runs without a GPU: it only combines committed measurements with the layer shapes.

    python scripts/roofline.py profiles/profile_step_v18.txt [--batch 16] [--out profiles/roofline_step_v18.md]

Per kernel: FLOPs of the GEMM-shaped work (2*M*N*K), the bytes that MUST cross HBM (compulsory traffic: every input read once,
every output written once; bf16 activations, 8-bit input volume, fp32 weight gradients), the time each would take at the measured
bf16 rate / copy bandwidth, and the achieved fraction of the binding one.  Compulsory traffic is a lower bound of what the kernel
really moves, so a fraction is an upper bound of how close the kernel is to ITS roofline only when re-reads are absent -- the ncu
DRAM counters in ``profiles/ncu_all_kernels_v2.md`` give the real bytes.  ``conv_in_alg_wgrad`` (first-layer weight gradient from pooled
tensors) is CUDA-core work bound by shared-memory loads: its row compares against fp32 FMA issue and HBM only for orientation.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def layer_table(batch: int, in_shape=(121, 145, 121)):
    from neuroimagedisttraining_b200.models import AlexNet3D_Dropout
    from neuroimagedisttraining_b200.models.volumetric import trace_shapes
    model = AlexNet3D_Dropout(1, in_shape=in_shape)
    rows, prev = [], (1,) + tuple(in_shape)
    for spec, (conv_out, pool_out) in zip(model.plan, trace_shapes(model.plan, in_shape)):
        vox_in = prev[1] * prev[2] * prev[3]
        vox_conv = conv_out[0] * conv_out[1] * conv_out[2]
        vox_pool = pool_out[0] * pool_out[1] * pool_out[2]
        rows.append(dict(cin=spec.cin, cout=spec.cout, k=spec.k, M=batch * vox_conv, K=spec.cin * spec.k ** 3,
                         in_elems=batch * vox_in * spec.cin, conv_elems=batch * vox_conv * spec.cout,
                         pool_elems=batch * vox_pool * spec.cout, pooled=spec.pool_k > 0))
        prev = (spec.cout,) + tuple(pool_out)
    return rows


def kernel_model(name: str, idx, L):
    """-> (flops, compulsory bytes) of one launch; None when the kernel is latency-sized (no meaningful roofline)."""
    if idx is None or idx >= len(L):
        return None
    r = L[idx]
    gemm = 2.0 * r["M"] * r["cout"] * r["K"]
    in_b = r["in_elems"] * (1 if idx == 0 else 2)          # block 0 reads the 8-bit volume
    w_b = r["cout"] * r["K"] * 2
    if name.startswith(("conv_in_s2d_fprop", "conv3d_fprop_tc")):
        return gemm, in_b + w_b + r["conv_elems"] * 2
    if name.startswith("conv3d_dgrad_tc"):
        return gemm, r["conv_elems"] * 2 + w_b + r["in_elems"] * 2
    if name.startswith(("conv3d_wgrad_tc", "conv_in_s2d_wgrad")):
        return gemm, r["conv_elems"] * 2 + in_b + r["cout"] * r["K"] * 4
    if name.startswith("bn_relu_pool_fwd"):
        out = r["pool_elems"] * 2 * (2 if r["pooled"] else 1)          # pooled value + the pre-pool maximum kept for backward
        return 0.0, r["conv_elems"] * 2 + out
    if name.startswith("bn_relu_pool_bwd"):
        return 0.0, r["conv_elems"] * 2 * 2 + r["pool_elems"] * 2 * 2  # read y, write dy (conv resolution); read pooled dy + max
    if name.startswith("bn_pool3_bwd_stats"):
        return 0.0, r["pool_elems"] * 2 * 2                           # the reductions read the pooled winner values and gradients only
    if name.startswith("conv_in_alg_wgrad"):
        # one 125-tap patch per pooling window and channel; the input volume (8 bit), winner taps (8 bit) and pooled gradients
        # (bf16) are read once.  The kernel's real bound is its shared-memory loads (75 per patch), see ncu_conv_in_alg_sparse.txt.
        return 2.0 * r["pool_elems"] * r["K"], r["in_elems"] + r["pool_elems"] * 3
    return None


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("profile")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--peaks", default=os.path.join(ROOT, "MEASURED_PEAKS.json"))
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    peaks = json.load(open(a.peaks)) if os.path.exists(a.peaks) else {}
    bw = float(peaks.get("hbm_gbs", 6542.0)) * 1e9
    tf = float(peaks.get("bf16_tflops_sustained", 1368.0)) * 1e12
    L = layer_table(a.batch)
    lines = [f"# Roofline of one AlexNet3D step at batch {a.batch} (`{os.path.basename(a.profile)}`)", "",
             f"Denominators: copy bandwidth {bw / 1e9:.0f} GB/s, sustained bf16 {tf / 1e12:.0f} TFLOP/s (`MEASURED_PEAKS.json`).  "
             "Bytes are compulsory HBM traffic (lower bound).", "",
             "| kernel | ms | GFLOP | MB (compulsory) | t_math ms | t_mem ms | bound | achieved / roofline |", "|---|---|---|---|---|---|---|---|"]
    total = modelled = 0.0
    for line in open(a.profile):
        m = re.match(r"^(\w+)(?:\[(\d+)\])?\s+([\d.]+)\s+([\d.]+)\s*$", line)
        if not m:
            continue
        name, idx, tmin = m.group(1), (int(m.group(2)) if m.group(2) else None), float(m.group(3))
        total += tmin
        km = kernel_model(name, idx, L)
        label = name + (f"[{idx}]" if idx is not None else "")
        if km is None:
            lines.append(f"| `{label}` | {tmin:.4f} | - | - | - | - | latency | - |")
            continue
        flops, nbytes = km
        t_math, t_mem = flops / tf * 1e3, nbytes / bw * 1e3
        roof = max(t_math, t_mem)
        modelled += tmin
        if name.startswith("conv_in_alg_wgrad"):
            # CUDA-core kernel: 75 32-bit shared-memory loads per (window, channel) patch, one 32-lane wavefront per clock and SM
            sm_clk = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
            t_smem = L[idx]["pool_elems"] * 75 / 32 / 148 / sm_clk * 1e3
            lines.append(f"| `{label}` | {tmin:.4f} | {flops / 1e9:.1f} (fp32 FMA) | {nbytes / 1e6:.1f} | - | {t_mem:.4f} (HBM), {t_smem:.4f} (shared-memory loads) | "
                         f"shared memory | {t_smem / tmin:.2f} |")
            continue
        lines.append(f"| `{label}` | {tmin:.4f} | {flops / 1e9:.1f} | {nbytes / 1e6:.1f} | {t_math:.4f} | {t_mem:.4f} | "
                     f"{'math' if t_math >= t_mem else 'memory'} | {roof / tmin:.2f} |")
    lines += ["", f"Sum of kernel minima {total:.3f} ms, of which {modelled:.3f} ms in kernels with a bandwidth / math roofline; the rest is "
              "launch-latency-sized (BN finalise, head, weight re-pack, blocks 2-4 at 2 800 output rows)."]
    text = "\n".join(lines) + "\n"
    if a.out:
        open(a.out, "w").write(text)
    print(text)
    return text


if __name__ == "__main__":
    main()
