#!/usr/bin/env python
"""Latency / bandwidth model of the aggregation sweeps (``profiles/agg_sweep_n*_v2.jsonl``): least-squares fit of
``t(S) = alpha + S / beta`` per variant and GPU count.  ``alpha`` is what the FL-sized (10 MB) aggregation pays -- the two cross-rank
flag barriers plus launch -- and ``beta`` the asymptotic algorithmic bandwidth, to be compared with the NVLink roofline
``770 GB/s * P / (P - 1)`` (each rank moves S (P-1)/P bytes per direction).  Runs without a GPU.

    python scripts/agg_model.py [--out profiles/agg_model.md]
"""
from __future__ import annotations

import argparse
import glob
import json
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
VARIANTS = (("fused P2P", "fused_p2p_ms"), ("fused NVLS multimem", "fused_multimem_ms"), ("NCCL all_reduce + mask + copy", "nccl_allreduce_mask_copy_ms"))
LINK_GBS = 770.0      # measured peer bandwidth per direction


def fit(sizes_mb, times_ms):
    """Weighted least squares (relative error): alpha in ms, beta in GB/s."""
    s, t = np.asarray(sizes_mb, float) * 1e6, np.asarray(times_ms, float) * 1e-3
    A = np.stack([np.ones_like(s), s], 1) / t[:, None]
    (alpha, inv_beta), *_ = np.linalg.lstsq(A, np.ones_like(t), rcond=None)
    return alpha * 1e3, 1e-9 / inv_beta


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--out", default=None)
    a = ap.parse_args(argv)
    lines = ["# Aggregation sweep: t(S) = alpha + S / beta", "",
             "| GPUs | variant | alpha (us) | beta (GB/s, algorithmic) | NVLink roofline beta (GB/s) | beta / roofline | implied link rate per direction (GB/s) | t(10 MB) us |",
             "|---|---|---|---|---|---|---|---|"]
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "agg_sweep_n*_v2.jsonl")), key=lambda p: int(p.split("_n")[1].split("_")[0])):
        rows = [json.loads(l) for l in open(path) if l.strip()]
        P = rows[0]["n_gpus"]
        roof = LINK_GBS * P / (P - 1)
        for label, key in VARIANTS:
            pts = [(r["size_mb"], r[key]) for r in rows if r.get(key)]
            if len(pts) < 3:
                continue
            alpha, beta = fit(*zip(*pts))
            t10 = dict(pts).get(10)
            link = beta * (P - 1) / P                      # bytes one rank pulls (and pushes) over NVLink per second of slope time
            lines.append(f"| {P} | {label} | {alpha * 1e3:.0f} | {beta:.0f} | {roof:.0f} | {beta / roof:.2f} | {link:.0f} | "
                         + (f"{t10 * 1e3:.0f} |" if t10 else "- |"))
    lines += ["", "The implied per-direction link rate is nearly the same at 2, 4 and 8 GPUs (about 215-260 GB/s of the 770 GB/s a plain peer-read "
              "kernel reaches): the slope is set by how one GPU issues its remote loads and stores, not by the number of peers or the fabric.  "
              "First experiments for the next GPU session: the reduce phase alone (peer loads, local stores) and the broadcast phase alone (local "
              "loads, peer stores) at the same grid, to see which side caps; then larger per-thread batches of loads before the first store.", "",
              "Reading: at the FL model size the step is all alpha (barrier + launch latency), where the fused kernels lead; the large-S slope "
              "is where the 60 %-of-roofline target is missed (DESIGN.md §6 item 4 has the plan: bulk-async staging of peer chunks, split CTAs "
              "between the reduce-scatter and all-gather phases)."]
    text = "\n".join(lines) + "\n"
    if a.out:
        open(a.out, "w").write(text)
    print(text)
    return text


if __name__ == "__main__":
    main()
