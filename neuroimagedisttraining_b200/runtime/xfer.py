"""Byte counters for the copies that cross PCIe on the FL hot path.

``H2D`` is kept by the stagers (``BatchStager.h2d_bytes``); device->host reads (epoch losses, evaluation
counters, reduced metrics) go through :func:`to_host` so ``bench.py`` reports what was actually copied
instead of a formula.  The reference reads three scalars per evaluation batch and one loss per SGD step
(``sailentgrads/my_model_trainer.py:226,270-273``).
"""
from __future__ import annotations

import torch

D2H_BYTES = 0
D2H_COPIES = 0


def to_host(t: torch.Tensor) -> torch.Tensor:
    """``t.cpu()`` with the bytes counted when ``t`` lives on an accelerator."""
    global D2H_BYTES, D2H_COPIES
    if t.device.type != "cpu":
        D2H_BYTES += t.numel() * t.element_size()
        D2H_COPIES += 1
        return t.cpu()
    return t
