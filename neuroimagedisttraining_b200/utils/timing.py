"""Device-timed phases (CUDA events on the launching stream; wall clock on CPU).  The reference has
no timers at all."""
from __future__ import annotations

import time
from collections import defaultdict
from contextlib import contextmanager

import torch


class PhaseTimer:
    def __init__(self, device):
        self.cuda = torch.device(device).type == "cuda"
        self.totals = defaultdict(float)
        self.counts = defaultdict(int)
        self._pending = []

    @contextmanager
    def phase(self, name: str):
        if self.cuda:
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                torch.cuda.nvtx.range_push(name)
                yield
            finally:
                torch.cuda.nvtx.range_pop()
                b.record()
                self._pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self.totals[name] += time.perf_counter() - t0
                self.counts[name] += 1

    def flush(self):
        if self._pending:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.totals[name] += a.elapsed_time(b) / 1e3
                self.counts[name] += 1
            self._pending = []

    def summary(self):
        self.flush()
        return {k: {"total_s": v, "calls": self.counts[k]} for k, v in self.totals.items()}
