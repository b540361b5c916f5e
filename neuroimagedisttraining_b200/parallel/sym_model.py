"""Host-side model of the fused symmetric-memory aggregation (``csrc/sym_agg.cu``): an in-memory "fake communicator" for W virtual
ranks and a checker of the kernel's flag protocol.

Two uses:

* :class:`FakeSymmetricWorld` / :func:`fake_reduce_mask_broadcast` run the kernel's algorithm -- chunk ``r`` of the flat slab is
  owned by rank ``r``, summed over the ranks' partials in rank order, multiplied by the mask prefix and stored into every rank's
  result buffer -- on CPU tensors, so the orchestration around the kernel (staging of partials, masks shorter than the slab,
  padding, epochs) is testable without GPUs and the expected bits of the P2P path are available to the multi-GPU test.
* :func:`check_flag_protocol` executes the per-rank programme of ``sym_agg_kernel`` (entry barrier ``2e+1`` -> reduce / broadcast ->
  exit barrier ``2e+2`` -> write-back; flags are monotonically increasing epochs compared as ``(int32)(flag - value) >= 0``) as
  interleaved atomic steps under a seeded random scheduler, over several back-to-back launches, with every buffer element stamped
  by the epoch that produced it.  It reports reads of data that is not yet (or no longer) the current epoch's, i.e. exactly the
  races the two barriers exist to exclude; dropping either barrier in the model makes the checker fail, and so does a flag
  comparison that is not wrap-safe once ``2e+2`` passes 2^32.
"""
from __future__ import annotations

import random
from dataclasses import dataclass, field
from typing import Iterator, List, Optional

import numpy as np
import torch

K_MAX_RANKS = 16            # csrc/sym_agg.cu: kMaxRanks (flag pad stride)


def chunk_bounds(n: int, world: int, rank: int) -> tuple:
    """Element range ``[lo, hi)`` rank ``rank`` reduces: the kernel splits the ``n/4`` float4 vectors into ``world`` equal parts."""
    assert n % 4 == 0
    n4 = n // 4
    per = (n4 + world - 1) // world
    lo = min(per * rank, n4)
    return 4 * lo, 4 * min(n4, lo + per)


# ------------------------------------------------------------------------------------------- fake communicator
class FakeSymmetricWorld:
    """``partial`` / ``result`` buffers of W virtual ranks, laid out like ``SymmetricAggregator`` (result padded to 1024 floats)."""

    def __init__(self, world: int, numel: int):
        assert 1 <= world <= K_MAX_RANKS and numel % 4 == 0
        self.world, self.numel = world, numel
        self.pad = (numel + 1023) // 1024 * 1024
        self.buf = [torch.zeros(2 * self.pad, dtype=torch.float32) for _ in range(world)]
        self.epoch = 0

    def partial(self, r: int) -> torch.Tensor:
        return self.buf[r][:self.numel]

    def result(self, r: int) -> torch.Tensor:
        return self.buf[r][self.pad:self.pad + self.numel]


def fake_reduce_mask_broadcast(w: FakeSymmetricWorld, mask: Optional[torch.Tensor] = None) -> List[torch.Tensor]:
    """What one launch of ``sym_agg_kernel`` on every rank leaves in the result buffers (P2P path: fp32 adds in rank order)."""
    w.epoch += 1
    n_mask = 0 if mask is None else min(mask.numel(), w.numel)
    for r in range(w.world):
        lo, hi = chunk_bounds(w.numel, w.world, r)
        if hi <= lo:
            continue
        acc = w.partial(0)[lo:hi].clone()
        for p in range(1, w.world):
            acc += w.partial(p)[lo:hi]
        if n_mask > lo:
            m_hi = min(hi, n_mask)
            acc[:m_hi - lo] *= mask[lo:m_hi]
        for p in range(w.world):
            w.result(p)[lo:hi] = acc
    return [w.result(r).clone() for r in range(w.world)]


# ------------------------------------------------------------------------------------------- protocol checker
@dataclass
class ProtocolReport:
    ok: bool
    launches: int
    steps: int
    violations: List[str] = field(default_factory=list)

    def __bool__(self):
        return self.ok


def _i32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & 0x80000000 else x


class _World:
    def __init__(self, world: int, chunks_per_rank: int):
        self.world = world
        self.flags = np.zeros((world, K_MAX_RANKS), dtype=np.uint64)          # flags[dst][src], slot 0
        self.partial_epoch = [-1] * world                                       # which launch's partial a rank's buffer holds
        self.result_epoch = [[-1] * (world * chunks_per_rank) for _ in range(world)]
        self.violations: List[str] = []


def _rank_program(w: _World, rank: int, first_epoch: int, launches: int, chunks_per_rank: int, entry_barrier: bool, exit_barrier: bool,
                  wrap_safe: bool) -> Iterator[bool]:
    """Yields after every atomic step; yields False while blocked on a flag (the scheduler may then run someone else)."""
    def passed(dst_src_flag: int, value: int) -> bool:
        if wrap_safe:
            return _i32(int(dst_src_flag) - value) >= 0
        return int(dst_src_flag) >= (value & 0xFFFFFFFF)

    def barrier(value: int):
        for peer in range(w.world):                      # st.release.sys into every peer's pad
            w.flags[peer][rank] = value & 0xFFFFFFFF
            yield True
        for peer in range(w.world):                      # ld.acquire.sys on my own pad
            while not passed(w.flags[rank][peer], value):
                yield False
            yield True

    for launch in range(launches):
        e = first_epoch + launch
        # stream-ordered producer before the kernel: this rank's weighted partial of launch `launch`
        w.partial_epoch[rank] = launch
        yield True
        if entry_barrier:
            yield from barrier(2 * e + 1)
        for c in range(chunks_per_rank):                 # reduce my chunk piece by piece, broadcast each piece
            for p in range(w.world):
                if w.partial_epoch[p] != launch:
                    w.violations.append(f"launch {launch}: rank {rank} read rank {p}'s partial holding launch {w.partial_epoch[p]}")
                yield True
            for p in range(w.world):
                w.result_epoch[p][rank * chunks_per_rank + c] = launch
                yield True
        if exit_barrier:
            yield from barrier(2 * e + 2)
        for c, got in enumerate(w.result_epoch[rank]):   # write-back of the full result into the caller's slab
            if got != launch:
                w.violations.append(f"launch {launch}: rank {rank} copied result piece {c} holding launch {got}")
        yield True


def check_flag_protocol(world: int = 4, launches: int = 3, first_epoch: int = 1, chunks_per_rank: int = 2, seed: int = 0, schedules: int = 50,
                        entry_barrier: bool = True, exit_barrier: bool = True, wrap_safe: bool = True, stall_rank: Optional[int] = None,
                        max_steps: int = 2_000_000) -> ProtocolReport:
    """Runs ``schedules`` random interleavings (``stall_rank`` is scheduled 20x less often: a slow GPU).  A run that cannot make
    progress is reported as a deadlock."""
    rng = random.Random(seed)
    total_steps = 0
    violations: List[str] = []
    for s in range(schedules):
        w = _World(world, chunks_per_rank)
        # flag pads persist across launches; earlier launches left epoch first_epoch-1's exit value everywhere
        if first_epoch > 0:
            w.flags[:, :world] = (2 * (first_epoch - 1) + 2) & 0xFFFFFFFF
        progs = [_rank_program(w, r, first_epoch, launches, chunks_per_rank, entry_barrier, exit_barrier, wrap_safe) for r in range(world)]
        alive = list(range(world))
        blocked_streak = 0
        while alive:
            weights = [0.05 if r == stall_rank else 1.0 for r in alive]
            r = rng.choices(alive, weights)[0]
            try:
                progressed = next(progs[r])
            except StopIteration:
                alive.remove(r)
                continue
            total_steps += 1
            blocked_streak = 0 if progressed else blocked_streak + 1
            if blocked_streak > 200 * world or total_steps > max_steps:
                w.violations.append(f"schedule {s}: no progress (deadlock) with ranks {alive} still running")
                break
        violations += [f"[schedule {s}] {v}" for v in w.violations[:5]]
        if len(violations) > 20:
            break
    return ProtocolReport(ok=not violations, launches=launches, steps=total_steps, violations=violations)
