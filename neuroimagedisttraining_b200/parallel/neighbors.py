"""Neighbour averaging over peer memory.

Decentralised algorithms (D-PSGD, DisPFL, FedFomo) mix a client's model with a few *other clients'* models every round
(``fedml_api/standalone/dpsgd/dpsgd_api.py:169-178``).  With clients spread over ranks, replicating every model on every rank
and broadcasting all of them once per round moves O(clients x model) bytes through NCCL.  Here every rank keeps ONLY its own
clients' models, in a symmetric-memory arena that all peers have mapped, and the mixing kernel
(``nidt_weighted_aggregate``: multiply and add in list order) reads exactly the neighbours it needs straight out of their
owners' memory with P2P loads over NVLink -- a gather fused with the weighted sum, no staging copy, no collective call.

Two generations of the arena alternate (round parity), so a rank that is already publishing round r+1 never overwrites what a
slower rank still reads for round r; one stream-ordered one-element all-reduce per round lines the ranks up.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Sequence

import torch
import torch.distributed as dist

from ..ops import native
from .context import DistContext


class PeerModelArena:
    def __init__(self, ctx: DistContext, numel: int, n_clients: int, device):
        import torch.distributed._symmetric_memory as symm
        self.ctx, self.device, self.numel = ctx, torch.device(device), int(numel)
        self.pad = (self.numel + 1023) // 1024 * 1024
        owned: Dict[int, list] = {r: [] for r in range(ctx.world_size)}
        for c in range(n_clients):
            owned[ctx.owner(c)].append(c)
        self.slot = {c: i for r in owned for i, c in enumerate(owned[r])}
        self.slots = max(1, max(len(v) for v in owned.values()))
        group = dist.group.WORLD
        self.buf = symm.empty(2 * self.slots * self.pad, dtype=torch.float32, device=self.device)
        self.buf.zero_()
        self.hdl = symm.rendezvous(self.buf, group=group)
        self.base = [int(p) for p in self.hdl.buffer_ptrs]
        self._align = torch.zeros(1, dtype=torch.float32, device=self.device)
        torch.cuda.synchronize(self.device)
        dist.barrier(device_ids=[ctx.local_rank])
        self.lib = native.load()
        self.bytes_pulled = 0

    def _offset(self, cid: int, parity: int) -> int:
        return ((parity & 1) * self.slots + self.slot[int(cid)]) * self.pad

    def local_view(self, cid: int, parity: int) -> torch.Tensor:
        assert self.ctx.owns(int(cid))
        o = self._offset(cid, parity)
        return self.buf[o:o + self.numel]

    def publish(self, cid: int, flat: torch.Tensor, parity: int) -> None:
        """The owner makes ``flat`` the generation-``parity`` model of client ``cid``."""
        self.local_view(cid, parity).copy_(flat[:self.numel])

    def align(self) -> None:
        """Every rank's publishes of this generation are complete and visible before anyone reads them (stream-ordered)."""
        dist.all_reduce(self._align)

    def average(self, out: torch.Tensor, cids: Sequence[int], weights: Sequence[float], parity: int) -> torch.Tensor:
        """``out = sum_j weights[j] * model(cids[j])`` in list order; models are read in place from their owners (P2P loads)."""
        k = len(cids)
        ptrs = (C.c_void_p * k)(*[self.base[self.ctx.owner(int(c))] + 4 * self._offset(c, parity) for c in cids])
        wts = (C.c_float * k)(*[float(w) for w in weights])
        native.check(self.lib.nidt_weighted_aggregate(native.ptr(out), ptrs, wts, k, None, 0, self.numel, native.stream_ptr(self.device)),
                     "nidt_weighted_aggregate(peer)")
        native.count_launches(1 + (k - 1) // 64)
        self.bytes_pulled += 4 * self.numel * sum(1 for c in cids if not self.ctx.owns(int(c)))
        return out


def make_arena(ctx: DistContext, numel: int, n_clients: int, device):
    """The arena when this is a multi-rank CUDA run with symmetric memory on every rank, else None (callers then replicate and
    broadcast)."""
    if not ctx.is_distributed or torch.device(device).type != "cuda":
        return None
    arena, ok = None, 1.0
    try:
        arena = PeerModelArena(ctx, numel, n_clients, device)
    except Exception:                                      # noqa: BLE001
        ok = 0.0
    t = torch.tensor([ok], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return arena if t.item() >= 1.0 else None
