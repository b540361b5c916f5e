"""The data-path choice of the fused aggregation (parallel/symmetric.py::SymmetricAggregator.pick) as a table: which kernel variant
a slab of a given size takes at a given rank count, and what the environment / constructor overrides do.  (The variants themselves
are checked against ordered sums and NCCL on GPUs: tests/dist_worker_sym_agg.py, tests/test_sym_world1.py.)"""
import types

import pytest

from neuroimagedisttraining_b200.parallel.symmetric import SymmetricAggregator

MB = (1 << 20) // 4          # floats per MiB


def _agg(world, multicast=True, forced=None, bulk_max=8, mm_max=256 << 20):
    lib = types.SimpleNamespace(nidt_sym_bulk_max_ranks=lambda: bulk_max)
    return types.SimpleNamespace(ctx=types.SimpleNamespace(world_size=world), lib=lib, has_multicast=multicast, _mm_forced=forced,
                                 multimem_max_bytes=mm_max)


def _pick(a, numel):
    return SymmetricAggregator.pick(a, numel)


def test_choice_by_size_and_rank_count(monkeypatch):
    monkeypatch.delenv("NIDT_AGG_MODE", raising=False)
    for world in (2, 4):                                           # up to 4 ranks: shared-memory P2P at every size
        for mb in (1, 10, 256, 1024):
            assert _pick(_agg(world), mb * MB) == ("bulk", 0.0, 0)
    a = _agg(8)
    assert _pick(a, 10 * MB) == ("multimem", 0.0, 0)               # the FL model: NVLS
    assert _pick(a, 64 * MB) == ("multimem", 0.0, 148)             # from 64 MB on one CTA per SM
    assert _pick(a, 256 * MB) == ("multimem", 0.0, 148)
    assert _pick(a, 1024 * MB) == ("bulk", 0.5, 0)                 # beyond: P2P with half of every chunk through the switch
    assert _pick(_agg(8, multicast=False), 10 * MB) == ("bulk", 0.0, 0)
    assert _pick(_agg(16, multicast=False), 10 * MB) == ("p2p", 0.0, 0)     # more ranks than the bulk variant stages for


def test_overrides(monkeypatch):
    monkeypatch.delenv("NIDT_AGG_MODE", raising=False)
    assert _pick(_agg(2, forced=True), 10 * MB)[0] == "multimem"
    assert _pick(_agg(8, forced=False), 10 * MB) == ("bulk", 0.0, 0)
    monkeypatch.setenv("NIDT_AGG_MODE", "p2p")
    assert _pick(_agg(8), 10 * MB) == ("p2p", 0.0, 0)
    monkeypatch.setenv("NIDT_AGG_MODE", "multimem")
    assert _pick(_agg(8), 10 * MB) == ("multimem", 0.0, 0)
    assert _pick(_agg(8, multicast=False), 10 * MB) == ("bulk", 0.0, 0)      # asked for, not available: the choice by size
    monkeypatch.setenv("NIDT_AGG_MODE", "bulk")
    assert _pick(_agg(16), 10 * MB)[0] == "multimem"                          # bulk not available at 16 ranks: the choice by size
