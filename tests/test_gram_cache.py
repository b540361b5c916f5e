"""Bookkeeping of the per-sample Gram-matrix cache (ops/engine_b200.py::GramCache) with the kernels replaced by a recorder: which
samples are computed when, where they are stored, what the per-batch sum reads -- misses, repeats inside a batch, growth, the memory
bound, a change of the volume store.  (The kernels are checked on the GPU: tests/test_alg_bwd0_gpu.py.)"""
import torch

from neuroimagedisttraining_b200.ops.engine_b200 import GramCache


class _Lib:
    def __init__(self, max_list=8):
        self.max_list = max_list
        self.computed, self.summed = [], []

    def nidt_conv_in_gram_floats(self):
        return 16

    def nidt_conv_in_gram_max_samples(self):
        return self.max_list

    def nidt_conv_in_gram_scratch_floats(self, n, di):
        return n * 4

    def nidt_conv_in_gram_samples(self, x, rows, slots, n, cache, scratch, d, h, w, scale, stream):
        self.computed.append((list(rows)[:n], list(slots)[:n]))
        return 0

    def nidt_conv_in_gram_sum(self, cache, slots, n, out, stream):
        self.summed.append(list(slots)[:n])
        return 0


def _cache(monkeypatch, mb=None, max_list=8):
    if mb is not None:
        monkeypatch.setenv("NIDT_GRAM_CACHE_MB", str(mb))
    lib = _Lib(max_list)
    return GramCache(lib, torch.device("cpu"), (9, 9, 9), 1 / 255.0, 4), lib


def test_misses_hits_repeats_and_growth(monkeypatch):
    g, lib = _cache(monkeypatch)
    x = torch.zeros((4, 9, 9, 9), dtype=torch.uint8)
    assert g.sum_for([10, 11, 12, 13], x, 0)
    assert lib.computed == [([0, 1, 2, 3], [0, 1, 2, 3])] and lib.summed == [[0, 1, 2, 3]] and g.built == 4
    assert g.sum_for([13, 10, 12, 11], x, 0)                       # all cached: only the sum runs, slots follow the ids
    assert len(lib.computed) == 1 and lib.summed[-1] == [3, 0, 2, 1]
    assert g.sum_for([10, 99, 99, 11], x, 0)                       # one new sample that occurs twice: computed once (from its first row)
    assert lib.computed[-1] == ([1], [4]) and lib.summed[-1] == [0, 4, 4, 1] and g.built == 5
    assert not g.sum_for([1, 2, 3], x, 0)                          # ids do not match the batch
    assert not g.sum_for(list(range(9)), torch.zeros((9, 9, 9, 9), dtype=torch.uint8), 0)      # more samples than one launch takes
    # growth keeps what is cached
    many = list(range(1000, 1300))
    for a in range(0, 300, 4):
        assert g.sum_for(many[a:a + 4], x, 0)
    assert g.cap >= 305 and g.slots[10] == 0 and g.slots[1299] == 304
    assert g.sum_for([10, 11, 1299, 1000], x, 0) and lib.summed[-1] == [0, 1, 304, 5]


def test_memory_bound_and_store_change(monkeypatch):
    g, lib = _cache(monkeypatch, mb=0)                             # bound: the minimum of 64 matrices
    assert g.limit == 64
    x = torch.zeros((4, 9, 9, 9), dtype=torch.uint8)
    for a in range(0, 64, 4):
        assert g.sum_for(list(range(a, a + 4)), x, 0)
    assert len(g.slots) == 64 and g.built == 64
    assert g.sum_for([500, 501, 0, 1], x, 0)                       # full: starts over, everything in this batch is recomputed
    assert len(g.slots) == 4 and lib.computed[-1] == ([0, 1, 2, 3], [0, 1, 2, 3]) and g.built == 68
    # ids index a volume store: another store means other samples
    g.use_store("a")
    assert len(g.slots) == 0
    assert g.sum_for([0, 1, 2, 3], x, 0) and g.built == 72
    g.use_store("a")
    assert g.sum_for([3, 2, 1, 0], x, 0) and g.built == 72
    g.use_store("b")
    assert g.sum_for([3, 2, 1, 0], x, 0) and g.built == 76
