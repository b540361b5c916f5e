"""Span table of the zero-run-skipping uploads (runtime/staging.py::SparsePack): the library's threaded host scan against the numpy
statement of it, and both against the definition (first / last non-zero byte of every image row)."""
import numpy as np
import pytest
import torch

from neuroimagedisttraining_b200.runtime.staging import SparsePack


def _volumes(n=7, shape=(5, 6, 31), seed=0):
    g = np.random.default_rng(seed)
    v = g.integers(0, 256, size=(n,) + shape, dtype=np.uint8)
    v[g.random((n,) + shape) < 0.3] = 0                       # interior zeros stay inside the span
    for i in range(n):                                        # background at both ends, some empty rows, some full rows
        lo, hi = g.integers(0, 10), g.integers(20, 31)
        v[i, :, :, :lo] = 0
        v[i, :, :, hi:] = 0
    v[1, 2] = 0
    v[3, :, 1, :] = g.integers(1, 256, size=(shape[0], shape[2]), dtype=np.uint8)
    v[5] = 0                                                  # a completely empty volume
    return torch.from_numpy(v)


def _expand(sp, n, shape):
    out = np.zeros((n, shape[0] * shape[1], shape[2]), np.uint8)
    tab, off, packed = sp.tab.numpy(), sp.off.numpy().view(np.uint32), sp.packed.numpy()
    for i in range(n):
        for r in range(out.shape[1]):
            s, ln = int(tab[i, r, 0]), int(tab[i, r, 1])
            o = int(sp.vol_off[i]) + int(off[i, r])
            out[i, r, s:s + ln] = packed[o:o + ln]
    return out.reshape((n,) + shape)


def test_native_scan_equals_the_numpy_statement_and_is_lossless():
    cache = _volumes()
    n, shape = cache.shape[0], tuple(cache.shape[1:])
    a = SparsePack(cache, native=False)
    try:
        b = SparsePack(cache, native=True)
    except RuntimeError:
        pytest.skip("library not built")
    assert torch.equal(a.tab, b.tab) and torch.equal(a.off, b.off) and torch.equal(a.packed, b.packed)
    assert np.array_equal(a.vol_off, b.vol_off) and a.max_vol_bytes == b.max_vol_bytes and a.ratio == b.ratio
    assert np.array_equal(_expand(b, n, shape), cache.numpy())
    # the definition: first / last non-zero byte per row
    v = cache.numpy().reshape(n, -1, shape[2])
    for i, r in [(0, 0), (1, 2 * shape[1]), (3, 1), (5, 4), (6, 17)]:
        nzc = np.nonzero(v[i, r])[0]
        want = (int(nzc[0]), int(nzc[-1] - nzc[0] + 1)) if len(nzc) else (0, 0)
        assert (int(b.tab[i, r, 0]), int(b.tab[i, r, 1])) == want
    assert int(b.vol_off[6]) - int(b.vol_off[5]) == 0         # the empty volume packs to nothing
    assert b.ratio < 1.0
