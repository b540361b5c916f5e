"""Property tests (hypothesis) of the host-side semantics every device kernel restates: partitioners, the round-seeded client
sampler, global top-k with kept ties, aggregation independent of client->rank packing, slab <-> state_dict round trips, the
DisPFL select-k and the secure-aggregation field arithmetic.  SURVEY §4 item 2."""
import math
import types

import numpy as np
import torch
from hypothesis import given, settings, strategies as st

from neuroimagedisttraining_b200 import ops
from neuroimagedisttraining_b200.data.partition import partition_labels, rescale_partition, site_partition
from neuroimagedisttraining_b200.fl.base import FedAPIBase
from neuroimagedisttraining_b200.models import Tiny3DCNN
from neuroimagedisttraining_b200.runtime.slab import ParamSlab
from neuroimagedisttraining_b200.utils import mpc
from neuroimagedisttraining_b200.utils.topology import AsymmetricTopologyManager, SymmetricTopologyManager

from helpers import quiet_logger

FAST = settings(max_examples=25, deadline=None, derandomize=True, database=None)      # same examples on every run


# ------------------------------------------------------------------------------ partitioners
@FAST
@given(st.lists(st.integers(5, 40), min_size=1, max_size=8), st.floats(0.05, 0.5))
def test_site_partition_is_a_disjoint_cover_with_the_stated_split(sizes, ratio):
    sites = np.repeat(np.arange(len(sizes)), sizes)
    np.random.RandomState(0).shuffle(sites)
    train, test, ids = site_partition(sites, split_ratio=ratio)
    assert ids == sorted(set(sites.tolist()))
    seen = np.concatenate(train + test)
    assert len(seen) == len(sites) == len(set(seen.tolist()))
    for s, tr, te in zip(ids, train, test):
        assert (sites[tr] == s).all() and (sites[te] == s).all()
        assert len(te) == int((sites == s).sum() * ratio)
    again = site_partition(sites, split_ratio=ratio)
    assert all(np.array_equal(a, b) for a, b in zip(train + test, again[0] + again[1]))       # seed 42 every time


@FAST
@given(st.lists(st.integers(6, 30), min_size=2, max_size=6), st.integers(1, 12))
def test_rescale_partition_cuts_equal_contiguous_shards(sizes, n_clients):
    sites = np.repeat(np.arange(len(sizes)), sizes)
    train, test, _ = site_partition(sites)
    tr, te = rescale_partition(train, test, n_clients)
    cat = np.concatenate(train)
    assert len(tr) == len(te) == n_clients
    assert len({len(t) for t in tr}) == 1 and len(tr[0]) == len(cat) // n_clients
    assert np.array_equal(np.concatenate(tr), cat[:len(tr[0]) * n_clients])                    # site order kept, remainder dropped


@FAST
@given(st.sampled_from(["homo", "hetero", "dir", "n_cls", "my_part"]), st.integers(2, 8), st.integers(0, 1000))
def test_label_partitioners_are_deterministic_and_in_range(method, n_clients, seed):
    y = np.random.RandomState(seed).randint(0, 10, size=600)
    alpha = {"homo": 0, "hetero": 0.5, "dir": 0.3, "n_cls": 3, "my_part": 2}[method]
    a = partition_labels(y, method, n_clients, alpha, seed=seed)
    b = partition_labels(y, method, n_clients, alpha, seed=seed)
    assert sorted(a) == list(range(n_clients))
    allidx = np.concatenate([a[c] for c in range(n_clients)])
    assert allidx.min() >= 0 and allidx.max() < len(y)
    if method not in ("n_cls", "my_part"):               # those two refill exhausted class pools (the reference's behaviour)
        assert len(allidx) == len(set(allidx.tolist()))
    assert all(np.array_equal(a[c], b[c]) for c in range(n_clients))


# ---------------------------------------------------------------------------------- sampler
@FAST
@given(st.integers(2, 40), st.integers(0, 500), st.data())
def test_client_sampling_depends_only_on_the_round(total, round_idx, data):
    per_round = data.draw(st.integers(1, total))
    api = types.SimpleNamespace(logger=quiet_logger())
    np.random.seed(123)
    first = np.asarray(FedAPIBase._client_sampling(api, round_idx, total, per_round))
    np.random.seed(999)                                   # whatever the stream did in between
    np.random.rand(17)
    second = np.asarray(FedAPIBase._client_sampling(api, round_idx, total, per_round))
    assert np.array_equal(first, second)
    assert len(first) == per_round == len(set(first.tolist())) and first.min() >= 0 and first.max() < total
    if per_round == total:
        assert first.tolist() == list(range(total))
    else:                                                 # the reference's rule, restated
        np.random.seed(round_idx)
        assert np.array_equal(first, np.random.choice(range(total), per_round, replace=False))


# -------------------------------------------------------------------------- top-k with ties
@FAST
@given(st.lists(st.integers(0, 6), min_size=4, max_size=200), st.floats(0.05, 0.95))
def test_global_topk_keeps_every_tie_at_the_threshold(vals, keep):
    """Scores drawn from a handful of integers force ties: the kept set is everything >= the k-th largest (``snip.py:85-98``), so at
    least k entries survive, entries strictly above the threshold always survive and non-scored positions are ones."""
    s = torch.tensor(vals, dtype=torch.float32)
    if float(s.sum()) == 0:
        s[0] = 1.0
    pad = 3
    scores = torch.cat([s, torch.full((pad,), 123.0)])    # trailing non-prunable region
    mask = torch.empty_like(scores)
    ops.reference.global_topk_mask(scores, [(0, s.numel())], keep, mask)
    k = int(s.numel() * keep)
    kept = mask[:s.numel()].bool()
    assert (mask[s.numel():] == 1).all()
    if k == 0:
        assert not kept.any()
        return
    kth = torch.sort(s, descending=True)[0][k - 1]
    assert torch.equal(kept, s >= kth)
    assert int(kept.sum()) >= k


# --------------------------------------------------------- aggregation vs. packing / ordering
@FAST
@given(st.integers(2, 9), st.integers(1, 4), st.integers(0, 10_000))
def test_aggregation_is_independent_of_client_to_rank_packing(n_clients, n_ranks, seed):
    """Each rank pre-reduces its clients with the global weights n_i / sum n, ranks are then summed: any packing gives the plain
    weighted mean up to fp32 reassociation, and exactly the same bits when the packing keeps the client order on one rank."""
    g = torch.Generator().manual_seed(seed)
    xs = [torch.randn(257, generator=g) for _ in range(n_clients)]
    n = torch.randint(1, 50, (n_clients,), generator=g).tolist()
    w = [v / sum(n) for v in n]
    exact = ops.reference.weighted_aggregate(torch.empty(257), xs, w)
    oracle = sum(x.double() * wt for x, wt in zip(xs, w))
    torch.testing.assert_close(exact.double(), oracle, rtol=1e-6, atol=1e-6)
    parts = [list(range(r, n_clients, n_ranks)) for r in range(n_ranks)]
    total = torch.zeros(257)
    for p in parts:
        if p:
            total += ops.reference.weighted_aggregate(torch.empty(257), [xs[i] for i in p], [w[i] for i in p])
    torch.testing.assert_close(total, exact, rtol=1e-5, atol=1e-6)
    if n_ranks == 1:
        assert torch.equal(total, exact)
    m = (torch.rand(200, generator=g) > 0.5).float()      # the SalientGrads mask covers the parameter prefix only
    masked = ops.reference.weighted_aggregate(torch.empty(257), xs, w, mask=m)
    assert torch.equal(masked[:200], exact[:200] * m) and torch.equal(masked[200:], exact[200:])


# ---------------------------------------------------------------------------------- slabs
@FAST
@given(st.integers(0, 10_000))
def test_slab_state_dict_round_trip_and_aligned_views(seed):
    torch.manual_seed(seed)
    model = Tiny3DCNN(1, in_shape=(32, 32, 32), width=4)
    for b in model.buffers():
        if b.dtype == torch.long:
            b.fill_(seed % 7)
    ref = {k: v.clone() for k, v in model.state_dict().items()}
    slab = ParamSlab.from_module(model, bind=False)
    sd = slab.state_dict()
    assert list(sd) == list(ref)
    for k in ref:
        assert sd[k].dtype == ref[k].dtype and torch.equal(sd[k], ref[k]), k
    for e in slab.layout.entries.values():
        assert (e.offset * 4) % 16 == 0                   # every tensor starts on a 16-byte boundary
    other = slab.zeros_like()
    other.load_state_dict(sd)
    assert torch.equal(other.flat, slab.flat)
    covered = sum(e.numel for e in slab.layout.entries.values())
    assert int((slab.flat != 0).sum()) <= covered         # alignment gaps stay zero


# ------------------------------------------------------------------------ DisPFL select-k
@FAST
@given(st.integers(1, 300), st.floats(0.0, 1.0), st.integers(0, 10_000))
def test_select_k_preserves_counts_and_picks_extremes(n, frac, seed):
    g = torch.Generator().manual_seed(seed)
    v = torch.randn(n, generator=g)
    mask = (torch.rand(n, generator=g) > 0.5).float()
    active, inactive = int(mask.sum()), n - int(mask.sum())
    k = math.ceil(frac * active)
    fired = ops.mask_select_k(v, mask.clone(), k, "fire")
    assert int(fired.sum()) == active - k and (fired <= mask).all()
    if 0 < k < active:
        assert v.abs()[(mask - fired).bool()].max() <= v.abs()[fired.bool()].min()
    k2 = min(k, n - int(fired.sum()))
    grown = ops.mask_select_k(v, fired.clone(), k2, "regrow")
    assert int(grown.sum()) == active - k + k2 and (grown >= fired).all()
    new = (grown - fired).bool()
    rest = (grown == 0)
    if new.any() and rest.any():
        assert v.abs()[new].min() >= v.abs()[rest].max()
    assert inactive + active == n


# -------------------------------------------------------------------------------- MPC field
@FAST
@given(st.integers(1, 2**31 - 2))
def test_modular_inverse(a):
    p = 2**31 - 1
    assert (a * mpc.modular_inv(a, p)) % p == 1


@FAST
@given(st.integers(3, 7), st.integers(0, 10_000))
def test_bgw_and_quantisation_round_trips(n_workers, seed):
    p = 2**31 - 1
    rng = np.random.RandomState(seed)
    T = (n_workers - 1) // 2
    X = rng.randint(0, p, size=(4, 3)).astype(np.int64)
    shares = mpc.BGW_encoding(X, n_workers, T, p, rng)
    idx = list(range(T + 1))                              # any T+1 shares reconstruct
    assert np.array_equal(mpc.BGW_decoding(shares[idx], idx, p)[0] % p, X % p)
    x = rng.randn(50).astype(np.float64)
    q = mpc.quantize(x, 2**16, p)
    np.testing.assert_allclose(mpc.dequantize(q, 2**16, p), x, atol=1.0 / 2**16)
    ss = mpc.Gen_Additive_SS(5, n_workers, p, rng)
    assert (np.mod(ss.sum(axis=0), p) == 0).all()


# -------------------------------------------------------------------------------- topology
@FAST
@given(st.integers(4, 16), st.integers(0, 2))
def test_topology_mixing_matrices_are_row_stochastic(n, extra):
    k = 2 + 2 * extra
    if k >= n:
        k = 2
    sym = SymmetricTopologyManager(n, k)
    sym.generate_topology()
    W = np.asarray(sym.topology)
    np.testing.assert_allclose(W.sum(axis=1), 1.0, atol=1e-6)
    assert ((W > 0) == (W.T > 0)).all()
    assert all(i in sym.get_in_neighbor_idx_list(i) or W[i, i] > 0 for i in range(n))
    asym = AsymmetricTopologyManager(n, k, max(1, extra))
    asym.generate_topology()
    np.testing.assert_allclose(np.asarray(asym.topology).sum(axis=1), 1.0, atol=1e-6)
