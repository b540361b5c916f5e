"""Client partitioners.

* ``site_partition``        one client per acquisition site, per-site 80/20 split after a
                            ``np.random.seed(42)`` shuffle (``ABCD/data_loader.py:67-102``).
* ``rescale_partition``     the site-ordered concatenation cut into N equal contiguous shards
                            (``load_partition_data_abcd_rescale``, ``ABCD/data_loader.py:216-315``).
* label-skew partitioners for the image datasets: ``dir``, ``n_cls``, ``my_part``
  (``cifar10/data_loader.py:75-194``), ``homo``, ``hetero`` (``cifar10/data_val_loader.py:83-237``)
  and FedML's LDA with min-size retry (``fedml_core/non_iid_partition/noniid_partition.py:6-102``).
"""
from __future__ import annotations

import random
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np


# ------------------------------------------------------------------------------ ABCD / site
def site_partition(sites: np.ndarray, split_ratio: float = 0.2, seed: int = 42) -> Tuple[List[np.ndarray], List[np.ndarray], List[int]]:
    """Per site: shuffle with ``seed`` (re-seeded for every site, as the reference does), last
    ``int(n*split_ratio)`` samples are the test split."""
    train, test, ids = [], [], []
    for s in np.unique(sites):
        idx = np.where(sites == s)[0]
        n_test = int(len(idx) * split_ratio)
        n_train = len(idx) - n_test
        rs = np.random.RandomState(seed)   # == np.random.seed(seed); np.random.shuffle(idx)
        rs.shuffle(idx)
        train.append(idx[:n_train].copy())
        test.append(idx[n_train:].copy())
        ids.append(int(s))
    return train, test, ids


def rescale_partition(train: Sequence[np.ndarray], test: Sequence[np.ndarray], client_number: int) -> Tuple[List[np.ndarray], List[np.ndarray]]:
    """Concatenate the per-site splits in site order and cut them into ``client_number`` equal
    contiguous shards (remainder samples are dropped, as in the reference)."""
    tr = np.concatenate(list(train))
    te = np.concatenate(list(test))
    ntr, nte = len(tr) // client_number, len(te) // client_number
    if ntr < 1:
        raise ValueError(f"{len(tr)} training samples cannot feed {client_number} clients")
    return ([tr[i * ntr:(i + 1) * ntr] for i in range(client_number)],
            [te[i * nte:(i + 1) * nte] for i in range(client_number)])


# ------------------------------------------------------------------------- label-skew family
def record_net_data_stats(y: np.ndarray, net_dataidx_map: Dict[int, np.ndarray]) -> Dict[int, Dict[int, int]]:
    out = {}
    for cid, idx in net_dataidx_map.items():
        u, c = np.unique(y[idx], return_counts=True)
        out[cid] = {int(a): int(b) for a, b in zip(u, c)}
    return out


def homo_partition(n: int, n_clients: int, rng: np.random.RandomState) -> Dict[int, np.ndarray]:
    idxs = rng.permutation(n)
    return {i: part for i, part in enumerate(np.array_split(idxs, n_clients))}


def hetero_dirichlet_partition(y: np.ndarray, n_clients: int, alpha: float, rng: np.random.RandomState,
                               min_size_required: int = 10) -> Dict[int, np.ndarray]:
    """Per-class Dirichlet over clients with the "balance" cap and min-size retry (FedML LDA)."""
    n = len(y)
    classes = np.unique(y)
    min_size = 0
    while min_size < min(min_size_required, max(1, n // (n_clients * 2))):
        idx_batch: List[List[int]] = [[] for _ in range(n_clients)]
        for k in classes:
            idx_k = np.where(y == k)[0]
            rng.shuffle(idx_k)
            p = rng.dirichlet(np.repeat(alpha, n_clients))
            p = np.array([pj * (len(b) < n / n_clients) for pj, b in zip(p, idx_batch)])
            p = p / p.sum()
            cuts = (np.cumsum(p) * len(idx_k)).astype(int)[:-1]
            idx_batch = [b + part.tolist() for b, part in zip(idx_batch, np.split(idx_k, cuts))]
        min_size = min(len(b) for b in idx_batch)
    out = {}
    for j in range(n_clients):
        rng.shuffle(idx_batch[j])
        out[j] = np.asarray(idx_batch[j], dtype=np.int64)
    return out


def non_iid_partition_with_dirichlet_distribution(label_list, client_num, classes, alpha, seed: int = 0):
    """FedML entry-point name kept for drop-in use (``noniid_partition.py:6-73``)."""
    return hetero_dirichlet_partition(np.asarray(label_list), client_num, alpha, np.random.RandomState(seed))


_MAX_CLASS_REDRAWS = 200_000_000     # per call, over all samples: below it the stream is the reference's, draw for draw


def _lognormal_sizes(n: int, n_clients: int, rng: np.random.RandomState, sigma: float = 0.0) -> np.ndarray:
    sizes = rng.lognormal(mean=np.log(n / n_clients), sigma=sigma, size=n_clients)
    return (sizes / sizes.sum() * n).astype(int)


def _redraw_until_available(rng, cum_j, prior_j, remaining, redraws) -> Tuple[int, int]:
    """-> (class, redraws so far); class -1 when nothing is left anywhere."""
    if not (remaining > 0).any():
        return -1, redraws
    block = 256
    while redraws < _MAX_CLASS_REDRAWS:
        state = rng.get_state()
        u = rng.uniform(size=block)
        cls = np.searchsorted(cum_j, u, side="left")          # == argmax(u <= cum): first class whose cumulative prior reaches u
        cls[cls >= len(cum_j)] = 0                            # argmax of an all-False row
        ok = remaining[cls] > 0
        if ok.any():
            i = int(np.argmax(ok))
            rng.set_state(state)
            rng.uniform(size=i + 1)
            return int(cls[i]), redraws + i
        redraws += block
        block = min(block * 4, 1 << 16)
    # the prior has (next to) no mass on what is left: fall back to the prior restricted to the available classes
    left = np.where(remaining > 0)[0]
    pr = prior_j[left]
    return (int(left[rng.choice(len(left), p=pr / pr.sum())]) if pr.sum() > 0 else int(left[rng.randint(len(left))])), redraws


def _draw_by_prior(y: np.ndarray, priors: np.ndarray, sizes: np.ndarray, rng, refill: str) -> Dict[int, np.ndarray]:
    """The sampling loop shared by ``dir`` / ``n_cls`` / ``my_part`` (``cifar10/data_loader.py:98-114,134-149,177-194``), draw for
    draw: pick a random client that still has quota, draw a class from its prior, take the next unused sample of that class.
    When the drawn class has run dry the reference redraws, after -- depending on the method -- leaving the pool empty (``dir``:
    ``refill="none"``, partitions are disjoint), resetting it to a random fill level (``n_cls``: ``"random"``) or to full
    (``my_part``: ``"full"``); with a refill, samples are handed out more than once, as in the reference."""
    n_clients, n_cls = priors.shape
    cum = np.cumsum(priors, axis=1)
    pools = [np.where(y == c)[0] for c in range(n_cls)]
    remaining = [len(p) for p in pools]
    out: Dict[int, List[int]] = {j: [] for j in range(n_clients)}
    sizes = sizes.copy()
    redraws = 0
    while sizes.sum() != 0:
        j = rng.randint(n_clients)
        if sizes[j] <= 0:
            continue
        sizes[j] -= 1
        while True:
            c = int(np.argmax(rng.uniform() <= cum[j]))
            if remaining[c] <= 0:
                redraws += 1
                if refill == "random" and len(pools[c]):
                    remaining[c] = int(rng.randint(0, len(pools[c])))
                elif refill == "full" and len(pools[c]):
                    remaining[c] = len(pools[c])
                if refill != "none":
                    # (a class without any sample can never be refilled: give up on this draw when the client's prior
                    # only covers such classes -- the reference would raise / spin there)
                    if len(pools[c]) == 0 and not any(priors[j][k] > 0 and len(pools[k]) for k in range(n_cls)):
                        break
                    if redraws < _MAX_CLASS_REDRAWS:
                        continue
                    break
                # ``dir``: the reference redraws one uniform at a time until a class with samples left comes up, which takes
                # millions of draws once a client's prior sits on exhausted classes.  Same draws, taken in blocks: on a hit the
                # generator is rewound to the start of the block and advanced by exactly the number of draws the loop would have used.
                c, redraws = _redraw_until_available(rng, cum[j], priors[j], np.asarray(remaining), redraws)
                if c < 0:
                    break
            remaining[c] -= 1
            out[j].append(int(pools[c][remaining[c]]))
            break
    return {j: np.asarray(v, dtype=np.int64) for j, v in out.items()}


def dir_partition(y: np.ndarray, n_clients: int, alpha: float, rng: np.random.RandomState, n_cls: int = None) -> Dict[int, np.ndarray]:
    """``dir``: each client draws a class prior ~ Dir(alpha) and fills its (lognormal, sigma=0) quota by sampling classes from that
    prior; exhausted classes are redrawn (``data_loader.py:118-149``)."""
    n_cls = n_cls or int(y.max()) + 1
    sizes = _lognormal_sizes(len(y), n_clients, rng)
    priors = rng.dirichlet([alpha] * n_cls, size=n_clients)
    return _draw_by_prior(y, priors, sizes, rng, "none")


def n_cls_partition(y: np.ndarray, n_clients: int, n_cls_per_client: int, rng: np.random.RandomState, n_cls: int = None,
                    py_rng: Optional[random.Random] = None) -> Dict[int, np.ndarray]:
    """``n_cls``: every client holds ``alpha`` (here ``n_cls_per_client``) classes with equal prior; the classes come from python's
    ``random.sample`` as in the reference (``data_loader.py:80-114``), hence the second stream ``py_rng``."""
    n_cls = n_cls or int(y.max()) + 1
    py_rng = py_rng or random
    sizes = _lognormal_sizes(len(y), n_clients, rng)
    priors = np.zeros((n_clients, n_cls))
    for j in range(n_clients):
        priors[j][py_rng.sample(range(n_cls), int(n_cls_per_client))] = 1.0 / n_cls_per_client
    return _draw_by_prior(y, priors, sizes, rng, "random")


def my_part_partition(y: np.ndarray, n_clients: int, n_shards: int, rng: np.random.RandomState, n_cls: int = None,
                      shard_alpha: float = 0.3) -> Dict[int, np.ndarray]:
    """``my_part``: clients are grouped into ``n_shards`` shards; all clients of a shard share one Dir(0.3) class prior
    (``data_loader.py:151-194``; it draws ``n_shards * n_clients`` priors and uses the first ``n_shards`` of them)."""
    n_cls = n_cls or int(y.max()) + 1
    sizes = _lognormal_sizes(len(y), n_clients, rng)
    tmp = rng.dirichlet([shard_alpha] * n_cls, size=int(n_shards * n_clients))
    per = max(1, int(n_clients / n_shards))
    priors = np.stack([tmp[int(j / per)] for j in range(n_clients)])
    return _draw_by_prior(y, priors, sizes, rng, "full")


def partition_labels(y: np.ndarray, method: str, n_clients: int, alpha: float, seed: int = 0,
                     rng: Optional[np.random.RandomState] = None, py_rng: Optional[random.Random] = None, n_cls: Optional[int] = None) -> Dict[int, np.ndarray]:
    """``rng``: continue an existing numpy stream (the loaders go on to draw the clients' test subsets from it, as the reference
    does from the global stream); default: a fresh ``RandomState(seed)`` == ``np.random.seed(seed)``.  ``py_rng``: likewise for python's
    ``random`` (``n_cls`` picks classes with it, the validation split samples with it).  ``n_cls``: number of classes of the dataset
    (the reference hard-codes 10 / 100 / 200); default: largest label + 1."""
    rng = rng if rng is not None else np.random.RandomState(seed)
    if method == "homo":
        return homo_partition(len(y), n_clients, rng)
    if method == "hetero":
        return hetero_dirichlet_partition(y, n_clients, alpha, rng)
    if method == "dir":
        return dir_partition(y, n_clients, alpha, rng, n_cls)
    if method == "n_cls":
        return n_cls_partition(y, n_clients, int(alpha), rng, n_cls, py_rng=py_rng or random.Random(seed))
    if method == "my_part":
        return my_part_partition(y, n_clients, int(alpha), rng, n_cls)
    raise ValueError(f"unknown partition method {method!r}")
