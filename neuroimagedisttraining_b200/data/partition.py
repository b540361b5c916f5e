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

from typing import Dict, List, Sequence, Tuple

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


_MAX_CLASS_REDRAWS = 10_000


def _lognormal_sizes(n: int, n_clients: int, rng: np.random.RandomState, sigma: float = 0.0) -> np.ndarray:
    sizes = rng.lognormal(mean=np.log(n / n_clients), sigma=sigma, size=n_clients)
    return (sizes / sizes.sum() * n).astype(int)


def dir_partition(y: np.ndarray, n_clients: int, alpha: float, rng: np.random.RandomState, n_cls: int = None) -> Dict[int, np.ndarray]:
    """``dir``: each client draws a class prior ~ Dir(alpha) and fills its (lognormal, sigma=0)
    quota by sampling classes from that prior until class pools run dry (``data_loader.py:118-149``)."""
    n_cls = n_cls or int(y.max()) + 1
    sizes = _lognormal_sizes(len(y), n_clients, rng)
    priors = rng.dirichlet([alpha] * n_cls, size=n_clients)
    cum = np.cumsum(priors, axis=1)
    pools = [list(np.where(y == c)[0]) for c in range(n_cls)]
    remaining = np.array([len(p) for p in pools])
    out: Dict[int, List[int]] = {j: [] for j in range(n_clients)}
    sizes = sizes.copy()
    while sizes.sum() > 0 and remaining.sum() > 0:
        j = rng.randint(n_clients)
        if sizes[j] <= 0:
            continue
        sizes[j] -= 1
        redraws = 0
        while True:
            c = int(np.argmax(rng.uniform() <= cum[j]))
            if remaining[c] <= 0:
                if remaining.sum() <= 0:
                    break
                redraws += 1
                if redraws < _MAX_CLASS_REDRAWS:
                    continue
                # The reference redraws for ever here; when the client's prior has (next to) no mass on the classes that still
                # have samples that never ends.  Same stream up to this point, then the prior restricted to what is left.
                left = np.where(remaining > 0)[0]
                pr = priors[j][left]
                c = int(left[rng.choice(len(left), p=pr / pr.sum())]) if pr.sum() > 0 else int(left[rng.randint(len(left))])
            remaining[c] -= 1
            out[j].append(pools[c][remaining[c]])
            break
    return {j: np.asarray(v, dtype=np.int64) for j, v in out.items()}


def n_cls_partition(y: np.ndarray, n_clients: int, n_cls_per_client: int, rng: np.random.RandomState, n_cls: int = None) -> Dict[int, np.ndarray]:
    """``n_cls``: every client holds ``alpha`` (here ``n_cls_per_client``) classes with equal prior."""
    n_cls = n_cls or int(y.max()) + 1
    priors = np.zeros((n_clients, n_cls))
    for j in range(n_clients):
        priors[j, rng.choice(n_cls, int(n_cls_per_client), replace=False)] = 1.0 / n_cls_per_client
    return _fill_from_priors(y, priors, _lognormal_sizes(len(y), n_clients, rng), rng)


def my_part_partition(y: np.ndarray, n_clients: int, n_shards: int, rng: np.random.RandomState, n_cls: int = None,
                      shard_alpha: float = 0.3) -> Dict[int, np.ndarray]:
    """``my_part``: clients are grouped into ``n_shards`` shards; all clients of a shard share one
    Dir(0.3) class prior (``data_loader.py:151-194``)."""
    n_cls = n_cls or int(y.max()) + 1
    shard_priors = rng.dirichlet([shard_alpha] * n_cls, size=int(n_shards))
    per = max(1, n_clients // int(n_shards))
    priors = np.stack([shard_priors[min(j // per, int(n_shards) - 1)] for j in range(n_clients)])
    return _fill_from_priors(y, priors, _lognormal_sizes(len(y), n_clients, rng), rng)


def _fill_from_priors(y, priors, sizes, rng) -> Dict[int, np.ndarray]:
    n_clients, n_cls = priors.shape
    cum = np.cumsum(priors, axis=1)
    pools = [list(np.where(y == c)[0]) for c in range(n_cls)]
    remaining = np.array([len(p) for p in pools])
    out: Dict[int, List[int]] = {j: [] for j in range(n_clients)}
    sizes = sizes.copy()
    guard = 0
    while sizes.sum() > 0 and remaining.sum() > 0 and guard < 50 * len(y):
        guard += 1
        j = rng.randint(n_clients)
        if sizes[j] <= 0:
            continue
        avail = (priors[j] > 0) & (remaining > 0)
        if not avail.any():
            sizes[j] = 0
            continue
        c = int(np.argmax(rng.uniform() <= cum[j]))
        if remaining[c] <= 0:
            continue
        sizes[j] -= 1
        remaining[c] -= 1
        out[j].append(pools[c][remaining[c]])
    return {j: np.asarray(v, dtype=np.int64) for j, v in out.items()}


def partition_labels(y: np.ndarray, method: str, n_clients: int, alpha: float, seed: int = 0) -> Dict[int, np.ndarray]:
    rng = np.random.RandomState(seed)
    if method == "homo":
        return homo_partition(len(y), n_clients, rng)
    if method == "hetero":
        return hetero_dirichlet_partition(y, n_clients, alpha, rng)
    if method == "dir":
        return dir_partition(y, n_clients, alpha, rng)
    if method == "n_cls":
        return n_cls_partition(y, n_clients, int(alpha), rng)
    if method == "my_part":
        return my_part_partition(y, n_clients, int(alpha), rng)
    raise ValueError(f"unknown partition method {method!r}")
