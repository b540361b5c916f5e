"""Light-weight batch iterators and the FedML 8-tuple builders.

``load_partition_data_abcd`` returns the same structure as the reference
(``ABCD/data_loader.py:157-212``): ``[train_num, test_num, train_global, test_global,
local_num_dict, train_local_dict, test_local_dict, class_num]`` whose local loaders yield
``(sample_index, label, site)`` batches -- the volumes themselves are fetched from a
:class:`~.volumes.VolumeStore` by index when the batch is staged for the device.
"""
from __future__ import annotations

import math
import random
from typing import Iterator, Optional, Sequence, Tuple

import numpy as np
import torch

from .partition import partition_labels, rescale_partition, site_partition
from .volumes import VolumeStore


class IndexLoader:
    """Iterates ``(idx, y, site)`` mini-batches over a fixed index set.  ``shuffle`` draws a new
    permutation per epoch from a private, seeded generator (so runs are reproducible and
    independent of global RNG state).  ``len()`` = number of batches; ``.dataset`` mimics the
    attribute the reference reads for sample counts."""

    def __init__(self, indices, labels, sites, batch_size: int, shuffle: bool, seed: int = 0, drop_last: bool = False):
        self.indices = torch.as_tensor(np.asarray(indices), dtype=torch.int64)
        self.labels = torch.as_tensor(np.asarray(labels), dtype=torch.float32)
        self.sites = torch.as_tensor(np.asarray(sites), dtype=torch.float32)
        self.batch_size = int(batch_size)
        self.shuffle = shuffle
        self.drop_last = drop_last
        self._gen = torch.Generator()
        self._gen.manual_seed(int(seed))
        self.dataset = self.indices  # len(loader.dataset) == sample count

    def __len__(self) -> int:
        n = self.indices.numel()
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    @property
    def num_samples(self) -> int:
        return self.indices.numel()

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor, torch.Tensor]]:
        n = self.indices.numel()
        order = torch.randperm(n, generator=self._gen) if self.shuffle else torch.arange(n)
        for b in range(len(self)):
            sel = order[b * self.batch_size:(b + 1) * self.batch_size]
            yield self.indices[sel], self.labels[sel], self.sites[sel]

    def state(self):
        return self._gen.get_state()

    def set_state(self, s):
        self._gen.set_state(s)


class BatchAugment:
    """The reference's training transform for CIFAR / Tiny-ImageNet -- ``RandomCrop(size, padding=4)`` + ``RandomHorizontalFlip``
    (``cifar10/data_loader.py:43-60``), plus the ``Cutout(16)`` it keeps commented out -- as one batched tensor op on the already
    normalised images.  The crop pads with zeros in normalised space (torchvision pads the raw image with black, i.e. ``-mean/std``
    after normalisation); the stream is the loader's own seeded generator, so runs are reproducible."""

    def __init__(self, padding: int = 4, flip: bool = True, cutout: int = 0):
        self.padding, self.flip, self.cutout = int(padding), bool(flip), int(cutout)

    def __call__(self, x: torch.Tensor, gen: torch.Generator) -> torch.Tensor:
        n, _, h, w = x.shape
        if self.padding:
            p = self.padding
            xp = torch.nn.functional.pad(x, (p, p, p, p))
            dy = torch.randint(0, 2 * p + 1, (n,), generator=gen)
            dx = torch.randint(0, 2 * p + 1, (n,), generator=gen)
            rows = (dy.view(n, 1) + torch.arange(h).view(1, h)).view(n, 1, h, 1).expand(n, x.shape[1], h, w + 2 * p)
            xp = xp.gather(2, rows)
            cols = (dx.view(n, 1) + torch.arange(w).view(1, w)).view(n, 1, 1, w).expand(n, x.shape[1], h, w)
            x = xp.gather(3, cols)
        if self.flip:
            f = torch.rand(n, generator=gen) < 0.5
            x = torch.where(f.view(n, 1, 1, 1), x.flip(3), x)
        if self.cutout:
            c = self.cutout
            cy = torch.randint(0, h, (n,), generator=gen).view(n, 1, 1)
            cx = torch.randint(0, w, (n,), generator=gen).view(n, 1, 1)
            yy, xx = torch.arange(h).view(1, h, 1), torch.arange(w).view(1, 1, w)
            hole = ((yy >= cy - c // 2) & (yy < cy + c // 2) & (xx >= cx - c // 2) & (xx < cx + c // 2)).view(n, 1, h, w)
            x = x.masked_fill(hole, 0.0)
        return x


class TensorBatchLoader:
    """``(x, y)`` mini-batches over in-memory tensors (image datasets); ``augment`` is applied to training batches."""

    def __init__(self, x: torch.Tensor, y: torch.Tensor, batch_size: int, shuffle: bool, seed: int = 0, augment: Optional[BatchAugment] = None):
        self.x, self.y = x, y
        self.batch_size = int(batch_size)
        self.shuffle = shuffle
        self.augment = augment
        self._gen = torch.Generator()
        self._gen.manual_seed(int(seed))
        self.dataset = y

    def __len__(self):
        return (self.y.numel() + self.batch_size - 1) // self.batch_size

    @property
    def num_samples(self):
        return self.y.numel()

    def __iter__(self):
        n = self.y.shape[0]
        order = torch.randperm(n, generator=self._gen) if self.shuffle else torch.arange(n)
        for b in range(len(self)):
            sel = order[b * self.batch_size:(b + 1) * self.batch_size]
            xb = self.x[sel]
            yield (self.augment(xb, self._gen) if self.augment is not None else xb), self.y[sel]

    def state(self):
        return self._gen.get_state()

    def set_state(self, s):
        self._gen.set_state(s)


def _abcd_tuple(store: VolumeStore, train: Sequence[np.ndarray], test: Sequence[np.ndarray], batch_size: int, seed: int):
    local_num, train_local, test_local = {}, {}, {}
    for cid, (tr, te) in enumerate(zip(train, test)):
        train_local[cid] = IndexLoader(tr, store.labels[tr], store.sites[tr], batch_size, shuffle=True, seed=seed * 100003 + cid)
        test_local[cid] = IndexLoader(te, store.labels[te], store.sites[te], batch_size, shuffle=False)
        local_num[cid] = len(tr)
    # class_num 2 mirrors the reference's return value (``data_loader.py:211-212``); the model has 1 logit
    return [None, None, None, None, local_num, train_local, test_local, 2]


def load_partition_data_abcd(store: VolumeStore, partition_method: str = "site", partition_alpha: float = 0.3,
                             client_number: Optional[int] = None, batch_size: int = 16, logger=None, seed: int = 0):
    """One client per site.  The reference hard-codes 21 clients (``data_loader.py:176``); here the
    count is data-driven and ``client_number`` (if given) keeps the first N sites, which is what the
    reference's default ``client_num_in_total=4`` effectively does."""
    train, test, ids = site_partition(store.sites)
    if client_number is not None:
        if client_number > len(train):
            raise ValueError(f"{client_number} clients requested but the data has {len(train)} sites; "
                             "use load_partition_data_abcd_rescale")
        train, test = train[:client_number], test[:client_number]
    if logger is not None:
        logger.info("*********partition data based on site***************")
        for cid, (tr, te) in enumerate(zip(train, test)):
            logger.info("client_idx = %d, local_train_sample_number = %d, local_test_sample_number = %d" % (cid, len(tr), len(te)))
    return _abcd_tuple(store, train, test, batch_size, seed)


def load_partition_data_abcd_rescale(store: VolumeStore, partition_method: str = "site", partition_alpha: float = 0.3,
                                     client_number: int = 84, batch_size: int = 16, logger=None, seed: int = 0):
    train, test, _ = site_partition(store.sites)
    tr, te = rescale_partition(train, test, client_number)
    if logger is not None:
        for cid in range(client_number):
            logger.info("client_idx = %d, local_train_sample_number = %d, local_test_sample_number = %d" % (cid, len(tr[cid]), len(te[cid])))
    return _abcd_tuple(store, tr, te, batch_size, seed)


# ------------------------------------------------------------------------- image datasets
def synthetic_image_dataset(n_train: int, n_test: int, n_cls: int, shape=(3, 32, 32), seed: int = 0):
    """Class-conditional Gaussian blobs standing in for CIFAR/Tiny-ImageNet when the files are absent."""
    g = torch.Generator().manual_seed(seed)
    protos = torch.randn((n_cls,) + tuple(shape), generator=g) * 0.5
    def make(n):
        y = torch.randint(0, n_cls, (n,), generator=g)
        x = protos[y] + torch.randn((n,) + tuple(shape), generator=g) * 0.7
        return x, y
    return make(n_train), make(n_test)


_CIFAR_STATS = {"cifar10": ((0.49139968, 0.48215827, 0.44653124), (0.24703233, 0.24348505, 0.26158768)),
                "cifar100": ((0.5071, 0.4865, 0.4409), (0.2673, 0.2564, 0.2762))}


def _load_cifar(dataset: str, data_dir: str):
    """The python-pickle distribution of CIFAR-10 / CIFAR-100 (``cifar-10-batches-py/{data_batch_1..5,test_batch}``,
    ``cifar-100-python/{train,test}``; what torchvision downloads for the reference, ``cifar10/datasets.py``), read directly:
    uint8 ``[N, 3072]`` rows -> float ``[N, 3, 32, 32]`` normalised with the reference's channel statistics
    (``cifar10/data_loader.py:44-45``).  ``data_dir`` may be the dataset folder itself or its parent."""
    import os
    import pickle
    sub = "cifar-10-batches-py" if dataset == "cifar10" else "cifar-100-python"
    root = os.path.join(data_dir, sub) if os.path.isdir(os.path.join(data_dir, sub)) else data_dir
    train_files = [f"data_batch_{i}" for i in range(1, 6)] if dataset == "cifar10" else ["train"]
    test_files = ["test_batch"] if dataset == "cifar10" else ["test"]
    label_key = b"labels" if dataset == "cifar10" else b"fine_labels"
    mean, std = (torch.tensor(v).view(1, 3, 1, 1) for v in _CIFAR_STATS[dataset])

    def read(files):
        xs, ys = [], []
        for name in files:
            path = os.path.join(root, name)
            if not os.path.exists(path):
                raise FileNotFoundError(f"{dataset}: {path} not found (expected the python-pickle layout {sub}/{name})")
            with open(path, "rb") as f:
                d = pickle.load(f, encoding="bytes")
            xs.append(np.asarray(d[b"data"], dtype=np.uint8).reshape(-1, 3, 32, 32))
            ys.append(np.asarray(d[label_key], dtype=np.int64))
        x = torch.from_numpy(np.concatenate(xs)).float() / 255.0
        return (x - mean) / std, torch.from_numpy(np.concatenate(ys))
    return read(train_files), read(test_files)


def load_partition_data_images(dataset: str, data_dir: Optional[str], partition_method: str, partition_alpha: float,
                               client_number: int, batch_size: int, logger=None, seed: int = 0,
                               val_split: float = 0.0, synthetic_size: Tuple[int, int] = (2000, 500), augment: bool = False,
                               cutout: int = 0):
    """CIFAR-10/100/Tiny-ImageNet 8-tuple (``cifar10/data_loader.py:208-249``; the ``_val`` variant
    with a 10 % validation split, ``cifar10/data_val_loader.py:260-326``, via ``val_split``).
    Test data of a client follows the client's own label distribution, as in the reference.  ``augment`` turns the reference's
    training transform on (random crop with 4-pixel padding + horizontal flip, optional Cutout) for the clients' training loaders."""
    n_cls = {"cifar10": 10, "cifar100": 100, "tiny": 200}[dataset]
    shape = (3, 64, 64) if dataset == "tiny" else (3, 32, 32)
    if data_dir is None:                                 # no files: seeded synthetic stand-in of the same shape
        (xtr, ytr), (xte, yte) = synthetic_image_dataset(synthetic_size[0], synthetic_size[1], n_cls, shape, seed)
    elif dataset == "tiny":                              # a given directory must load -- never a silent switch to synthetic data
        from .tiny_imagenet import TinyImageNet
        (xtr, ytr), (xte, yte) = TinyImageNet(data_dir, True).tensors(), TinyImageNet(data_dir, False).tensors()
    else:
        (xtr, ytr), (xte, yte) = _load_cifar(dataset, data_dir)
    ytr_np, yte_np = ytr.numpy(), yte.numpy()
    rng = np.random.RandomState(seed)                 # one stream for the partition and the test subsets, like the reference's global one
    py_rng = random.Random(seed)
    idx_map = partition_labels(ytr_np, partition_method, client_number, partition_alpha, seed, rng=rng, py_rng=py_rng, n_cls=n_cls)
    stats = {c: {int(k): int(v) for k, v in zip(*np.unique(ytr_np[i], return_counts=True))} for c, i in idx_map.items()}
    local_num, train_local, test_local, val_local = {}, {}, {}, {}
    te_pools = {c: np.where(yte_np == c)[0] for c in range(n_cls)}
    n_te = math.ceil(len(yte_np) / client_number)
    aug = BatchAugment(4, True, cutout) if augment else None
    val_idx = {}
    if val_split > 0:
        # data_val_loader.py:271-274: a python-random sample of int(0.1 * len(client 0's data)) indices per client becomes its
        # validation set, the rest (as a set, so repeats collapse) stays for training
        for cid in range(client_number):
            nv = int(val_split * len(idx_map[0]))          # evaluated per client as in the reference: client 0 has shrunk by then
            pool = [int(i) for i in idx_map[cid]]
            val_idx[cid] = np.asarray(py_rng.sample(pool, min(nv, len(pool))), dtype=np.int64)
            idx_map[cid] = np.asarray(sorted(set(pool) - set(val_idx[cid].tolist())), dtype=np.int64)
    for cid in range(client_number):
        idx = idx_map[cid]
        if val_split > 0:
            val_local[cid] = TensorBatchLoader(xtr[val_idx[cid]], ytr[val_idx[cid]], batch_size, False)
        # test subset with the client's own label mix (cifar10/data_loader.py:222-236): per label, the first
        # ceil(share * ceil(n_test / clients)) entries of a fresh permutation of that label's test samples
        total = sum(stats[cid].values())
        picks = []
        for c in range(n_cls):
            want = math.ceil(stats[cid].get(c, 0) / total * n_te) if total else 0
            perm = rng.permutation(len(te_pools[c]))
            picks.append(te_pools[c][perm[:want]])
        te_idx = np.concatenate(picks).astype(np.int64)
        if len(te_idx) == 0:                              # a tiny test set without any of the client's classes
            te_idx = np.arange(min(n_te, len(yte_np)))
        train_local[cid] = TensorBatchLoader(xtr[idx], ytr[idx], batch_size, True, seed=seed * 100003 + cid, augment=aug)
        test_local[cid] = TensorBatchLoader(xte[te_idx], yte[te_idx], batch_size, False)
        local_num[cid] = len(idx)
    train_global = TensorBatchLoader(xtr, ytr, batch_size, True, seed)
    test_global = TensorBatchLoader(xte, yte, batch_size, False)
    out = [len(ytr), len(yte), train_global, test_global, local_num, train_local, test_local, stats]
    if val_split > 0:
        out.append(val_local)
    return out
