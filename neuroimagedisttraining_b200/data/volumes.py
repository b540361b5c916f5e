"""Volume stores: where a batch of 3-D scans comes from.

The reference keeps only *indices* in its loaders and re-opens one HDF5 file per batch to fetch
``X[sorted idx]`` as fp32 CUDA tensors (``sailentgrads/my_model_trainer.py:185-199``,
``client.py:55-72``; 8.5 MB/sample over PCIe).  Here a store hands out 8-bit volumes from pinned
host memory (2.1 MB/sample) and normalisation happens on the device, in the first conv's loader.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import numpy as np
import torch

ABCD_SHAPE = (121, 145, 121)


class VolumeStore:
    """Abstract source of ``(uint8 volumes [B,D,H,W], float labels [B])`` by sample index."""
    shape: Tuple[int, int, int]
    labels: np.ndarray
    sites: np.ndarray
    scale: float = 1.0 / 255.0   # device-side normalisation x = u8 * scale

    def __len__(self) -> int:
        return len(self.labels)

    def fetch(self, indices, out: Optional[torch.Tensor] = None) -> Tuple[torch.Tensor, torch.Tensor]:
        raise NotImplementedError

    def fetch_sorted(self, indices, out=None):
        """The reference sorts the indices of a batch before reading (``my_model_trainer.py:190-195``)."""
        idx = np.sort(np.asarray(indices).astype(np.int64))
        return self.fetch(idx, out)

    def to_float(self, x_u8: torch.Tensor) -> torch.Tensor:
        return x_u8.to(torch.float32) * self.scale


class InMemoryVolumeStore(VolumeStore):
    def __init__(self, volumes_u8: torch.Tensor, labels, sites, pin: bool = False):
        assert volumes_u8.dtype == torch.uint8 and volumes_u8.dim() == 4
        self.volumes = volumes_u8.pin_memory() if (pin and torch.cuda.is_available()) else volumes_u8
        self.labels = np.asarray(labels, dtype=np.float32)
        self.sites = np.asarray(sites, dtype=np.int64)
        self.shape = tuple(volumes_u8.shape[1:])

    def fetch(self, indices, out=None):
        idx = torch.as_tensor(np.asarray(indices).astype(np.int64))
        if out is None:
            x = self.volumes.index_select(0, idx)
        else:
            x = out[:idx.numel()]
            torch.index_select(self.volumes, 0, idx, out=x)
        return x, torch.from_numpy(self.labels[idx.numpy()])


class SyntheticABCD(VolumeStore):
    """Seeded synthetic stand-in for the ABCD grey-matter maps (11 573 subjects x 121x145x121,
    8-bit quantised, binary label, 21+ acquisition sites -- ``Preprocess_ABCD.ipynb`` cells 6-32).

    Volumes come from a small pool of smooth "brain-like" base fields; subject ``i`` is
    ``pool[i % pool] `` modulated by a label-dependent regional contrast and a per-subject gain, so
    the classification task is learnable and every subject is distinct.  Site sizes are uneven.
    """

    def __init__(self, n_subjects: int = 420, shape=ABCD_SHAPE, n_sites: int = 21, seed: int = 0,
                 pool: int = 8, pin: bool = False, materialize: bool = True, site_sizes: Optional[Sequence[int]] = None,
                 label_noise: float = 0.0, signal_gain: float = 1.0):
        """``label_noise``: probability that a subject's stored label is the opposite of the one its volume was generated from
        (an irreducible error: no model can score above ``1 - label_noise`` on held-out subjects); ``signal_gain`` scales the
        label-dependent regional contrast.  The defaults give the cleanly separable task of the throughput benchmarks."""
        rng = np.random.RandomState(seed)
        self.shape = tuple(shape)
        self.n_sites = n_sites
        if site_sizes is None:
            w = rng.dirichlet(np.full(n_sites, 5.0))
            sizes = np.maximum((w * n_subjects).astype(int), 5)
            sizes[-1] = max(5, n_subjects - sizes[:-1].sum())
        else:
            sizes = np.asarray(site_sizes, dtype=int)
        self.sites = np.repeat(np.arange(n_sites), sizes).astype(np.int64)
        n = len(self.sites)
        self.labels = (rng.rand(n) < 0.5).astype(np.float32)
        self._gen_labels = self.labels.copy()              # what the volumes are generated from
        self._signal_gain = float(signal_gain)
        self._gain = (0.85 + 0.3 * rng.rand(n)).astype(np.float32)
        self._pool_id = rng.randint(0, pool, size=n)
        d, h, w_ = self.shape
        zz, yy, xx = np.meshgrid(np.linspace(-1, 1, d), np.linspace(-1, 1, h), np.linspace(-1, 1, w_), indexing="ij")
        r2 = (zz / 0.8) ** 2 + (yy / 0.9) ** 2 + (xx / 0.8) ** 2
        brain = np.clip(1.2 - r2, 0, 1).astype(np.float32)
        self._pool = []
        for p in range(pool):
            f = rng.rand(3) * 6 + 2
            ph = rng.rand(3) * 6.28
            tex = 0.5 + 0.25 * (np.sin(f[0] * zz + ph[0]) * np.sin(f[1] * yy + ph[1]) * np.sin(f[2] * xx + ph[2]))
            self._pool.append((brain * tex).astype(np.float32))
        # label signal: a regional contrast (one hemisphere-ish blob brighter for label 1)
        self._signal = (np.exp(-(((zz - 0.2) ** 2 + (yy + 0.1) ** 2 + (xx - 0.3) ** 2) / 0.08)) * brain).astype(np.float32)
        if label_noise > 0:
            flip = np.random.RandomState(seed + 7919).rand(n) < label_noise
            self.labels = np.where(flip, 1.0 - self.labels, self.labels).astype(np.float32)
        self._cache = None
        self._pin = pin
        if materialize:
            self.materialize()

    def _make(self, i: int) -> np.ndarray:
        v = self._pool[self._pool_id[i]] * self._gain[i] + self._signal_gain * (0.25 if self._gen_labels[i] > 0.5 else -0.05) * self._signal
        return np.clip(v * 255.0, 0, 255).astype(np.uint8)

    def materialize(self) -> None:
        n = len(self.labels)
        buf = torch.empty((n,) + self.shape, dtype=torch.uint8)
        if self._pin and torch.cuda.is_available():
            buf = buf.pin_memory()
        arr = buf.numpy()
        for i in range(n):
            arr[i] = self._make(i)
        self._cache = buf

    def fetch(self, indices, out=None):
        idx = np.asarray(indices).astype(np.int64)
        y = torch.from_numpy(self.labels[idx])
        if self._cache is not None:
            ti = torch.from_numpy(idx)
            if out is None:
                return self._cache.index_select(0, ti), y
            x = out[:len(idx)]
            torch.index_select(self._cache, 0, ti, out=x)
            return x, y
        x = out[:len(idx)] if out is not None else torch.empty((len(idx),) + self.shape, dtype=torch.uint8)
        xn = x.numpy()
        for j, i in enumerate(idx):
            xn[j] = self._make(int(i))
        return x, y


def hdf5_backend():
    """``h5py`` when a working one is installed, else the in-tree reader for plain contiguous files (what the reference's notebook
    writes).  (A module named h5py without ``File`` -- e.g. an import stub put into ``sys.modules`` -- does not count.)"""
    try:
        import h5py  # type: ignore
        if hasattr(h5py, "File"):
            return h5py
    except ImportError:
        pass
    from . import hdf5_mini
    return hdf5_mini


class HDF5VolumeStore(VolumeStore):
    """Reader for the reference's files: ``X`` volumes in one HDF5 (``alldatain8bitsnormalized.h5``),
    ``y``/``site`` in another (``dataset_all_labels_site.h5``; ``ABCD/data_loader.py:105-141``).
    The file is opened once, not per batch.  Uses ``h5py`` when installed, else the built-in reader (:mod:`.hdf5_mini`: contiguous
    and chunked / gzip / shuffle / fletcher32 layouts, chunk row-blocks inflated on several threads).

    ``preload=True`` reads every volume once into host memory as 8 bits (pinned when CUDA is there) and serves batches from it: the
    stagers then treat the store like the synthetic one -- row copies straight out of the pinned cache, zero-run-skipping uploads,
    per-loader device caches -- instead of going back to the file (and, for compressed files, inflating again) for every batch.
    The reference re-reads every batch through h5py (``ABCD/data_loader.py:105-141``).  Every rank holds the whole cache: meant for
    datasets that fit in host memory ``world_size`` times (``--preload``)."""

    def __init__(self, volumes_path: str, labels_path: Optional[str] = None, scale: Optional[float] = None, preload: bool = False,
                 pin: bool = True, block_rows: int = 64):
        h5py = hdf5_backend()
        self._f = h5py.File(volumes_path, "r")
        self._X = self._f["X"]
        lab = h5py.File(labels_path, "r") if labels_path else self._f
        self.labels = np.asarray(lab["y"][()], dtype=np.float32)
        self.sites = np.asarray(lab["site"][()], dtype=np.int64) if "site" in lab else np.zeros(len(self.labels), np.int64)
        self.shape = tuple(self._X.shape[1:])
        self._is_u8 = self._X.dtype == np.uint8
        if scale is not None:
            self.scale = scale
        elif not self._is_u8:
            self.scale = 1.0 / 255.0
        self._cache = None
        if preload:
            n = len(self.labels)
            buf = torch.empty((n,) + self.shape, dtype=torch.uint8)
            if pin and torch.cuda.is_available():
                buf = buf.pin_memory()
            for a in range(0, n, block_rows):
                b = min(n, a + block_rows)
                xb = np.asarray(self._X[a:b])
                if not self._is_u8:
                    xb = np.clip(np.rint(xb * 255.0), 0, 255).astype(np.uint8)
                buf[a:b] = torch.from_numpy(xb)
            self._cache = buf

    def fetch(self, indices, out=None):
        """Rows in the order requested.  HDF5 fancy indexing wants increasing, unique indices: read sorted, then un-permute."""
        idx = np.asarray(indices).astype(np.int64)
        if self._cache is not None:
            ti = torch.from_numpy(idx)
            y = torch.from_numpy(self.labels[idx])
            if out is None:
                return self._cache.index_select(0, ti), y
            x = out[:len(idx)]
            torch.index_select(self._cache, 0, ti, out=x)
            return x, y
        uniq, inverse = np.unique(idx, return_inverse=True)
        xb = np.asarray(self._X[uniq])[inverse]
        if not self._is_u8:
            xb = np.clip(np.rint(xb * 255.0), 0, 255).astype(np.uint8)
        x = torch.from_numpy(xb)
        if out is not None:
            out[:len(idx)].copy_(x)
            x = out[:len(idx)]
        return x, torch.from_numpy(self.labels[idx])
