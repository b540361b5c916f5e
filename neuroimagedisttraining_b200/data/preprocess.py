"""ABCD preprocessing (reference: ``Preprocess_ABCD.ipynb`` cells 2-37): SPM grey-matter maps ``Sm6mwc1pT1.nii`` of
121x145x121 voxels -> brain mask where the group mean exceeds 0.2 -> per-subject min-max normalisation -> 8-bit quantisation
-> one container with ``X`` (uint8 volumes), ``y`` (``female``), ``site`` (label-encoded ``abcd_site``).

NIfTI / HDF5 I/O needs ``nibabel`` / ``h5py`` (neither is in this image): the numerical steps are plain numpy and the
``.npz`` container works everywhere."""
from __future__ import annotations

from typing import Iterable, Optional, Sequence, Tuple

import numpy as np
import torch

from .volumes import InMemoryVolumeStore


def group_mean_mask(volumes: Iterable[np.ndarray], threshold: float = 0.2) -> np.ndarray:
    """Boolean brain mask: voxels whose mean over subjects exceeds ``threshold`` (streams over the subjects)."""
    acc, n = None, 0
    for v in volumes:
        acc = v.astype(np.float64) if acc is None else acc + v
        n += 1
    return (acc / max(n, 1)) > threshold


def minmax_quantize_u8(volume: np.ndarray, mask: Optional[np.ndarray] = None) -> np.ndarray:
    """Per-subject min-max over the masked voxels, then 8-bit quantisation; voxels outside the mask become 0."""
    v = volume.astype(np.float32)
    inside = v[mask] if mask is not None else v.ravel()
    lo, hi = float(inside.min()), float(inside.max())
    out = np.zeros(v.shape, dtype=np.uint8)
    if hi > lo:
        q = np.clip(np.rint((v - lo) / (hi - lo) * 255.0), 0, 255).astype(np.uint8)
        if mask is not None:
            out[mask] = q[mask]
        else:
            out = q
    return out


def encode_sites(site_names: Sequence) -> Tuple[np.ndarray, list]:
    """Label-encode site identifiers (sorted unique order, as ``sklearn.LabelEncoder`` does)."""
    classes = sorted(set(site_names))
    lut = {s: i for i, s in enumerate(classes)}
    return np.asarray([lut[s] for s in site_names], dtype=np.int64), classes


def build_store(volumes: Sequence[np.ndarray], labels: Sequence, site_names: Sequence, threshold: float = 0.2) -> InMemoryVolumeStore:
    mask = group_mean_mask(volumes, threshold)
    x = np.stack([minmax_quantize_u8(v, mask) for v in volumes])
    sites, _ = encode_sites(site_names)
    return InMemoryVolumeStore(torch.from_numpy(x), np.asarray(labels, dtype=np.float32), sites)


def load_nifti(path: str) -> np.ndarray:  # pragma: no cover - optional dependency
    try:
        import nibabel as nib  # type: ignore
    except ImportError as e:
        raise ImportError("reading NIfTI files needs nibabel") from e
    return np.asarray(nib.load(path).get_fdata(), dtype=np.float32)


def save_npz(store: InMemoryVolumeStore, path: str) -> None:
    np.savez_compressed(path, X=store.volumes.numpy(), y=store.labels, site=store.sites)


def load_npz(path: str, pin: bool = False) -> InMemoryVolumeStore:
    d = np.load(path)
    return InMemoryVolumeStore(torch.from_numpy(d["X"]), d["y"], d["site"], pin=pin)


def save_hdf5(store: InMemoryVolumeStore, path: str) -> None:  # pragma: no cover - optional dependency
    try:
        import h5py  # type: ignore
    except ImportError as e:
        raise ImportError("writing the reference's HDF5 layout needs h5py") from e
    with h5py.File(path, "w") as f:
        f.create_dataset("X", data=store.volumes.numpy(), chunks=(1,) + tuple(store.shape), compression="lzf")
        f.create_dataset("y", data=store.labels)
        f.create_dataset("site", data=store.sites)
