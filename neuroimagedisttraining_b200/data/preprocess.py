"""ABCD preprocessing (reference: ``Preprocess_ABCD.ipynb`` cells 2-37): SPM grey-matter maps ``Sm6mwc1pT1.nii`` of
121x145x121 voxels -> brain mask where the group mean exceeds 0.2 -> per-subject min-max normalisation -> 8-bit quantisation
-> one container with ``X`` (uint8 volumes), ``y`` (``female``), ``site`` (label-encoded ``abcd_site``).

NIfTI files are read with ``nibabel`` when present, else with the built-in single-file NIfTI-1 reader; the reference's HDF5 container is
read / written with ``h5py`` when present, else with :mod:`.hdf5_mini` (plain contiguous files); the numerical steps are plain numpy and the ``.npz`` container works everywhere."""
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


# ------------------------------------------------------------------------------------------------------ NIfTI-1
_NIFTI_DTYPES = {2: np.uint8, 4: np.int16, 8: np.int32, 16: np.float32, 64: np.float64, 256: np.int8, 512: np.uint16, 768: np.uint32,
                 1024: np.int64, 1280: np.uint64}


def _read_nifti1(raw: bytes) -> np.ndarray:
    """Single-file NIfTI-1 (``.nii``): 348-byte header (dims at 40, datatype at 70, vox_offset at 108, scl_slope / scl_inter at 112 /
    116, magic ``n+1`` at 344), voxels in Fortran order from ``vox_offset``.  Either byte order; the intensity scaling is applied
    when the slope is set, which is what ``nibabel``'s ``get_fdata()`` returns."""
    if len(raw) < 348:
        raise ValueError("not a NIfTI-1 file: shorter than its header")
    endian = "<" if int(np.frombuffer(raw, "<i4", 1, 0)[0]) == 348 else ">"
    if int(np.frombuffer(raw, endian + "i4", 1, 0)[0]) != 348 or raw[344:347] not in (b"n+1", b"ni1"):
        raise ValueError("not a NIfTI-1 file (sizeof_hdr / magic)")
    if raw[344:347] == b"ni1":
        raise ValueError("two-file NIfTI (.hdr/.img) is not supported; convert to single-file .nii")
    dim = np.frombuffer(raw, endian + "i2", 8, 40)
    shape = tuple(int(d) for d in dim[1:1 + int(dim[0])])
    code = int(np.frombuffer(raw, endian + "i2", 1, 70)[0])
    if code not in _NIFTI_DTYPES:
        raise ValueError(f"unsupported NIfTI datatype code {code}")
    dt = np.dtype(_NIFTI_DTYPES[code]).newbyteorder(endian)
    vox_offset, slope, inter = (float(v) for v in np.frombuffer(raw, endian + "f4", 3, 108))
    off = int(vox_offset) if vox_offset >= 348 else 352
    n = int(np.prod(shape))
    data = np.frombuffer(raw, dt, n, off).reshape(shape, order="F").astype(np.float32)
    if slope != 0.0 and np.isfinite(slope) and np.isfinite(inter) and (slope != 1.0 or inter != 0.0):
        data = data * np.float32(slope) + np.float32(inter)
    while data.ndim > 3 and data.shape[-1] == 1:          # SPM writes 3-D maps with trailing singleton dims
        data = data[..., 0]
    return data


def load_nifti(path: str) -> np.ndarray:
    """Volume of a ``.nii`` / ``.nii.gz`` file as float32 (the notebook's ``nib.load(p).get_fdata()``, cells 6-11).  Uses ``nibabel``
    when it is installed, otherwise the built-in NIfTI-1 reader."""
    try:
        import nibabel as nib  # type: ignore
        return np.asarray(nib.load(path).get_fdata(), dtype=np.float32)
    except ImportError:
        pass
    import gzip
    with open(path, "rb") as f:
        raw = f.read()
    if raw[:2] == b"\x1f\x8b":
        raw = gzip.decompress(raw)
    return _read_nifti1(raw)


def save_nifti(volume: np.ndarray, path: str, voxel_size=(1.5, 1.5, 1.5)) -> None:
    """Minimal single-file NIfTI-1 writer (float32 / uint8 / int16 volumes; ``.gz`` suffix compresses) -- enough to exchange
    preprocessed maps and to test the reader."""
    v = np.asarray(volume)
    code = {np.dtype(np.float32): 16, np.dtype(np.uint8): 2, np.dtype(np.int16): 4, np.dtype(np.float64): 64}.get(v.dtype)
    if code is None:
        v, code = v.astype(np.float32), 16
    hdr = bytearray(352)
    hdr[0:4] = np.int32(348).tobytes()
    dim = np.ones(8, np.int16)
    dim[0] = v.ndim
    dim[1:1 + v.ndim] = v.shape
    hdr[40:56] = dim.tobytes()
    hdr[70:72] = np.int16(code).tobytes()
    hdr[72:74] = np.int16(v.dtype.itemsize * 8).tobytes()
    pix = np.ones(8, np.float32)
    pix[1:4] = voxel_size
    hdr[76:108] = pix.tobytes()
    hdr[108:112] = np.float32(352).tobytes()
    hdr[112:116] = np.float32(1.0).tobytes()
    hdr[344:348] = b"n+1\x00"
    blob = bytes(hdr) + v.astype(v.dtype.newbyteorder("<")).tobytes(order="F")
    if path.endswith(".gz"):
        import gzip
        blob = gzip.compress(blob, 1)
    with open(path, "wb") as f:
        f.write(blob)


def build_store_from_nifti(paths: Sequence[str], labels: Sequence, site_names: Sequence, threshold: float = 0.2) -> InMemoryVolumeStore:
    """The notebook end to end: read every subject's map, group-mean mask, per-subject min-max, 8-bit, label-encoded sites."""
    return build_store([load_nifti(p) for p in paths], labels, site_names, threshold)


def save_npz(store: InMemoryVolumeStore, path: str) -> None:
    np.savez_compressed(path, X=store.volumes.numpy(), y=store.labels, site=store.sites)


def load_npz(path: str, pin: bool = False) -> InMemoryVolumeStore:
    d = np.load(path)
    return InMemoryVolumeStore(torch.from_numpy(d["X"]), d["y"], d["site"], pin=pin)


def save_hdf5(store: InMemoryVolumeStore, path: str) -> None:
    """``X`` / ``y`` / ``site`` in one HDF5 file, the container the reference's loaders read (``ABCD/data_loader.py:105-141``).
    With ``h5py``: one chunk per volume, LZF-compressed.  Without: the built-in writer (contiguous, uncompressed)."""
    from .volumes import hdf5_backend
    h5py = hdf5_backend()
    if not hasattr(h5py, "special_dtype"):               # the in-tree subset
        h5py.write_file(path, {"X": store.volumes.numpy(), "y": np.asarray(store.labels), "site": np.asarray(store.sites)})
        return
    with h5py.File(path, "w") as f:
        f.create_dataset("X", data=store.volumes.numpy(), chunks=(1,) + tuple(store.shape), compression="lzf")
        f.create_dataset("y", data=store.labels)
        f.create_dataset("site", data=store.sites)
