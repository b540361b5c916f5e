"""Built-in HDF5 subset (``data/hdf5_mini.py``): round trips, memory-mapped row access, the volume store and the SalientGrads entry
point on a file, and clear errors for what lies outside the subset.  (No HDF5 library is available here: reader and writer are
checked against each other and against byte-level expectations from the format specification.)"""
import os
import struct

import numpy as np
import pytest
import torch

from neuroimagedisttraining_b200.data import HDF5VolumeStore, SyntheticABCD
from neuroimagedisttraining_b200.data import hdf5_mini as H
from neuroimagedisttraining_b200.data import preprocess as P
from neuroimagedisttraining_b200.data.volumes import InMemoryVolumeStore


def test_round_trip_dtypes_shapes_and_row_access(tmp_path):
    rng = np.random.RandomState(0)
    data = {"X": rng.randint(0, 256, size=(9, 4, 5, 3)).astype(np.uint8), "y": (rng.rand(9) > 0.5).astype(np.float64),
            "site": np.repeat(np.arange(3), 3).astype(np.int64), "f32": rng.rand(2, 3).astype(np.float32), "i16": np.arange(-3, 4, dtype=np.int16),
            "scalar": np.float64(2.5), "empty": np.zeros((0, 4), np.float32)}
    path = str(tmp_path / "t.h5")
    H.write_file(path, data)
    raw = open(path, "rb").read()
    assert raw[:8] == b"\x89HDF\r\n\x1a\n" and raw[8] == 0 and raw[13:15] == b"\x08\x08"           # signature, superblock v0, 8-byte offsets
    assert struct.unpack_from("<Q", raw, 40)[0] == len(raw)                                           # end-of-file address
    assert raw.count(b"TREE") >= 1 and raw.count(b"SNOD") >= 1 and raw.count(b"HEAP") >= 1
    with H.File(path) as f:
        assert sorted(f.keys()) == sorted(data) and "X" in f and "nope" not in f
        for k, v in data.items():
            got = f[k][()]
            assert got.dtype == np.asarray(v).dtype and got.shape == np.asarray(v).shape and np.array_equal(got, v), k
        X = f["X"]
        assert X.shape == (9, 4, 5, 3) and len(X) == 9 and X.dtype == np.uint8
        assert np.array_equal(X[[1, 4, 8]], data["X"][[1, 4, 8]]) and np.array_equal(X[2:5], data["X"][2:5]) and np.array_equal(X[7], data["X"][7])
        with pytest.raises(TypeError, match="increasing order"):
            X[[4, 1]]
        assert isinstance(X._a, np.memmap)                                                            # rows are read on demand
    H.write_file(path, data, split_headers=True)            # layout messages in object-header continuation blocks
    with H.File(path) as f:
        assert all(np.array_equal(f[k][()], v) for k, v in data.items())


def test_volume_store_and_cli_on_a_file(tmp_path):
    src = SyntheticABCD(24, (32, 32, 32), n_sites=3, seed=2, site_sizes=[8, 8, 8])
    mem = InMemoryVolumeStore(src.fetch(np.arange(24))[0], src.labels, src.sites)
    path = str(tmp_path / "abcd.h5")
    P.save_hdf5(mem, path)
    store = HDF5VolumeStore(path)
    assert store.shape == (32, 32, 32) and np.array_equal(store.labels, src.labels) and np.array_equal(store.sites, src.sites)
    x, y = store.fetch([20, 3, 11])
    want, _ = src.fetch([20, 3, 11])
    assert torch.equal(x, want) and list(y) == list(src.labels[[20, 3, 11]])
    # volumes and labels in separate files, as the reference keeps them
    H.write_file(str(tmp_path / "vol.h5"), {"X": mem.volumes.numpy()})
    H.write_file(str(tmp_path / "lab.h5"), {"y": src.labels.astype(np.float64), "site": src.sites})
    two = HDF5VolumeStore(str(tmp_path / "vol.h5"), str(tmp_path / "lab.h5"))
    assert torch.equal(two.fetch([5])[0], src.fetch([5])[0]) and np.array_equal(two.sites, src.sites)
    from neuroimagedisttraining_b200.cli import main_sailentgrads
    stat = main_sailentgrads.main(["--model", "tiny3DCNN", "--in_shape", "32,32,32", "--data_dir", str(tmp_path / "vol.h5"), "--labels_file",
                                   str(tmp_path / "lab.h5"), "--client_num_in_total", "3", "--batch_size", "4", "--epochs", "1", "--comm_round", "1",
                                   "--device", "cpu", "--log_dir", str(tmp_path / "LOG")])
    assert len(stat["global_test_acc"]) == 2
    # --preload: the same job served from the in-memory copy of the file gives the same result
    stat2 = main_sailentgrads.main(["--model", "tiny3DCNN", "--in_shape", "32,32,32", "--data_dir", str(tmp_path / "vol.h5"), "--labels_file",
                                    str(tmp_path / "lab.h5"), "--client_num_in_total", "3", "--batch_size", "4", "--epochs", "1", "--comm_round", "1",
                                    "--device", "cpu", "--log_dir", str(tmp_path / "LOG2"), "--preload"])
    assert [float(a) for a in stat2["global_test_acc"]] == [float(a) for a in stat["global_test_acc"]]


def test_features_outside_the_subset_are_refused_with_a_reason(tmp_path):
    path = str(tmp_path / "t.h5")
    H.write_file(path, {"X": np.arange(12, dtype=np.int32).reshape(3, 4)})
    raw = bytearray(open(path, "rb").read())
    with pytest.raises(H.UnsupportedHDF5, match="not an HDF5 file"):
        (tmp_path / "junk.h5").write_bytes(b"junk" * 40)
        H.File(str(tmp_path / "junk.h5"))
    v2 = bytearray(raw); v2[8] = 2
    (tmp_path / "v2.h5").write_bytes(bytes(v2))
    with pytest.raises(H.UnsupportedHDF5, match="superblock version 2"):
        H.File(str(tmp_path / "v2.h5"))
    # an unknown data layout class (3 = virtual): version 3 layout message = type 8, body starts 03 01
    i = bytes(raw).index(struct.pack("<HH", 8, 24)) + 8
    assert raw[i] == 3 and raw[i + 1] == 1
    ch = bytearray(raw); ch[i + 1] = 3
    (tmp_path / "virtual.h5").write_bytes(bytes(ch))
    with pytest.raises(H.UnsupportedHDF5, match="layout class 3"):
        H.File(str(tmp_path / "virtual.h5"))
    # a filter the reader does not implement (szip = 4) is refused when the dataset is opened, not decoded into garbage
    H.write_file(path, {"X": np.arange(12, dtype=np.int32).reshape(3, 4)}, chunks={"X": (1, 4)}, compression="gzip")
    z = bytearray(open(path, "rb").read())
    j = bytes(z).index(struct.pack("<HHHH", H.FILTER_DEFLATE, 0, 0, 1))
    z[j] = 4
    (tmp_path / "szip.h5").write_bytes(bytes(z))
    with pytest.raises(H.UnsupportedHDF5, match="filter 4"):
        H.File(str(tmp_path / "szip.h5"))
    (tmp_path / "short.h5").write_bytes(bytes(raw[:200]))
    with pytest.raises(H.UnsupportedHDF5, match="truncated"):
        H.File(str(tmp_path / "short.h5"))
    with pytest.raises(H.UnsupportedHDF5, match="subset"):
        H.write_file(path, {"s": np.array(["a", "b"])})


@pytest.mark.parametrize("compression,shuffle,chunk", [(None, False, (1, 9, 11, 7)), ("gzip", True, (4, 5, 11, 4)), ("gzip", False, (8, 9, 6, 7))])
def test_chunked_and_compressed_datasets_round_trip(tmp_path, compression, shuffle, chunk):
    """Chunked storage with the shuffle + deflate pipeline (what ``h5py.create_dataset(..., chunks=..., compression='gzip', shuffle=True)``
    writes): full reads, single rows, slices and sorted index lists decode only the chunks they need; a chunk index with more than 64
    chunks is a two-level B-tree; the volume store serves batches from such a file."""
    from neuroimagedisttraining_b200.data.volumes import HDF5VolumeStore
    rng = np.random.RandomState(0)
    X = (rng.rand(37, 9, 11, 7) * 255).astype(np.uint8)
    y = (rng.rand(37) > 0.5).astype(np.float32)
    site = rng.randint(0, 3, 37).astype(np.int64)
    Z = rng.randn(300, 5)
    path = str(tmp_path / "c.h5")
    H.write_file(path, {"X": X, "y": y, "site": site, "Z": Z}, chunks={"X": chunk, "Z": (2, 3)}, compression=compression, shuffle=shuffle)
    with H.File(path) as f:
        assert isinstance(f["X"], H.ChunkedDataset) and f["X"].chunks == chunk and f["X"].shape == X.shape
        assert np.array_equal(f["X"][()], X) and np.array_equal(f["Z"][()], Z) and np.array_equal(f["y"][()], y)
        assert np.array_equal(f["X"][[1, 5, 6, 30, 36]], X[[1, 5, 6, 30, 36]])
        assert np.array_equal(f["X"][3], X[3]) and np.array_equal(f["X"][-1], X[-1]) and np.array_equal(f["X"][10:20], X[10:20])
        assert len(f["Z"]._index) == 300
        with pytest.raises(TypeError):
            f["X"][[5, 1]]
    if compression == "gzip":
        assert os.path.getsize(path) < X.nbytes + Z.nbytes + 40000
    store = HDF5VolumeStore(path)
    xb, yb = store.fetch([2, 7, 11])
    assert np.array_equal(xb.numpy(), X[[2, 7, 11]]) and np.array_equal(yb.numpy(), y[[2, 7, 11]])


def test_preloaded_store_serves_the_same_volumes_and_parallel_decode_matches(tmp_path):
    """``HDF5VolumeStore(preload=True)`` (the whole file read once into host memory, chunk row-blocks inflated on several threads)
    hands out exactly what the per-batch reader does, in any index order; the threaded decode equals the one-thread decode."""
    import time
    from neuroimagedisttraining_b200.data import hdf5_mini
    from neuroimagedisttraining_b200.data.volumes import HDF5VolumeStore
    g = np.random.default_rng(3)
    n, shape = 24, (20, 24, 22)
    X = g.integers(0, 256, size=(n,) + shape, dtype=np.uint8)
    X[:, :, :, :6] = 0
    y = (g.random(n) > 0.5).astype(np.float32)
    site = np.repeat(np.arange(4), 6).astype(np.int64)
    p = str(tmp_path / "vol.h5")
    hdf5_mini.write_file(p, {"X": X, "y": y, "site": site}, chunks={"X": (2,) + shape}, compression="gzip", shuffle=True)
    lazy = HDF5VolumeStore(p)
    t0 = time.time()
    pre = HDF5VolumeStore(p, preload=True, block_rows=8)
    t_pre = time.time() - t0
    assert pre._cache is not None and tuple(pre._cache.shape) == (n,) + shape and t_pre < 60
    for idx in ([5, 1, 1, 23, 0], list(range(n)), [7]):
        a, ya = lazy.fetch(idx)
        b, yb = pre.fetch(idx)
        assert torch.equal(a, b) and torch.equal(ya, yb)
        assert np.array_equal(b.numpy(), X[np.asarray(idx)])
    out = torch.empty((8,) + shape, dtype=torch.uint8)
    b, _ = pre.fetch([3, 2], out=out)
    assert np.array_equal(b.numpy(), X[[3, 2]]) and b.data_ptr() == out.data_ptr()
    # threaded == single-threaded decode
    ds = hdf5_mini.File(p)["X"]
    rows = np.arange(0, n, 3)
    many = ds[rows]
    old, hdf5_mini.READ_THREADS = hdf5_mini.READ_THREADS, 1
    try:
        one = hdf5_mini.File(p)["X"][rows]
    finally:
        hdf5_mini.READ_THREADS = old
    assert np.array_equal(many, one) and np.array_equal(many, X[rows])
