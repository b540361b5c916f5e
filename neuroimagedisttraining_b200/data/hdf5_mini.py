"""A small reader / writer for the subset of HDF5 that ``h5py.File(...).create_dataset(...)`` produces with the default library
version: version-0 superblock, old-style groups (symbol table: v1 B-tree + local heap), version-1 object headers, and numeric
datasets stored contiguously, compactly, or CHUNKED with the shuffle / deflate (gzip) / fletcher32 filters (version-1 chunk B-tree).
That covers the reference's data files (``Preprocess_ABCD.ipynb`` cell 30: ``dataset_all_labels_site.h5``;
``alldatain8bitsnormalized.h5`` with ``X`` / ``y``) however they were written: plain, or with ``chunks=`` / ``compression='gzip'``.

It exists because ``h5py`` is not available everywhere this framework runs; :class:`~.volumes.HDF5VolumeStore` and
``preprocess.save_hdf5`` use ``h5py`` when it is installed and this module otherwise.  Contiguous datasets are memory-mapped and
chunked ones decode only the chunks a request touches, so indexing a 24 GB volume file reads only the requested rows.  Anything
outside the subset (other filters such as szip / lzf, new-style groups from ``libver='latest'``, compound / string types) raises
:class:`UnsupportedHDF5` with the reason, not garbage.

NOTE: no HDF5 library exists in the image this was developed in, so the reader has only been exercised against files from the
writer below, which follows the published file-format specification (HDF5 File Format Specification, version 1.1 structures).
"""
from __future__ import annotations

import os
import struct
import zlib
from concurrent.futures import ThreadPoolExecutor
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np

SIGNATURE = b"\x89HDF\r\n\x1a\n"
UNDEF = 0xFFFFFFFFFFFFFFFF
MSG_NIL, MSG_DATASPACE, MSG_DATATYPE, MSG_FILL, MSG_LAYOUT, MSG_CONTINUATION, MSG_SYMBOL_TABLE = 0x0, 0x1, 0x3, 0x5, 0x8, 0x10, 0x11


class UnsupportedHDF5(ValueError):
    pass


# ------------------------------------------------------------------------------------------------ datatype message
def _encode_dtype(dt: np.dtype) -> bytes:
    dt = np.dtype(dt)
    if dt.byteorder == ">":
        raise UnsupportedHDF5("the writer stores little-endian data only")
    if dt.kind in "iu":
        bits0 = 0x08 if dt.kind == "i" else 0x00
        return struct.pack("<BBBBI", 0x10, bits0, 0, 0, dt.itemsize) + struct.pack("<HH", 0, 8 * dt.itemsize)
    if dt.kind == "f" and dt.itemsize in (4, 8):
        sign, eloc, esz, msz, bias = (31, 23, 8, 23, 127) if dt.itemsize == 4 else (63, 52, 11, 52, 1023)
        return struct.pack("<BBBBI", 0x11, 0x20, sign, 0, dt.itemsize) + struct.pack("<HHBBBBI", 0, 8 * dt.itemsize, eloc, esz, 0, msz, bias)
    raise UnsupportedHDF5(f"dtype {dt} is outside the supported subset (integers, float32, float64)")


def _decode_dtype(body: bytes) -> np.dtype:
    cls, size = body[0] & 0x0F, struct.unpack_from("<I", body, 4)[0]
    order = ">" if body[1] & 0x01 else "<"
    if cls == 0:
        kind = "i" if body[1] & 0x08 else "u"
        if size not in (1, 2, 4, 8):
            raise UnsupportedHDF5(f"{size}-byte integers")
        return np.dtype(f"{order}{kind}{size}")
    if cls == 1:
        if size not in (2, 4, 8):
            raise UnsupportedHDF5(f"{size}-byte floats")
        return np.dtype(f"{order}f{size}")
    names = {2: "time", 3: "string", 4: "bitfield", 5: "opaque", 6: "compound", 7: "reference", 8: "enum", 9: "variable-length", 10: "array"}
    raise UnsupportedHDF5(f"datatype class {cls} ({names.get(cls, '?')}) is outside the supported subset (integers and floats)")


# ------------------------------------------------------------------------------------------------------- reader
class Dataset:
    """Read-only, numpy-like view of one dataset (``shape``, ``dtype``, ``ds[()]``, ``ds[i]``, ``ds[a:b]``, ``ds[sorted_indices]``)."""

    def __init__(self, name: str, array: np.ndarray):
        self.name, self._a = name, array
        self.shape, self.dtype = array.shape, array.dtype

    def __len__(self):
        return self.shape[0]

    def __getitem__(self, key):
        if isinstance(key, tuple) and key == ():
            return np.array(self._a)
        if isinstance(key, (list, np.ndarray)):
            k = np.asarray(key)
            if k.ndim == 1 and k.dtype != bool and (np.diff(k) <= 0).any():
                raise TypeError("Indexing elements must be in increasing order")        # h5py's rule, kept so code stays portable
        return np.array(self._a[key])

    def __array__(self, dtype=None, copy=None):
        return np.asarray(self._a, dtype=dtype)


FILTER_DEFLATE, FILTER_SHUFFLE, FILTER_FLETCHER32 = 1, 2, 3


def _parse_filters(body: bytes) -> List[Tuple[int, Tuple[int, ...]]]:
    """Filter pipeline message (0x000B), versions 1 and 2 -> [(filter id, client data), ...] in the order applied on write."""
    version, n = body[0], body[1]
    q = 8 if version == 1 else 2
    out = []
    for _ in range(n):
        fid, = struct.unpack_from("<H", body, q)
        q += 2
        name_len = 0
        if version == 1 or fid >= 256:
            name_len, = struct.unpack_from("<H", body, q)
            q += 2
        _flags, ncd = struct.unpack_from("<HH", body, q)
        q += 4
        if name_len:
            q += name_len if version == 2 else (name_len + 7) // 8 * 8
        cd = struct.unpack_from(f"<{ncd}I", body, q)
        q += 4 * ncd
        if version == 1 and ncd % 2:
            q += 4
        if fid not in (FILTER_DEFLATE, FILTER_SHUFFLE, FILTER_FLETCHER32):
            raise UnsupportedHDF5(f"filter {fid} (only deflate / shuffle / fletcher32 are supported)")
        out.append((fid, tuple(cd)))
    return out


READ_THREADS = max(1, min(8, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))


class ChunkedDataset:
    """Read-only view of a chunked (optionally shuffled / deflated / checksummed) dataset with the same face as :class:`Dataset`.
    The chunk index (a version-1 B-tree) is walked once; ``ds[i]``, ``ds[a:b]`` and ``ds[sorted_indices]`` decode only the chunks that
    hold the requested rows -- the access pattern of the volume stores (``X[sorted batch indices]``)."""

    def __init__(self, name: str, file: "File", shape, dtype: np.dtype, chunk, btree_addr: int, filters):
        self.name, self._file = name, file
        self.shape, self.dtype = tuple(int(s) for s in shape), np.dtype(dtype)
        self.chunks = tuple(int(c) for c in chunk)
        self._filters = filters
        self._index: Dict[Tuple[int, ...], Tuple[int, int, int]] = {}
        if btree_addr != UNDEF:
            self._walk(btree_addr)
        self._cache: Tuple[Optional[Tuple[int, ...]], Optional[np.ndarray]] = (None, None)

    def __len__(self):
        return self.shape[0]

    def _walk(self, addr: int) -> None:
        f = self._file
        a = f._base + addr
        if f._read(a, 4) != b"TREE":
            raise UnsupportedHDF5("bad chunk B-tree node signature")
        node_type, level, used = struct.unpack("<BBH", f._read(a + 4, 4))
        if node_type != 1:
            raise UnsupportedHDF5("unexpected B-tree node type in a chunk index")
        rank1 = len(self.shape) + 1
        key = 8 + 8 * rank1
        body = f._read(a + 24, used * (key + 8) + key)
        for i in range(used):
            q = i * (key + 8)
            size, mask = struct.unpack_from("<II", body, q)
            offs = struct.unpack_from(f"<{rank1}Q", body, q + 8)
            child, = struct.unpack_from("<Q", body, q + key)
            if level > 0:
                self._walk(child)
            else:
                self._index[tuple(int(o) for o in offs[:-1])] = (int(child), int(size), int(mask))

    def _chunk(self, origin: Tuple[int, ...]) -> np.ndarray:
        if self._cache[0] == origin:
            return self._cache[1]
        arr = self._decode(origin)
        self._cache = (origin, arr)
        return arr

    def _decode(self, origin: Tuple[int, ...]) -> np.ndarray:
        """One chunk, read and run back through the filter pipeline (no shared state: callable from several threads)."""
        n = int(np.prod(self.chunks))
        ent = self._index.get(origin)
        if ent is None:                                   # never written: fill value (0)
            arr = np.zeros(self.chunks, self.dtype)
        else:
            addr, size, mask = ent
            buf = self._file._read(self._file._base + addr, size)
            for k in range(len(self._filters) - 1, -1, -1):      # undo the pipeline back to front; a set mask bit = filter skipped
                if mask & (1 << k):
                    continue
                fid = self._filters[k][0]
                if fid == FILTER_DEFLATE:
                    buf = zlib.decompress(buf)
                elif fid == FILTER_FLETCHER32:
                    buf = buf[:-4]
                elif fid == FILTER_SHUFFLE:
                    isz = self.dtype.itemsize
                    buf = np.frombuffer(buf, np.uint8).reshape(isz, -1).T.tobytes() if isz > 1 else buf
            arr = np.frombuffer(buf, self.dtype, count=n).reshape(self.chunks)
        return arr

    def _rows(self, rows: np.ndarray) -> np.ndarray:
        out = np.empty((len(rows),) + self.shape[1:], self.dtype)
        if len(rows) == 0 or out.size == 0:
            return out
        c0 = self.chunks[0]
        grids = [range(0, s, c) for s, c in zip(self.shape[1:], self.chunks[1:])]
        blocks: Dict[int, List[Tuple[int, int]]] = {}                 # first row of a chunk row-block -> [(position in out, row)]
        for j, r in enumerate(rows):
            r = int(r)
            if not 0 <= r < self.shape[0]:
                raise IndexError(f"row {r} out of range")
            blocks.setdefault(r // c0 * c0, []).append((j, r))

        def fill(base0: int, cached: bool) -> None:
            for rest in np.ndindex(*[len(g) for g in grids]) if grids else [()]:
                origin = (base0,) + tuple(g[i] for g, i in zip(grids, rest))
                ch = self._chunk(origin) if cached else self._decode(origin)
                sl_out = tuple(slice(o, min(o + c, s)) for o, c, s in zip(origin[1:], self.chunks[1:], self.shape[1:]))
                sl_in = tuple(slice(0, s.stop - s.start) for s in sl_out)
                for j, r in blocks[base0]:
                    out[(j,) + sl_out] = ch[(r - base0,) + sl_in]

        threads = min(READ_THREADS, len(blocks))
        if threads > 1 and self._filters:             # inflating is the cost, and zlib releases the GIL: decode row-blocks side by side
            with ThreadPoolExecutor(threads) as pool:
                list(pool.map(lambda b: fill(b, False), blocks))
        else:
            for b in blocks:
                fill(b, True)
        return out

    def __getitem__(self, key):
        n = self.shape[0] if self.shape else 1
        if isinstance(key, tuple) and key == ():
            return self._rows(np.arange(n)) if self.shape else self._chunk(())[()]
        if isinstance(key, (int, np.integer)):
            return self._rows(np.array([int(key) % n if key < 0 else int(key)]))[0]
        if isinstance(key, slice):
            return self._rows(np.arange(*key.indices(n)))
        k = np.asarray(key)
        if k.ndim == 1 and k.dtype != bool:
            if (np.diff(k) <= 0).any():
                raise TypeError("Indexing elements must be in increasing order")
            return self._rows(k)
        if k.dtype == bool:
            return self._rows(np.nonzero(k)[0])
        raise UnsupportedHDF5("only row indexing (int, slice, sorted index list) is supported on chunked datasets")

    def __array__(self, dtype=None, copy=None):
        return np.asarray(self[()], dtype=dtype)


class File:
    """``File(path)[name]`` -> :class:`Dataset`; also ``keys()``, ``in``, iteration, context manager.  Read-only."""

    def __init__(self, path: str, mode: str = "r"):
        if mode != "r":
            raise UnsupportedHDF5("hdf5_mini.File opens files read-only; use write_file() to create one")
        self.path = path
        self._f = open(path, "rb")
        self._datasets: Dict[str, Dataset] = {}
        self._parse()

    # -- low-level helpers
    def _read(self, off: int, n: int) -> bytes:
        b = os.pread(self._f.fileno(), n, off)             # positional: safe from several decode threads at once
        if len(b) != n:
            raise UnsupportedHDF5(f"truncated file: wanted {n} bytes at offset {off}")
        return b

    def _parse(self) -> None:
        head = self._read(0, 8)
        if head != SIGNATURE:
            raise UnsupportedHDF5("not an HDF5 file (signature; user blocks are not supported)")
        sb_version = self._read(8, 1)[0]
        if sb_version not in (0, 1):
            raise UnsupportedHDF5(f"superblock version {sb_version} (written with libver='latest'?); only the default old-style layout is supported")
        size_off, size_len = self._read(13, 1)[0], self._read(14, 1)[0]
        if (size_off, size_len) != (8, 8):
            raise UnsupportedHDF5(f"{size_off}-byte offsets / {size_len}-byte lengths")
        p = 24 + (4 if sb_version == 1 else 0)            # v1 adds indexed-storage K + reserved
        base, _free, _eof, _drv = struct.unpack("<4Q", self._read(p, 32))
        self._base = base
        _name_off, root_header, _cache, _res = struct.unpack("<QQII", self._read(p + 32, 24))
        btree, heap = self._symbol_table_of(root_header)
        heap_data = self._heap_data(heap)
        for name, header in self._walk_group(btree, heap_data):
            ds = self._dataset(name, header)
            if ds is not None:
                self._datasets[name] = ds

    def _messages(self, header_addr: int) -> List[Tuple[int, bytes]]:
        a = self._base + header_addr
        version, _, nmsgs, _refs, size = struct.unpack("<BBHII", self._read(a, 12))
        if version != 1:
            raise UnsupportedHDF5(f"object header version {version}; only version-1 headers are supported")
        blocks, out = [(a + 16, size)], []
        while blocks and len(out) < nmsgs:
            off, length = blocks.pop(0)
            raw, q = self._read(off, length), 0
            while q + 8 <= length and len(out) < nmsgs:
                mtype, msize, _flags = struct.unpack_from("<HHB", raw, q)
                body = raw[q + 8:q + 8 + msize]
                if mtype == MSG_CONTINUATION:
                    c_off, c_len = struct.unpack_from("<QQ", body)
                    blocks.append((self._base + c_off, c_len))
                out.append((mtype, body))
                q += 8 + msize
        return out

    def _symbol_table_of(self, header_addr: int) -> Tuple[int, int]:
        for mtype, body in self._messages(header_addr):
            if mtype == MSG_SYMBOL_TABLE:
                return struct.unpack_from("<QQ", body)
        raise UnsupportedHDF5("group without a symbol table (new-style group); only old-style groups are supported")

    def _heap_data(self, heap_addr: int) -> bytes:
        a = self._base + heap_addr
        if self._read(a, 4) != b"HEAP":
            raise UnsupportedHDF5("bad local heap signature")
        seg_size, _free, seg_addr = struct.unpack("<QQQ", self._read(a + 8, 24))
        return self._read(self._base + seg_addr, seg_size)

    def _walk_group(self, node_addr: int, heap: bytes) -> Iterator[Tuple[str, int]]:
        a = self._base + node_addr
        sig = self._read(a, 4)
        if sig == b"TREE":
            node_type, _level, used = struct.unpack("<BBH", self._read(a + 4, 4))
            if node_type != 0:
                raise UnsupportedHDF5("unexpected B-tree node type in a group")
            body = self._read(a + 24, 8 + 16 * used)      # key0, (child, key) * used
            for i in range(used):
                child = struct.unpack_from("<Q", body, 8 + 16 * i)[0]
                yield from self._walk_group(child, heap)
        elif sig == b"SNOD":
            count = struct.unpack("<H", self._read(a + 6, 2))[0]
            for i in range(count):
                name_off, header = struct.unpack("<QQ", self._read(a + 8 + 40 * i, 16))
                end = heap.index(b"\x00", name_off)
                yield heap[name_off:end].decode(), header
        else:
            raise UnsupportedHDF5(f"unexpected group node signature {sig!r}")

    def _dataset(self, name: str, header_addr: int) -> Optional[Dataset]:
        shape = dtype = layout = None
        filters: List[Tuple[int, Tuple[int, ...]]] = []
        for mtype, body in self._messages(header_addr):
            if mtype == MSG_SYMBOL_TABLE:
                return None                                # a sub-group: not part of the flat files this reads
            if mtype == MSG_DATASPACE:
                version, rank = body[0], body[1]
                if version == 1:
                    shape = struct.unpack_from(f"<{rank}Q", body, 8) if rank else ()
                elif version == 2:
                    shape = struct.unpack_from(f"<{rank}Q", body, 4) if rank else ()
                else:
                    raise UnsupportedHDF5(f"dataspace message version {version}")
            elif mtype == MSG_DATATYPE:
                dtype = _decode_dtype(body)
            elif mtype == MSG_LAYOUT:
                layout = body
            elif mtype == 0x0B:
                filters = _parse_filters(body)
        if shape is None or dtype is None or layout is None:
            return None
        version, cls = layout[0], layout[1]
        if version != 3:
            raise UnsupportedHDF5(f"dataset {name!r}: data layout message version {version}")
        n = int(np.prod(shape, dtype=np.int64)) if len(shape) else 1
        if cls == 1:
            addr, _size = struct.unpack_from("<QQ", layout, 2)
            if addr == UNDEF or n == 0:
                arr = np.zeros(shape, dtype)
            else:
                arr = np.memmap(self.path, dtype=dtype, mode="r", offset=self._base + addr, shape=tuple(int(s) for s in shape), order="C")
        elif cls == 0:
            size = struct.unpack_from("<H", layout, 2)[0]
            arr = np.frombuffer(layout[4:4 + size], dtype=dtype, count=n).reshape(shape)
        elif cls == 2:
            rank1 = layout[2]
            if rank1 != len(shape) + 1:
                raise UnsupportedHDF5(f"dataset {name!r}: chunk rank {rank1 - 1} does not match the dataspace rank {len(shape)}")
            btree, = struct.unpack_from("<Q", layout, 3)
            dims = struct.unpack_from(f"<{rank1}I", layout, 11)
            return ChunkedDataset(name, self, shape, dtype, dims[:-1], btree, filters)
        else:
            raise UnsupportedHDF5(f"dataset {name!r}: data layout class {cls}")
        if filters:
            raise UnsupportedHDF5(f"dataset {name!r}: a filter pipeline on a non-chunked dataset")
        return Dataset(name, arr)

    # -- mapping face
    def keys(self):
        return self._datasets.keys()

    def __iter__(self):
        return iter(self._datasets)

    def __contains__(self, name):
        return name in self._datasets

    def __getitem__(self, name) -> Dataset:
        return self._datasets[name]

    def close(self):
        self._f.close()

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()


# ------------------------------------------------------------------------------------------------------- writer
def _pad8(b: bytes) -> bytes:
    return b + b"\x00" * (-len(b) % 8)


def _message(mtype: int, body: bytes, flags: int = 0) -> bytes:
    body = _pad8(body)
    return struct.pack("<HHBBBB", mtype, len(body), flags, 0, 0, 0) + body


def _object_header(messages: List[bytes]) -> bytes:
    data = b"".join(messages)
    return struct.pack("<BBHII", 1, 0, len(messages), 1, len(data)) + b"\x00" * 4 + data


def _chunk_payloads(a: np.ndarray, chunk: Tuple[int, ...], compression: Optional[str], shuffle: bool, level: int):
    """[(origin, stored bytes)] of every chunk of ``a`` (edge chunks padded with zeros to the full chunk shape), filtered shuffle -> deflate."""
    out = []
    for idx in np.ndindex(*[-(-s // c) for s, c in zip(a.shape, chunk)]):
        origin = tuple(i * c for i, c in zip(idx, chunk))
        block = np.zeros(chunk, a.dtype)
        src = a[tuple(slice(o, min(o + c, s)) for o, c, s in zip(origin, chunk, a.shape))]
        block[tuple(slice(0, n) for n in src.shape)] = src
        buf = block.tobytes(order="C")
        if shuffle and a.dtype.itemsize > 1:
            buf = np.frombuffer(buf, np.uint8).reshape(-1, a.dtype.itemsize).T.tobytes()
        if compression == "gzip":
            buf = zlib.compress(buf, level)
        out.append((origin, buf))
    return out


def _chunk_btree(entries, rank: int, elem_size: int, shape, chunk, place, k: int = 32) -> Tuple[int, List[Tuple[int, bytes]]]:
    """Version-1 B-tree (node type 1) over ``entries`` = [(origin, address, stored size)]; ``place(nbytes)`` reserves file space and
    returns its address.  Returns (root address, [(address, node bytes)])."""
    nodes: List[Tuple[int, bytes]] = []
    key_size = 8 + 8 * (rank + 1)
    node_size = 24 + (2 * k) * (key_size + 8) + key_size
    end_key = struct.pack("<II", 0, 0) + b"".join(struct.pack("<Q", -(-s // c) * c) for s, c in zip(shape, chunk)) + struct.pack("<Q", 0)

    def key_of(origin, size):
        return struct.pack("<II", size, 0) + b"".join(struct.pack("<Q", o) for o in origin) + struct.pack("<Q", 0)
    level, items = 0, [(o, addr, size) for o, addr, size in entries]          # (first origin, child address, size for the key)
    while True:
        groups = [items[i:i + 2 * k] for i in range(0, len(items), 2 * k)] or [[]]
        addrs = [place(node_size) for _ in groups]
        nxt = []
        for gi, grp in enumerate(groups):
            body = b"".join(key_of(o, size) + struct.pack("<Q", addr) for o, addr, size in grp)
            last = key_of(groups[gi + 1][0][0], 0) if gi + 1 < len(groups) else end_key
            left = addrs[gi - 1] if gi > 0 else UNDEF
            right = addrs[gi + 1] if gi + 1 < len(groups) else UNDEF
            node = b"TREE" + struct.pack("<BBHQQ", 1, level, len(grp), left, right) + body + last
            nodes.append((addrs[gi], node + b"\x00" * (node_size - len(node))))
            nxt.append((grp[0][0] if grp else tuple([0] * rank), addrs[gi], 0))
        if len(groups) == 1:
            return addrs[0], nodes
        items, level = nxt, level + 1


def write_file(path: str, datasets: Dict[str, np.ndarray], leaf_k: int = 4, internal_k: int = 16, split_headers: bool = False,
               chunks: Optional[Dict[str, Tuple[int, ...]]] = None, compression: Optional[str] = None, shuffle: bool = False,
               compression_level: int = 4) -> None:
    """Writes a flat file (root group with one dataset per entry) in the layout described in the module docstring.
    Up to ``2 * leaf_k`` datasets (one symbol-table node) -- the data files of this project hold two or three.
    ``split_headers`` moves every dataset's layout message into an object-header continuation block, the form headers take in
    files that were modified after creation (exercises that reader path).  Datasets named in ``chunks`` are stored chunked with that
    chunk shape, optionally byte-shuffled and deflated (``compression='gzip'``) -- what ``h5py.create_dataset(..., chunks=...,
    compression='gzip', shuffle=True)`` produces: filter pipeline message, version-1 chunk B-tree, one stored block per chunk."""
    chunks = dict(chunks or {})
    if compression not in (None, "gzip"):
        raise UnsupportedHDF5(f"compression {compression!r} (gzip only)")
    names = sorted(datasets)
    if not 0 < len(names) <= 2 * leaf_k:
        raise UnsupportedHDF5(f"write_file handles 1..{2 * leaf_k} datasets per file")
    arrays = {n: np.asarray(datasets[n], order="C") for n in names}
    for n, a in arrays.items():
        if a.dtype.byteorder == ">":
            arrays[n] = a.astype(a.dtype.newbyteorder("<"))
    # local heap data segment: "" at offset 0, then the names, each NUL-terminated and 8-byte aligned
    heap_data, name_off = bytearray(8), {}
    for n in names:
        name_off[n] = len(heap_data)
        heap_data += _pad8(n.encode() + b"\x00")
    free_off = len(heap_data)
    heap_data += struct.pack("<QQ", 1, 16) + b"\x00" * 16          # one free block: (next = 1 "none", size), then room to grow
    heap_data = bytes(heap_data)

    sb_size, root_hdr_size = 96, 16 + 8 + 16
    btree_size = 24 + (2 * internal_k + 1) * 8 + 2 * internal_k * 8
    snod_size = 8 + 2 * leaf_k * 40
    a_root = sb_size
    a_btree = a_root + root_hdr_size
    a_heap = a_btree + btree_size
    a_heap_data = a_heap + 32
    a_snod = a_heap_data + len(heap_data)
    cursor = a_snod + snod_size
    headers, a_header, a_data = {}, {}, {}
    roots: Dict[str, int] = {}

    def layout_of(n):
        if n in chunks:
            ch = tuple(int(c) for c in chunks[n])
            return _message(MSG_LAYOUT, struct.pack("<BBBQ", 3, 2, len(ch) + 1, roots.get(n, 0)) +
                            b"".join(struct.pack("<I", c) for c in ch) + struct.pack("<I", arrays[n].dtype.itemsize))
        return _message(MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, a_data[n] if arrays[n].nbytes else UNDEF, arrays[n].nbytes))

    def filters_of(n):
        fl = []
        if n in chunks and shuffle and arrays[n].dtype.itemsize > 1:
            fl.append(struct.pack("<HHHH", FILTER_SHUFFLE, 0, 0, 1) + struct.pack("<I", arrays[n].dtype.itemsize) + b"\x00" * 4)
        if n in chunks and compression == "gzip":
            fl.append(struct.pack("<HHHH", FILTER_DEFLATE, 0, 0, 1) + struct.pack("<I", compression_level) + b"\x00" * 4)
        return [_message(0x0B, struct.pack("<BB", 1, len(fl)) + b"\x00" * 6 + b"".join(fl))] if fl else []

    for n in names:                                                   # object headers first, raw data after them
        a = arrays[n]
        dataspace = struct.pack("<BBBBI", 1, a.ndim, 0, 0, 0) + b"".join(struct.pack("<Q", s) for s in a.shape)
        fill = struct.pack("<BBBB", 2, 2, 2, 0)                       # version 2, late allocation, fill never written, undefined
        headers[n] = [_message(MSG_DATASPACE, dataspace), _message(MSG_DATATYPE, _encode_dtype(a.dtype), flags=1), _message(MSG_FILL, fill), None]
        a_header[n] = cursor
        if n in chunks:
            cursor += 16 + sum(len(m) for m in headers[n][:3]) + sum(len(m) for m in filters_of(n)) + len(layout_of(n))
        else:
            cursor += 16 + sum(len(m) for m in headers[n][:3]) + (8 + 16 if split_headers else 8 + 24)
    a_cont = {}
    if split_headers:
        for n in names:
            a_cont[n] = cursor
            cursor += 8 + 24
    chunked: Dict[str, dict] = {}
    for n in names:
        cursor += -cursor % 8
        if n in chunks:
            if split_headers:
                raise UnsupportedHDF5("split_headers with chunked datasets")
            a, ch = arrays[n], tuple(int(c) for c in chunks[n])
            if len(ch) != a.ndim or any(c < 1 for c in ch):
                raise UnsupportedHDF5(f"chunk shape {ch} does not fit dataset {n!r} of rank {a.ndim}")
            blobs = []

            def place(nbytes, _c=[cursor]):
                _c[0] += -_c[0] % 8
                at = _c[0]
                _c[0] += nbytes
                return at
            entries = []
            for origin, buf in _chunk_payloads(a, ch, compression, shuffle, compression_level):
                at = place(len(buf))
                blobs.append((at, buf))
                entries.append((origin, at, len(buf)))
            root, nodes = _chunk_btree(entries, a.ndim, a.dtype.itemsize, a.shape, ch, place)
            cursor = place(0)
            chunked[n] = {"chunk": ch, "root": root, "writes": blobs + nodes}
            roots[n] = root
            a_data[n] = root
        else:
            a_data[n] = cursor
            cursor += arrays[n].nbytes
    eof = cursor

    with open(path, "wb") as f:
        f.write(SIGNATURE + struct.pack("<BBBBBBBB", 0, 0, 0, 0, 0, 8, 8, 0) + struct.pack("<HHI", leaf_k, internal_k, 0))
        f.write(struct.pack("<QQQQ", 0, UNDEF, eof, UNDEF))
        f.write(struct.pack("<QQII", 0, a_root, 1, 0) + struct.pack("<QQ", a_btree, a_heap))      # root entry caches B-tree / heap
        f.write(_object_header([_message(MSG_SYMBOL_TABLE, struct.pack("<QQ", a_btree, a_heap))]))
        node = b"TREE" + struct.pack("<BBHQQ", 0, 0, 1, UNDEF, UNDEF) + struct.pack("<QQQ", 0, a_snod, name_off[names[-1]])
        f.write(node + b"\x00" * (btree_size - len(node)))
        f.write(b"HEAP" + struct.pack("<BBBBQQQ", 0, 0, 0, 0, len(heap_data), free_off, a_heap_data))
        f.write(heap_data)
        snod = b"SNOD" + struct.pack("<BBH", 1, 0, len(names))
        for n in names:
            snod += struct.pack("<QQII", name_off[n], a_header[n], 0, 0) + b"\x00" * 16
        f.write(snod + b"\x00" * (snod_size - len(snod)))
        for n in names:
            assert f.tell() == a_header[n]
            layout = layout_of(n)
            if n in chunks:
                f.write(_object_header(headers[n][:3] + filters_of(n) + [layout]))
            elif split_headers:
                first = headers[n][:3] + [_message(MSG_CONTINUATION, struct.pack("<QQ", a_cont[n], len(layout)))]
                data = b"".join(first)
                f.write(struct.pack("<BBHII", 1, 0, len(first) + 1, 1, len(data)) + b"\x00" * 4 + data)     # message count spans both blocks
            else:
                f.write(_object_header(headers[n][:3] + [layout]))
        if split_headers:
            for n in names:
                assert f.tell() == a_cont[n]
                f.write(_message(MSG_LAYOUT, struct.pack("<BBQQ", 3, 1, a_data[n] if arrays[n].nbytes else UNDEF, arrays[n].nbytes)))
        for n in names:
            if n in chunked:
                for at, blob in sorted(chunked[n]["writes"]):
                    f.write(b"\x00" * (at - f.tell()))
                    f.write(blob)
                continue
            f.write(b"\x00" * (a_data[n] - f.tell()))
            f.write(arrays[n].tobytes(order="C"))
        f.write(b"\x00" * (eof - f.tell()))
        assert f.tell() == eof
