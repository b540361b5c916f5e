"""Host -> device batch staging.

The reference builds every batch with a blocking ``torch.tensor(numpy, device="cuda")`` of fp32
volumes (``sailentgrads/my_model_trainer.py:185-199``: 8.5 MB/sample, pageable memory), again for
every epoch, and stages each client's test split twice per round (global, then personalised model:
``sailentgrads_api.py:247-256``).  Here volumes stay 8-bit (2.1 MB/sample) and

* ``stream``   mode copies batch by batch out of the store's pinned cache on a dedicated copy stream into a
               ring of device buffers (the copy of batch i+1.. overlaps the compute of batch i);
* ``client``   mode (default for index loaders) uploads the *set* of volumes a loader draws from once into a
               device slot -- every epoch's shuffled batches, and both evaluation passes, are then assembled
               on the device by one gather launch.  Uploads are queued ahead (:meth:`BatchStager.want`) so the
               next clients' data and the round's test splits cross PCIe while the current client trains;
* ``resident`` mode (:class:`ResidentStager`) keeps the whole store on the device: no copies at all inside a
               round (datasets that fit in HBM: 21 x 40 ABCD volumes are 1.8 GB of 180 GB).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Iterator, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..data.volumes import VolumeStore

_GATHER_ARG_ROWS = 64       # nidt_gather_rows_u8_args passes at most this many row numbers by value


class _Slot:
    """One device buffer of ``cap`` volumes + labels, filled by the copy stream, read by the compute stream."""

    def __init__(self, cap: int, shape, device, need_pin_x: bool):
        self.cap = cap
        self.x = torch.empty((cap,) + tuple(shape), dtype=torch.uint8, device=device)
        self.y = torch.empty(cap, dtype=torch.float32, device=device)
        self.pin_x = torch.empty((cap,) + tuple(shape), dtype=torch.uint8).pin_memory() if need_pin_x else None
        self.ready = torch.cuda.Event()
        self.consumed = torch.cuda.Event()
        self.busy = False
        self.packed = self.tab = self.off = None      # zero-run-skipping uploads: device copies of the spans / row tables


class SparsePack:
    """Zero-run-skipping transfer form of a pinned 8-bit volume cache ``[N, D, H, W]`` (W <= 255).

    Masked grey-matter maps are zero outside the brain (``Preprocess_ABCD.ipynb`` applies a group-mean mask), so an image row is
    ``[zeros | span | zeros]``.  Built once per store: the non-zero spans of every volume packed back to back (pinned), and per
    row a (start, length) byte pair and the span's offset inside its volume (uint32).  An upload then moves the spans and 6 bytes per
    row; ``nidt_expand_sparse_rows_u8`` rebuilds the dense volume on the device.  Lossless."""

    def __init__(self, cache: torch.Tensor, block: int = 32, native: Optional[bool] = None):
        n, d, h, w = cache.shape
        assert w <= 255
        self.rows, self.w = d * h, w
        self.dense_bytes = n * self.rows * w
        pin = torch.cuda.is_available()
        lib = self._native_lib() if native in (None, True) else None
        if native is True and lib is None:
            raise RuntimeError("SparsePack: the native scan is not available")
        if lib is not None and cache.is_contiguous() and cache.dtype == torch.uint8:
            self._build_native(lib, cache, n, pin)
        else:
            self._build_numpy(cache, n, block, pin)
        self.ratio = (int(self.vol_off[-1]) + self.tab.numel() + 4 * self.off.numel()) / max(1, self.dense_bytes)

    @staticmethod
    def _native_lib():
        try:
            from ..ops import native as _n
            lib = _n.load()
            i32, i64, vp = C.c_int, C.c_int64, C.c_void_p
            lib.nidt_sparse_scan_rows.argtypes = [vp, i64, i32, i32, vp, vp, vp, i32]
            lib.nidt_sparse_scan_rows.restype = i32
            lib.nidt_sparse_copy_rows.argtypes = [vp, i64, i32, i32, vp, vp, vp, vp, i32]
            lib.nidt_sparse_copy_rows.restype = i32
            return lib
        except Exception:
            return None

    def _build_native(self, lib, cache: torch.Tensor, n: int, pin: bool) -> None:
        """Span table + packed spans by the library's host code (a few threads; seconds -> tenths of a second per GB)."""
        from ..ops import native as _n
        threads = max(1, min(16, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))
        self.tab = torch.empty((n, self.rows, 2), dtype=torch.uint8)
        self.off = torch.empty((n, self.rows), dtype=torch.int32)                 # bit pattern of the uint32 offsets
        vol_bytes = np.zeros(n, np.int64)
        _n.check(lib.nidt_sparse_scan_rows(cache.data_ptr(), n, self.rows, self.w, self.tab.data_ptr(), self.off.data_ptr(), vol_bytes.ctypes.data,
                                           threads), "nidt_sparse_scan_rows")
        self.vol_off = np.concatenate([[0], np.cumsum(vol_bytes)]).astype(np.int64)
        self.max_vol_bytes = int(vol_bytes.max()) if n else 0
        self.packed = torch.empty(int(self.vol_off[-1]), dtype=torch.uint8)
        if pin:
            self.packed, self.tab, self.off = self.packed.pin_memory(), self.tab.pin_memory(), self.off.pin_memory()
        _n.check(lib.nidt_sparse_copy_rows(cache.data_ptr(), n, self.rows, self.w, self.tab.data_ptr(), self.off.data_ptr(), self.vol_off.ctypes.data,
                                           self.packed.data_ptr(), threads), "nidt_sparse_copy_rows")

    def _build_numpy(self, cache: torch.Tensor, n: int, block: int, pin: bool) -> None:
        w = self.w
        a = cache.numpy().reshape(n, self.rows, w)
        cols = np.arange(w)
        tab = np.zeros((n, self.rows, 2), np.uint8)
        off = np.zeros((n, self.rows), np.uint32)
        vol_bytes = np.zeros(n, np.int64)
        chunks = []
        for i0 in range(0, n, block):
            blk = a[i0:i0 + block]
            nz = blk != 0
            has = nz.any(-1)
            start = nz.argmax(-1)
            last = w - 1 - nz[..., ::-1].argmax(-1)
            ln = np.where(has, last - start + 1, 0)
            tab[i0:i0 + block, :, 0] = np.where(has, start, 0)
            tab[i0:i0 + block, :, 1] = ln
            cs = np.cumsum(ln, axis=1)
            off[i0:i0 + block] = cs - ln
            vol_bytes[i0:i0 + block] = cs[:, -1]
            inside = (cols >= np.where(has, start, 0)[..., None]) & (cols < (np.where(has, start, 0) + ln)[..., None])
            chunks.append(blk[inside])                     # row-major: volume by volume, row by row
        self.vol_off = np.concatenate([[0], np.cumsum(vol_bytes)]).astype(np.int64)
        self.max_vol_bytes = int(vol_bytes.max()) if n else 0
        self.packed = torch.from_numpy(np.concatenate(chunks) if chunks else np.zeros(0, np.uint8))
        self.tab = torch.from_numpy(tab)
        self.off = torch.from_numpy(off.view(np.int32))    # bit pattern of the uint32 offsets
        if pin:
            self.packed, self.tab, self.off = self.packed.pin_memory(), self.tab.pin_memory(), self.off.pin_memory()


class _Entry:
    """The volumes of one loader (or of a group of loaders) on their way to / resident in a slot."""

    def __init__(self, loaders: Sequence, kind: str):
        self.loaders = tuple(loaders)
        self.kind = kind
        self.ids: Dict[int, np.ndarray] = {}
        self.base: Dict[int, int] = {}
        rows = []
        off = 0
        for l in self.loaders:
            u = np.unique(np.asarray(l.indices).astype(np.int64))
            self.ids[id(l)] = u
            self.base[id(l)] = off
            rows.append(u)
            off += len(u)
        self.rows = np.ascontiguousarray(np.concatenate(rows) if rows else np.zeros(0, np.int64))
        self.slot: Optional[_Slot] = None
        self.opened = False

    def holds(self, loader) -> bool:
        return id(loader) in self.base


class BatchStager:
    def __init__(self, store: VolumeStore, device: torch.device, batch_size: int, depth: int = 4, sort_indices: bool = True,
                 train_slots: int = 3, eval_slots: int = 2, max_cache_bytes: int = 8 << 30):
        self.store = store
        self.device = torch.device(device)
        self.batch_size = batch_size
        self.depth = depth
        self.sort_indices = sort_indices
        self.cuda = self.device.type == "cuda"
        self.h2d_bytes = 0
        self.h2d_copies = 0
        self.gather_launches = 0
        self.last_ids = None              # store ids of the rows of the batch handed out last, in row order (engines key per-sample caches on them)
        self._slot_ids = {}
        self.fixed_output = False         # True: every batch is gathered into ONE buffer (stable addresses for CUDA-graph replays)
        self._row_bytes = int(np.prod(store.shape))
        self.max_cache_rows = max(1, max_cache_bytes // self._row_bytes)
        self._pool_depth = {"train": train_slots, "eval": eval_slots}
        self._pools: Dict[str, List[_Slot]] = {"train": [], "eval": []}
        self._queue: List[_Entry] = []          # wanted, in upload order; entries leave when they are closed
        self._ring_ready = False
        self._native = None
        self._pinned_cache = None
        self._sparse: Optional[SparsePack] = None
        if self.cuda:
            self.copy_stream = torch.cuda.Stream(device=self.device)
            cache = getattr(store, "_cache", None)
            if cache is None:
                cache = getattr(store, "volumes", None)
            self._pinned_cache = cache if (cache is not None and cache.is_pinned() and cache.is_contiguous()) else None
            self._kernel_blocks = int(os.environ.get("NIDT_STAGER_BLOCKS", "0"))      # > 0: stage with a copy kernel instead of the copy engine
            from ..ops import native
            lib = native.load()
            i32, i64, vp = C.c_int, C.c_int64, C.c_void_p
            lib.nidt_copy_rows_h2d_counted.argtypes = [vp, vp, i32, i64, vp, vp, C.POINTER(i32)]
            lib.nidt_copy_rows_h2d_counted.restype = i32
            lib.nidt_gather_rows_u8_capped.argtypes = [vp, vp, vp, i32, i64, i32, vp]
            lib.nidt_gather_rows_u8_capped.restype = i32
            lib.nidt_gather_rows_u8_args.argtypes = [vp, vp, vp, i32, i64, vp, vp, vp]
            lib.nidt_gather_rows_u8_args.restype = i32
            self._native = lib
            # labels are copied straight out of ONE pinned array of all labels (row copies, like the volumes): no per-slot pinned
            # bounce buffer that the host could overwrite while an earlier, still queued copy has yet to read it
            self._pin_labels = torch.from_numpy(np.ascontiguousarray(store.labels, dtype=np.float32)).pin_memory()
            lib.nidt_expand_sparse_rows_u8.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, vp]
            lib.nidt_expand_sparse_rows_u8.restype = i32
            # zero-run-skipping uploads for the per-loader device cache when the volumes are sparse enough (NIDT_SPARSE_H2D=0: dense)
            if (self._pinned_cache is not None and self._pinned_cache.dim() == 4 and self._pinned_cache.shape[-1] <= 255
                    and os.environ.get("NIDT_SPARSE_H2D", "1") != "0"):
                pack = getattr(store, "_sparse_pack", None)
                if pack is None:
                    pack = SparsePack(self._pinned_cache)
                    store._sparse_pack = pack                          # shared by every stager of this store
                self._sparse = pack if pack.ratio < 0.8 else None
            self._gx = [torch.empty((batch_size,) + tuple(store.shape), dtype=torch.uint8, device=self.device) for _ in range(2)]
            self._gy = [torch.empty(batch_size, dtype=torch.float32, device=self.device) for _ in range(2)]
            self._gk = 0

    # ------------------------------------------------------------------ copies
    def _copy_rows(self, ids: np.ndarray, dst_x: torch.Tensor, pin_x: Optional[torch.Tensor]) -> None:
        """Enqueue ``dst_x[j] = volume ids[j]`` on the copy stream (the caller holds the stream context)."""
        from ..ops import native
        n = len(ids)
        if n == 0:
            return
        if self._pinned_cache is not None:
            cnt = C.c_int(0)
            rows = np.ascontiguousarray(ids, dtype=np.int64)
            native.check(self._native.nidt_copy_rows_h2d_counted(self._pinned_cache.data_ptr(), rows.ctypes.data, n, self._row_bytes,
                                                                 dst_x.data_ptr(), self.copy_stream.cuda_stream, C.byref(cnt)), "nidt_copy_rows_h2d")
            self.h2d_copies += cnt.value
        else:
            self.store.fetch(ids, out=pin_x)
            dst_x[:n].copy_(pin_x[:n], non_blocking=True)
            self.h2d_copies += 1
        self.h2d_bytes += n * self._row_bytes

    def _copy_rows_sparse(self, ids: np.ndarray, slot: _Slot) -> None:
        """``slot.x[j] = volume ids[j]`` through the zero-run-skipping form: the packed spans and row tables of runs of adjacent
        volumes are copied (three copies per run), then expanded on the copy stream (the caller holds the stream context)."""
        from ..ops import native
        sp = self._sparse
        n = len(ids)
        if n == 0:
            return
        if getattr(slot, "packed", None) is None or slot.packed.numel() < slot.cap * sp.max_vol_bytes:
            slot.packed = torch.empty(max(1, slot.cap * sp.max_vol_bytes), dtype=torch.uint8, device=self.device)
            slot.tab = torch.empty((slot.cap, sp.rows, 2), dtype=torch.uint8, device=self.device)
            slot.off = torch.empty((slot.cap, sp.rows), dtype=torch.int32, device=self.device)
        bases = np.zeros(n, np.int64)
        j, poff = 0, 0
        while j < n:
            e = j + 1
            while e < n and ids[e] == ids[e - 1] + 1:
                e += 1
            a, b = int(ids[j]), int(ids[e - 1]) + 1
            lo, hi = int(sp.vol_off[a]), int(sp.vol_off[b])
            if hi > lo:
                slot.packed[poff:poff + hi - lo].copy_(sp.packed[lo:hi], non_blocking=True)
            slot.tab[j:e].copy_(sp.tab[a:b], non_blocking=True)
            slot.off[j:e].copy_(sp.off[a:b], non_blocking=True)
            bases[j:e] = poff + (sp.vol_off[a:b] - lo)
            self.h2d_bytes += (hi - lo) + (b - a) * sp.rows * 6
            self.h2d_copies += 3
            poff += hi - lo
            j = e
        st = self.copy_stream.cuda_stream
        for k in range(0, n, 32):
            m = min(32, n - k)
            part = np.ascontiguousarray(bases[k:k + m])
            native.check(self._native.nidt_expand_sparse_rows_u8(slot.packed.data_ptr(), part.ctypes.data, slot.tab[k:].data_ptr(), slot.off[k:].data_ptr(),
                                                                 slot.x[k:].data_ptr(), m, sp.rows, sp.w, st), "nidt_expand_sparse_rows_u8")
            native.count_launches(1)

    def _copy_labels(self, ids: np.ndarray, dst_y: torch.Tensor) -> None:
        """Enqueue ``dst_y[j] = label ids[j]`` on the copy stream (the caller holds the stream context)."""
        from ..ops import native
        n = len(ids)
        if n == 0:
            return
        rows = np.ascontiguousarray(ids, dtype=np.int64)
        cnt = C.c_int(0)
        native.check(self._native.nidt_copy_rows_h2d_counted(self._pin_labels.data_ptr(), rows.ctypes.data, n, 4, dst_y.data_ptr(),
                                                             self.copy_stream.cuda_stream, C.byref(cnt)), "nidt_copy_rows_h2d(labels)")
        self.h2d_bytes += 4 * n
        self.h2d_copies += cnt.value

    # ------------------------------------------------------------------ stream mode (ring of batch buffers)
    def _ring_init(self) -> None:
        shp = (self.batch_size,) + tuple(self.store.shape)
        d = self.depth
        self.dev_x = [torch.empty(shp, dtype=torch.uint8, device=self.device) for _ in range(d)]
        self.dev_y = [torch.empty(self.batch_size, dtype=torch.float32, device=self.device) for _ in range(d)]
        self.pin_x = [torch.empty(shp, dtype=torch.uint8).pin_memory() for _ in range(d)] if self._pinned_cache is None else [None] * d
        self.pin_idx = [torch.empty(self.batch_size, dtype=torch.int64).pin_memory() for _ in range(d)]
        self.ready = [torch.cuda.Event() for _ in range(d)]
        self.consumed = [torch.cuda.Event() for _ in range(d)]
        self._ring_ready = True

    def _issue(self, slot: int, idx: torch.Tensor) -> int:
        """Enqueue the H2D copies of one batch on the copy stream; returns the batch size."""
        from ..ops import native
        ids = idx.numpy().astype(np.int64)
        if self.sort_indices:
            ids = np.sort(ids)
        b = len(ids)
        self._slot_ids[slot] = ids
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])
            if self._pinned_cache is not None and self._kernel_blocks > 0:
                # a few CTAs read the pinned host rows over PCIe themselves (no copy-engine traffic next to the compute kernels)
                self.ready[slot].synchronize()             # the pinned index buffer is about to be overwritten by the host
                self.pin_idx[slot][:b] = torch.from_numpy(np.ascontiguousarray(ids, dtype=np.int64))
                native.check(self._native.nidt_gather_rows_u8_capped(self._pinned_cache.data_ptr(), self.pin_idx[slot].data_ptr(), self.dev_x[slot].data_ptr(),
                                                                     b, self._row_bytes, self._kernel_blocks, self.copy_stream.cuda_stream),
                             "nidt_gather_rows_u8_capped")
                self.h2d_bytes += b * self._row_bytes
                self.h2d_copies += 1
            else:
                if self.pin_x[slot] is not None:
                    self.ready[slot].synchronize()         # the pinned bounce buffer is about to be overwritten by the host
                self._copy_rows(ids, self.dev_x[slot], self.pin_x[slot])
            self._copy_labels(ids, self.dev_y[slot])
            self.ready[slot].record(self.copy_stream)
        return b

    def iterate(self, loader, lookahead=None) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        """Yields ``(x_u8 [b,D,H,W], y [b])`` on the device; the tensors of one iteration stay valid
        until ``depth - 1`` further iterations have been drawn.

        ``lookahead``: the index stream the NEXT call will iterate (the next client's batches).  When this stream
        runs dry its first batch is copied right away, so the next call starts with its data already on the way
        instead of exposing one host->device copy per client."""
        if not self.cuda:
            for batch in loader:
                ids = np.asarray(batch[0]).astype(np.int64)
                if self.sort_indices:
                    ids = np.sort(ids)
                x, y = self.store.fetch(ids)
                self.last_ids = ids
                yield x, y
            return
        if not self._ring_ready:
            self._ring_init()
        carry, self._carry = getattr(self, "_carry", None), None
        cur = torch.cuda.current_stream(self.device)
        if carry is not None and carry[0] is loader:
            _, it, pending, slot = carry
        else:
            it = iter(loader)
            pending = []
            slot = 0
            for _ in range(self.depth - 1):
                try:
                    b = next(it)
                except StopIteration:
                    break
                pending.append((slot, self._issue(slot, b[0])))
                slot = (slot + 1) % self.depth
        dry = False
        while pending:
            if not dry:
                try:
                    b = next(it)
                    pending.append((slot, self._issue(slot, b[0])))
                    slot = (slot + 1) % self.depth
                except StopIteration:
                    dry = True
                    if lookahead is not None:
                        la_it = iter(lookahead)
                        try:
                            b = next(la_it)
                            self._carry = (lookahead, la_it, [(slot, self._issue(slot, b[0]))], (slot + 1) % self.depth)
                        except StopIteration:
                            pass
            s, n = pending.pop(0)
            cur.wait_event(self.ready[s])
            self.last_ids = self._slot_ids.get(s)
            yield self.dev_x[s][:n], self.dev_y[s][:n]
            self.consumed[s].record(cur)

    # ------------------------------------------------------------------ client mode (device slot per loader / loader group)
    def cacheable(self, loader) -> bool:
        return self.cuda and hasattr(loader, "indices") and 0 < len(loader.indices) <= self.max_cache_rows

    def want(self, loaders: Sequence, kind: str = "train", group: bool = False) -> None:
        """Declare loaders that will be iterated soon, in order.  Their volumes are uploaded as slots become free
        (``train_slots`` / ``eval_slots`` deep), behind whatever the compute stream is doing.  ``group=True`` uploads
        all of them into ONE slot (the round's test splits, evaluated by several models)."""
        if not self.cuda:
            return
        loaders = [l for l in loaders if self.cacheable(l)]
        if not loaders:
            return
        if group:
            if not any(e.loaders == tuple(loaders) and not e.opened for e in self._queue):
                ent = _Entry(loaders, kind)
                if len(ent.rows) <= self.max_cache_rows:
                    self._queue.append(ent)
        else:
            for l in loaders:
                if not any(len(e.loaders) == 1 and e.holds(l) and not e.opened for e in self._queue):
                    self._queue.append(_Entry([l], kind))
        self._pump()

    def _free_slot(self, kind: str, rows: int, force: bool) -> Optional[_Slot]:
        pool = self._pools[kind]
        for s in pool:
            if not s.busy and s.cap >= rows:
                return s
        for i, s in enumerate(pool):
            if not s.busy:                                   # free but too small: replace it once the GPU is done with it
                s.consumed.synchronize()
                pool[i] = _Slot(rows, self.store.shape, self.device, self._pinned_cache is None)
                return pool[i]
        if len(pool) < self._pool_depth[kind]:
            cap = max([rows] + [len(e.rows) for e in self._queue if e.kind == kind])
            pool.append(_Slot(cap, self.store.shape, self.device, self._pinned_cache is None))
            return pool[-1]
        if force:
            # every slot is held by an upload nobody has opened (a mispredicted prefetch): evict the oldest and take its slot
            for e in self._queue:
                if e.kind == kind and e.slot is not None and not e.opened:
                    slot, e.slot = e.slot, None
                    self._queue.remove(e)
                    slot.busy = False
                    if slot.cap < rows:
                        slot.ready.synchronize()
                        pool[pool.index(slot)] = slot = _Slot(rows, self.store.shape, self.device, self._pinned_cache is None)
                    return slot
            pool.append(_Slot(rows, self.store.shape, self.device, self._pinned_cache is None))
            return pool[-1]
        return None

    def _pump(self, force_entry: Optional[_Entry] = None) -> None:
        """Issue queued uploads, oldest first, while slots are free."""
        for ent in list(self._queue):
            if ent.slot is not None or ent not in self._queue:
                continue
            slot = self._free_slot(ent.kind, len(ent.rows), force=ent is force_entry)
            if slot is None:
                if force_entry is None or ent is force_entry:
                    break
                continue
            slot.busy = True
            ent.slot = slot
            n = len(ent.rows)
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(slot.consumed)
                if slot.pin_x is not None:
                    slot.ready.synchronize()
                if self._sparse is not None:
                    self._copy_rows_sparse(ent.rows, slot)
                else:
                    self._copy_rows(ent.rows, slot.x, slot.pin_x)
                self._copy_labels(ent.rows, slot.y)
                slot.ready.record(self.copy_stream)

    def open(self, loaders, kind: str = "train") -> _Entry:
        """The entry holding ``loaders`` (one loader, or the exact group passed to ``want(group=True)``), uploaded or on its
        way; the current stream waits for the upload.  Pair with :meth:`close`."""
        group = isinstance(loaders, (list, tuple))
        key = tuple(loaders) if group else (loaders,)
        ent = next((e for e in self._queue if e.loaders == key and not e.opened), None)
        if ent is None:
            ent = _Entry(key, kind)
            self._queue.append(ent)
        if ent.slot is None:
            self._pump(force_entry=ent)
        ent.opened = True
        torch.cuda.current_stream(self.device).wait_event(ent.slot.ready)
        return ent

    def close(self, ent: _Entry) -> None:
        ent.slot.consumed.record(torch.cuda.current_stream(self.device))
        ent.slot.busy = False
        self._queue.remove(ent)
        self._pump()

    def _positions(self, ent: _Entry, loader, ids: np.ndarray) -> np.ndarray:
        return ent.base[id(loader)] + np.searchsorted(ent.ids[id(loader)], ids)

    def batch_from(self, ent: _Entry, loader, batch_idx) -> Tuple[torch.Tensor, torch.Tensor]:
        """One index batch of ``loader`` assembled from its slot: a view when the (sorted) rows are adjacent in the slot,
        otherwise one gather launch (volumes + labels, row numbers passed in the kernel parameters)."""
        from ..ops import native
        ids = np.asarray(batch_idx).astype(np.int64)
        if self.sort_indices:
            ids = np.sort(ids)
        pos = self._positions(ent, loader, ids)
        n = len(pos)
        slot = ent.slot
        self.last_ids = ids
        if not self.fixed_output and n > 0 and int(pos[-1]) - int(pos[0]) == n - 1 and (n == 1 or bool(np.all(np.diff(pos) == 1))):
            p0 = int(pos[0])
            return slot.x[p0:p0 + n], slot.y[p0:p0 + n]
        k = self._gk = 0 if self.fixed_output else (self._gk + 1) % len(self._gx)
        if n > self._gx[k].shape[0]:
            self._gx = [torch.empty((n,) + tuple(self.store.shape), dtype=torch.uint8, device=self.device) for _ in self._gx]
            self._gy = [torch.empty(n, dtype=torch.float32, device=self.device) for _ in self._gy]
        gx, gy = self._gx[k][:n], self._gy[k][:n]
        st = torch.cuda.current_stream(self.device).cuda_stream
        for o in range(0, n, _GATHER_ARG_ROWS):
            m = min(_GATHER_ARG_ROWS, n - o)
            part = np.ascontiguousarray(pos[o:o + m], dtype=np.int64)
            native.check(self._native.nidt_gather_rows_u8_args(slot.x.data_ptr(), part.ctypes.data, gx[o:o + m].data_ptr(), m, self._row_bytes,
                                                               slot.y.data_ptr(), gy[o:o + m].data_ptr(), st), "nidt_gather_rows_u8_args")
            native.count_launches(1)
            self.gather_launches += 1
        return gx, gy

    def cached_batches(self, loader, epochs: int = 1) -> Iterator[Tuple[int, torch.Tensor, torch.Tensor]]:
        """``(epoch, x, y)`` over ``epochs`` passes of ``loader`` with its volumes uploaded once."""
        ent = self.open(loader)
        try:
            for e in range(epochs):
                for b in loader:
                    x, y = self.batch_from(ent, loader, b[0])
                    yield e, x, y
        finally:
            self.close(ent)


class ResidentStager(BatchStager):
    """The whole store lives on the device (uploaded once, at construction): batches are views / gathers of it and
    nothing crosses PCIe afterwards.  For datasets that fit in HBM next to the activations."""

    def __init__(self, store: VolumeStore, device: torch.device, batch_size: int, **kw):
        super().__init__(store, device, batch_size, **kw)
        if not self.cuda:
            return
        n = len(store.labels)
        cache = getattr(store, "_cache", None)
        if cache is None:
            cache = getattr(store, "volumes", None)
        ent = _Entry([], "resident")
        slot = _Slot.__new__(_Slot)
        slot.cap = n
        slot.x = cache.to(self.device) if cache is not None else store.fetch(np.arange(n))[0].to(self.device)
        slot.y = torch.from_numpy(np.asarray(store.labels, dtype=np.float32)).to(self.device)
        slot.ready, slot.consumed, slot.busy = torch.cuda.Event(), torch.cuda.Event(), True
        slot.ready.record(torch.cuda.current_stream(self.device))
        ent.slot = slot
        self._all = ent

    def cacheable(self, loader) -> bool:
        return self.cuda and hasattr(loader, "indices")

    def want(self, loaders, kind: str = "train", group: bool = False) -> None:
        return None

    def _positions(self, ent: _Entry, loader, ids: np.ndarray) -> np.ndarray:
        return ids

    def open(self, loaders, kind: str = "train") -> _Entry:
        return self._all

    def close(self, ent: _Entry) -> None:
        return None

    def iterate(self, loader, lookahead=None):
        if not self.cuda:
            yield from super().iterate(loader, lookahead)
            return
        for b in loader:
            yield self.batch_from(self._all, None, b[0])


def make_stager(mode: str, store: VolumeStore, device, batch_size: int) -> BatchStager:
    """``stream`` | ``client`` (same class; the trainer decides per loader) | ``resident``."""
    if mode == "resident":
        return ResidentStager(store, device, batch_size)
    if mode not in ("stream", "client", "auto"):
        raise ValueError(f"unknown staging mode {mode!r} (stream | client | resident)")
    return BatchStager(store, device, batch_size)
