"""Host -> device batch staging.

The reference builds every batch with a blocking ``torch.tensor(numpy, device="cuda")`` of fp32
volumes (``sailentgrads/my_model_trainer.py:185-199``: 8.5 MB/sample, pageable memory).  Here batches
are 8-bit, copied sample-by-sample straight out of the store's pinned cache on a dedicated copy
stream into a ring of device buffers, so the copy of batch i+1 overlaps the compute of batch i.
"""
from __future__ import annotations

from typing import Iterator, Tuple

import os

import numpy as np
import torch

from ..data.volumes import VolumeStore


class BatchStager:
    def __init__(self, store: VolumeStore, device: torch.device, batch_size: int, depth: int = 2, sort_indices: bool = True):
        self.store = store
        self.device = torch.device(device)
        self.batch_size = batch_size
        self.depth = depth
        self.sort_indices = sort_indices
        self.cuda = self.device.type == "cuda"
        self.h2d_bytes = 0
        if self.cuda:
            self.copy_stream = torch.cuda.Stream(device=self.device)
            shp = (batch_size,) + tuple(store.shape)
            self.dev_x = [torch.empty(shp, dtype=torch.uint8, device=self.device) for _ in range(depth)]
            self.dev_y = [torch.empty(batch_size, dtype=torch.float32, device=self.device) for _ in range(depth)]
            self.pin_y = [torch.empty(batch_size, dtype=torch.float32).pin_memory() for _ in range(depth)]
            self.ready = [torch.cuda.Event() for _ in range(depth)]
            self.consumed = [torch.cuda.Event() for _ in range(depth)]
            cache = getattr(store, "_cache", None)
            if cache is None:
                cache = getattr(store, "volumes", None)
            self._pinned_cache = cache if (cache is not None and cache.is_pinned()) else None
            if self._pinned_cache is None:
                self.pin_x = [torch.empty(shp, dtype=torch.uint8).pin_memory() for _ in range(depth)]
            self._native = None
            self._kernel_blocks = int(os.environ.get("NIDT_STAGER_BLOCKS", "0"))      # > 0: stage with a copy kernel instead of the copy engine
            self.pin_idx = [torch.empty(batch_size, dtype=torch.int64).pin_memory() for _ in range(depth)]
            self._row_bytes = int(np.prod(store.shape))
            if self._pinned_cache is not None and self._pinned_cache.is_contiguous():
                try:
                    import ctypes as C
                    from ..ops import native
                    lib = native.load()
                    lib.nidt_copy_rows_h2d.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_void_p, C.c_void_p]
                    lib.nidt_copy_rows_h2d.restype = C.c_int
                    lib.nidt_gather_rows_u8_capped.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int64, C.c_int, C.c_void_p]
                    lib.nidt_gather_rows_u8_capped.restype = C.c_int
                    self._native = lib
                except Exception:
                    self._native = None

    def _issue(self, slot: int, idx: torch.Tensor) -> int:
        """Enqueue the H2D copies of one batch on the copy stream; returns the batch size."""
        ids = idx.numpy().astype(np.int64)
        if self.sort_indices:
            ids = np.sort(ids)
        b = len(ids)
        with torch.cuda.stream(self.copy_stream):
            self.copy_stream.wait_event(self.consumed[slot])
            if self._pinned_cache is not None:
                if self._native is not None and self._kernel_blocks > 0:
                    # a few CTAs read the pinned host rows over PCIe themselves (no copy-engine traffic next to the compute kernels)
                    self.pin_idx[slot][:b] = torch.from_numpy(np.ascontiguousarray(ids, dtype=np.int64))
                    rc = self._native.nidt_gather_rows_u8_capped(self._pinned_cache.data_ptr(), self.pin_idx[slot].data_ptr(), self.dev_x[slot].data_ptr(),
                                                                 b, self._row_bytes, self._kernel_blocks, self.copy_stream.cuda_stream)
                    if rc != 0:
                        raise RuntimeError("nidt_gather_rows_u8_capped failed")
                elif self._native is not None:      # one native call issues the per-sample copies (no per-copy framework overhead)
                    rows = np.ascontiguousarray(ids, dtype=np.int64)
                    rc = self._native.nidt_copy_rows_h2d(self._pinned_cache.data_ptr(), rows.ctypes.data, b, self._row_bytes,
                                                         self.dev_x[slot].data_ptr(), self.copy_stream.cuda_stream)
                    if rc != 0:
                        raise RuntimeError("nidt_copy_rows_h2d failed")
                else:
                    for j, i in enumerate(ids):
                        self.dev_x[slot][j].copy_(self._pinned_cache[int(i)], non_blocking=True)
            else:
                self.store.fetch(ids, out=self.pin_x[slot])
                self.dev_x[slot][:b].copy_(self.pin_x[slot][:b], non_blocking=True)
            self.pin_y[slot][:b] = torch.from_numpy(self.store.labels[ids])
            self.dev_y[slot][:b].copy_(self.pin_y[slot][:b], non_blocking=True)
            self.ready[slot].record(self.copy_stream)
        self.h2d_bytes += b * int(np.prod(self.store.shape)) + 4 * b
        return b

    def iterate(self, loader, lookahead=None) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        """Yields ``(x_u8 [b,D,H,W], y [b])`` on the device; the tensors of one iteration stay valid
        until the next-but-one iteration (ring depth 2).

        ``lookahead``: the index stream the NEXT call will iterate (the next client's batches).  When this stream
        runs dry its first batch is copied right away, so the next call starts with its data already on the way
        instead of exposing one host->device copy per client."""
        if not self.cuda:
            for batch in loader:
                ids = np.asarray(batch[0]).astype(np.int64)
                if self.sort_indices:
                    ids = np.sort(ids)
                x, y = self.store.fetch(ids)
                yield x, y
            return
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
            yield self.dev_x[s][:n], self.dev_y[s][:n]
            self.consumed[s].record(cur)
