"""Flat, device-resident parameter slabs.

The reference moves models around as CPU ``state_dict`` deep copies
(``fedml_api/standalone/sailentgrads/my_model_trainer.py:132-136``,
``sailentgrads_api.py:131-136``).  Here every model instance (global model, one per
client, momentum, gradient, mask) is ONE flat fp32 buffer with a name -> (offset, shape)
table, so aggregation / masked SGD / saliency / nnz are single flat kernels and a
"state_dict" is just a set of views.

Layout: all trainable parameters first (``named_parameters`` order), then all buffers
(BN running stats, ``num_batches_tracked`` stored as fp32 -- the reference's aggregation
promotes it to float anyway, ``sailentgrads_api.py:219-226``).  Every entry starts on a
16-byte boundary so kernels can use 128-bit accesses; padding elements are kept at zero.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, List, Optional, Tuple

import torch
import torch.nn as nn

ALIGN_ELEMS = 4  # 16 B of fp32


@dataclass(frozen=True)
class SlabEntry:
    name: str
    offset: int
    numel: int
    shape: Tuple[int, ...]
    kind: str          # "param" | "buffer"
    dtype: torch.dtype  # dtype in the exported state_dict
    prunable: bool     # Conv/Linear weight (the tensors SNIP scores and masks)


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class SlabLayout:
    """Name -> (offset, shape) table shared by every slab of one architecture."""

    def __init__(self, entries: List[SlabEntry], state_order: List[str]):
        self.entries: "OrderedDict[str, SlabEntry]" = OrderedDict((e.name, e) for e in entries)
        self.state_order = state_order
        self.param_names = [e.name for e in entries if e.kind == "param"]
        self.buffer_names = [e.name for e in entries if e.kind == "buffer"]
        self.param_numel = 0
        self.total_numel = 0
        for e in entries:
            end = _round_up(e.offset + e.numel, ALIGN_ELEMS)
            self.total_numel = max(self.total_numel, end)
            if e.kind == "param":
                self.param_numel = max(self.param_numel, end)
        self.prunable_names = [e.name for e in entries if e.prunable]

    @classmethod
    def from_module(cls, module: nn.Module) -> "SlabLayout":
        prunable = set()
        for mname, m in module.named_modules():
            if isinstance(m, (nn.Conv1d, nn.Conv2d, nn.Conv3d, nn.Linear)):
                prunable.add((mname + "." if mname else "") + "weight")
        entries: List[SlabEntry] = []
        off = 0
        for name, p in module.named_parameters():
            entries.append(SlabEntry(name, off, p.numel(), tuple(p.shape), "param", p.dtype, name in prunable))
            off = _round_up(off + p.numel(), ALIGN_ELEMS)
        pnames = {e.name for e in entries}
        for name, b in module.named_buffers():
            if name in pnames:
                continue
            entries.append(SlabEntry(name, off, b.numel(), tuple(b.shape), "buffer", b.dtype, False))
            off = _round_up(off + b.numel(), ALIGN_ELEMS)
        return cls(entries, list(module.state_dict().keys()))

    def __contains__(self, name: str) -> bool:
        return name in self.entries

    def __getitem__(self, name: str) -> SlabEntry:
        return self.entries[name]

    def numel_real(self, params_only: bool = False) -> int:
        return sum(e.numel for e in self.entries.values() if (e.kind == "param" or not params_only))

    def prunable_numel(self) -> int:
        return sum(self.entries[n].numel for n in self.prunable_names)


class ParamSlab:
    """One flat fp32 buffer + a layout.  ``flat[:layout.param_numel]`` is the trainable part."""

    def __init__(self, layout: SlabLayout, device=None, flat: Optional[torch.Tensor] = None,
                 params_only: bool = False, fill: float = 0.0):
        self.layout = layout
        self.params_only = params_only
        n = layout.param_numel if params_only else layout.total_numel
        if flat is None:
            flat = torch.full((n,), fill, dtype=torch.float32, device=device)
            if fill != 0.0:
                self._zero_padding(flat)
        assert flat.numel() == n and flat.dtype == torch.float32 and flat.is_contiguous()
        self.flat = flat

    # ------------------------------------------------------------------ views
    def _zero_padding(self, flat):
        prev_end = 0
        for e in self.layout.entries.values():
            if e.offset >= flat.numel():
                break
            if e.offset > prev_end:
                flat[prev_end:e.offset] = 0
            prev_end = e.offset + e.numel
        if prev_end < flat.numel():
            flat[prev_end:] = 0

    def view(self, name: str) -> torch.Tensor:
        e = self.layout[name]
        return self.flat[e.offset:e.offset + e.numel].view(e.shape)

    def views(self, names: Optional[Iterable[str]] = None) -> Dict[str, torch.Tensor]:
        names = list(names) if names is not None else [
            n for n, e in self.layout.entries.items() if not (self.params_only and e.kind == "buffer")]
        return OrderedDict((n, self.view(n)) for n in names)

    @property
    def params_flat(self) -> torch.Tensor:
        return self.flat[:self.layout.param_numel]

    @property
    def device(self):
        return self.flat.device

    # ------------------------------------------------------- module binding
    @classmethod
    def from_module(cls, module: nn.Module, device=None, bind: bool = True) -> "ParamSlab":
        """Build a slab from a module; with ``bind`` the module's parameters/buffers are
        re-pointed at views of the slab, so the module and the slab share storage."""
        layout = SlabLayout.from_module(module)
        device = device if device is not None else next(iter(module.parameters())).device
        slab = cls(layout, device=device)
        with torch.no_grad():
            for name, t in list(module.named_parameters()) + list(module.named_buffers()):
                slab.view(name).copy_(t.detach().to(torch.float32))
        if bind:
            slab.bind_module(module)
        return slab

    def bind_module(self, module: nn.Module, grad: Optional["ParamSlab"] = None) -> None:
        """Point ``module``'s tensors at this slab (and its ``.grad`` at ``grad``)."""
        for name, p in module.named_parameters():
            p.data = self.view(name)
            if grad is not None:
                p.grad = grad.view(name)
        for name, b in module.named_buffers():
            mod, _, leaf = name.rpartition(".")
            owner = module.get_submodule(mod) if mod else module
            e = self.layout[name]
            if e.dtype == torch.float32:
                owner._buffers[leaf] = self.view(name)
            # integer buffers (num_batches_tracked) stay module-local; synced on demand

    def sync_int_buffers_from(self, module: nn.Module) -> None:
        for name, b in module.named_buffers():
            if self.layout[name].dtype != torch.float32:
                self.view(name).copy_(b.to(torch.float32))

    def sync_int_buffers_to(self, module: nn.Module) -> None:
        for name, b in module.named_buffers():
            if self.layout[name].dtype != torch.float32:
                b.copy_(self.view(name).to(b.dtype))

    # ---------------------------------------------------------- state_dict
    def state_dict(self, cast_int: bool = True, clone: bool = True) -> "OrderedDict[str, torch.Tensor]":
        """Export with the reference's keys (``features.N.*`` ...).  ``cast_int`` restores
        int64 ``num_batches_tracked``; ``False`` keeps the float the reference's aggregation yields."""
        out = OrderedDict()
        for name in self.layout.state_order:
            e = self.layout[name]
            v = self.view(name)
            if cast_int and e.dtype != torch.float32:
                v = v.to(e.dtype)
            elif clone:
                v = v.clone()
            out[name] = v
        return out

    def load_state_dict(self, sd, strict: bool = True) -> None:
        with torch.no_grad():
            for name in self.layout.state_order:
                if name not in sd:
                    if strict:
                        raise KeyError(name)
                    continue
                self.view(name).copy_(sd[name].to(device=self.flat.device, dtype=torch.float32).reshape(self.layout[name].shape))

    # -------------------------------------------------------------- helpers
    def clone(self) -> "ParamSlab":
        return ParamSlab(self.layout, flat=self.flat.clone(), params_only=self.params_only)

    def copy_(self, other: "ParamSlab") -> "ParamSlab":
        n = min(self.flat.numel(), other.flat.numel())
        self.flat[:n].copy_(other.flat[:n], non_blocking=True)
        return self

    def zeros_like(self, params_only: Optional[bool] = None) -> "ParamSlab":
        po = self.params_only if params_only is None else params_only
        return ParamSlab(self.layout, device=self.flat.device, params_only=po)

    def to(self, device) -> "ParamSlab":
        return ParamSlab(self.layout, flat=self.flat.to(device), params_only=self.params_only)


def make_mask_slab(layout: SlabLayout, device=None) -> ParamSlab:
    """All-ones mask over the trainable region (biases / BN affine are never pruned,
    ``fedml_api/standalone/sailentgrads/snip.py:107-112``); padding stays zero."""
    return ParamSlab(layout, device=device, params_only=True, fill=1.0)
