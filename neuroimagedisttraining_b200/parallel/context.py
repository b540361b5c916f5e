"""Process-group bootstrap: one process per GPU, ``torch.distributed`` (NCCL on GPU, gloo on CPU)
for rendezvous, barriers and small metadata.  Bulk model traffic goes through
:mod:`.aggregation` (fused symmetric-memory kernels; NCCL all-reduce as the baseline)."""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class DistContext:
    rank: int = 0
    world_size: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    backend: Optional[str] = None
    group: Optional[object] = None

    @property
    def is_distributed(self) -> bool:
        return self.world_size > 1

    @property
    def is_main(self) -> bool:
        return self.rank == 0

    def barrier(self) -> None:
        if self.is_distributed:
            if self.backend == "nccl":
                dist.barrier(device_ids=[self.local_rank])
            else:
                dist.barrier()

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.is_distributed:
            dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.is_distributed:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t

    def broadcast(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.is_distributed:
            dist.broadcast(t, src=src)
        return t

    def owner(self, client_idx: int) -> int:
        """Static client -> rank placement (round-robin keeps sorted sampled subsets balanced)."""
        return int(client_idx) % self.world_size

    def owns(self, client_idx: int) -> bool:
        return self.owner(client_idx) == self.rank


def init_distributed(device: Optional[str] = None, backend: Optional[str] = None) -> DistContext:
    """Read RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* (torchrun) and join the process group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    dev = torch.device(f"cuda:{local_rank}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    if world == 1:
        return DistContext(0, 1, local_rank, dev, None)
    backend = backend or ("nccl" if use_cuda else "gloo")
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        kw = {"device_id": dev} if backend == "nccl" else {}
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return DistContext(rank, world, local_rank, dev, backend, dist.group.WORLD)


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
