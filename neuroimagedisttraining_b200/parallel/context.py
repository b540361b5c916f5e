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
    placement: Optional[list] = None      # client -> rank, set by place_clients(); None = round-robin

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

    def place_clients(self, sample_counts) -> None:
        """Static client -> rank placement by longest-processing-time over the clients' training-set sizes (real ABCD sites are
        uneven); equal sizes give round-robin.  The same on every rank: it only depends on the data partition."""
        from .placement import lpt_assign
        self.placement = lpt_assign(list(sample_counts), self.world_size) if self.world_size > 1 else None

    def owner(self, client_idx: int) -> int:
        """Static client -> rank placement (a client's personal model lives on its owner): LPT by size after
        :meth:`place_clients`, else round-robin (which keeps sorted sampled subsets balanced)."""
        if self.placement is not None and int(client_idx) < len(self.placement):
            return self.placement[int(client_idx)]
        return int(client_idx) % self.world_size

    def owns(self, client_idx: int) -> bool:
        return self.owner(client_idx) == self.rank


def bind_to_gpu_numa(dev: torch.device) -> Optional[int]:
    """Restrict this process to the CPUs NVML reports as local to ``dev`` (its NUMA node), so that the pinned host buffers
    allocated afterwards -- the volume cache every batch is copied from -- sit next to the GPU's PCIe root: with 8 ranks
    pulling ~50 GB/s each, remote-socket memory halves the host->device rate.  Returns the number of CPUs kept, or None when
    nothing was changed (no NVML, affinity not supported, ``NIDT_NO_NUMA_BIND=1``)."""
    if os.environ.get("NIDT_NO_NUMA_BIND", "0") == "1" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(dev).uuid)
        h = pynvml.nvmlDeviceGetHandleByUUID(uuid if uuid.startswith("GPU-") else "GPU-" + uuid)
        ncpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        local = {i for i in range(ncpu) if (int(words[i // 64]) >> (i % 64)) & 1}
        keep = os.sched_getaffinity(0) & local
        if not keep or keep == os.sched_getaffinity(0):
            return None
        os.sched_setaffinity(0, keep)
        return len(keep)
    except Exception:
        return None


@dataclass
class ReplicatedContext(DistContext):
    """Every rank runs the SAME federated programme on the same clients (the ranks cooperate inside each client's forward /
    backward, see :mod:`.spatial`): for the orchestrators this is a one-rank world -- no client placement, no cross-rank
    aggregation -- but only the real rank 0 writes logs, metrics and checkpoints."""
    real_rank: int = 0
    real_world_size: int = 1

    @property
    def is_main(self) -> bool:
        return self.real_rank == 0

    @classmethod
    def of(cls, ctx: DistContext) -> "ReplicatedContext":
        return cls(0, 1, ctx.local_rank, ctx.device, ctx.backend, ctx.group, real_rank=ctx.rank, real_world_size=ctx.world_size)


def init_distributed(device: Optional[str] = None, backend: Optional[str] = None) -> DistContext:
    """Read RANK/LOCAL_RANK/WORLD_SIZE/MASTER_* (torchrun) and join the process group."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and device != "cpu"
    dev = torch.device(f"cuda:{local_rank}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
        if world > 1:
            bind_to_gpu_numa(dev)
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
