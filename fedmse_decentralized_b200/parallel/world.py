"""Process-group plumbing: one process per GPU, ``torch.distributed`` for bootstrap and scalars.

The reference has no communication layer at all: its "peers" are Python objects
in one process and "sending a model" stores a dict reference
(``src/Trainer/client_trainer.py:136-151``).  Here clients are
block-partitioned over ranks; model parameters cross GPUs only inside kernels
through peer-mapped memory (``parallel/symmetric.py``), while this module covers
rendezvous and the tiny scalar exchanges (election scores, merge weights, per-
client AUC: the C4/C5 collectives of SURVEY.md §2.6).  With the ``gloo`` backend
the same code runs multi-process on a CPU box, which is how the host-side logic
is tested without GPUs.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def block_partition(n_global: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous block of clients owned by ``rank``: (first, count)."""
    base, rem = divmod(n_global, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


def owner_of(g: int, n_global: int, world: int) -> Tuple[int, int]:
    """(rank, local index) of global client ``g`` under ``block_partition``."""
    base, rem = divmod(n_global, world)
    cut = rem * (base + 1)
    if g < cut:
        return g // (base + 1), g % (base + 1)
    return rem + (g - cut) // base, (g - cut) % base


@dataclass
class Comm:
    rank: int = 0
    world: int = 1
    device: torch.device = torch.device("cpu")
    group: Optional[object] = None

    @property
    def is_dist(self) -> bool:
        return self.world > 1

    def sum_(self, t: torch.Tensor) -> torch.Tensor:
        if self.is_dist:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def max_(self, t: torch.Tensor) -> torch.Tensor:
        if self.is_dist:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t

    def all_gather(self, t: torch.Tensor) -> torch.Tensor:
        if not self.is_dist:
            return t.unsqueeze(0)
        out = torch.empty((self.world,) + tuple(t.shape), dtype=t.dtype, device=t.device)
        dist.all_gather_into_tensor(out.view(-1), t.contiguous().view(-1), group=self.group) \
            if t.is_cuda else dist.all_gather(list(out.unbind(0)), t.contiguous(), group=self.group)
        return out

    def all_gather_object(self, obj):
        if not self.is_dist:
            return [obj]
        out = [None] * self.world
        dist.all_gather_object(out, obj, group=self.group)
        return out

    def gather_rows(self, slab: torch.Tensor, sel: Sequence[int], first: int) -> torch.Tensor:
        """Selected models ``[n_sel, PP]`` assembled from their owners (host/gloo path only;
        the CUDA path reads peers' slabs from inside the merge kernel instead)."""
        n_local = slab.shape[0]
        rows = torch.zeros(len(sel), slab.shape[1], dtype=slab.dtype, device=slab.device)
        for k, g in enumerate(sel):
            if first <= g < first + n_local:
                rows[k] = slab[g - first]
        return self.sum_(rows)

    def barrier(self) -> None:
        if self.is_dist:
            dist.barrier(group=self.group)


def bind_to_gpu_numa(device_index: int) -> Optional[int]:
    """Pin this process to the CPU cores NVML reports as local to ``device_index`` (same NUMA node / PCIe root).

    Pinned host buffers are placed by first touch, so a rank that runs on the GPU's own socket stages its per-round
    uploads from local memory instead of across the socket interconnect -- with 8 ranks uploading 53 MB per round each
    that is the difference between the host side keeping up with the GPUs or not.  Best effort: returns the number of
    cores bound to, or None when NVML / the affinity call is unavailable (``FEDMSE_NO_NUMA_BIND=1`` disables it)."""
    if os.environ.get("FEDMSE_NO_NUMA_BIND") == "1" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            visible = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = device_index
            if visible:                                  # NVML indexes physical devices
                ids = [v.strip() for v in visible.split(",") if v.strip()]
                if device_index < len(ids) and ids[device_index].isdigit():
                    idx = int(ids[device_index])
            h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            n_cpu = os.cpu_count() or 64
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
            cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
            cpus &= set(os.sched_getaffinity(0))         # never leave the cores this process was given
            if not cpus:
                return None
            os.sched_setaffinity(0, cpus)
            return len(cpus)
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        return None


def init_distributed(backend: Optional[str] = None) -> Comm:
    """Create the process group from the torchrun environment (RANK / WORLD_SIZE / MASTER_*)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    use_cuda = torch.cuda.is_available()
    if use_cuda:
        torch.cuda.set_device(local % max(1, torch.cuda.device_count()))
        device = torch.device("cuda", torch.cuda.current_device())
        n_bound = bind_to_gpu_numa(device.index)
        if os.environ.get("FEDMSE_DEBUG_BIND") == "1":
            print(f"[world] rank {rank}: cuda:{device.index}, bound to {n_bound} local cores", flush=True)
    else:
        device = torch.device("cpu")
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if use_cuda else "gloo"
        kw = {}
        if backend == "nccl":
            kw["device_id"] = device
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kw)
    return Comm(rank=rank, world=world, device=device)


def shutdown() -> None:
    if dist.is_available() and dist.is_initialized():
        dist.destroy_process_group()
