"""Peer-mapped ("symmetric") device slabs.

Every rank allocates the same-sized slab with ``cudaMalloc`` through the C
runtime (``csrc/runtime.cu``), exports it with ``cudaIpcGetMemHandle``, and maps
every peer's slab with ``cudaIpcOpenMemHandle``.  The resulting table of base
pointers (one per rank, own slab included) is handed to the merge / gossip
kernels, which read peers' model parameters with ordinary loads that travel over
NVLink 5 / NVSwitch -- the framework's replacement for the reference's
"receive_model stores a dict reference" (``src/Trainer/client_trainer.py:148-151``)
and for an NCCL all-gather of models.

``torch.distributed`` is used only to exchange the 64-byte IPC handles.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch

from ..ops import _lib
from .world import Comm


class _RawCuda:
    """Zero-copy view of raw device memory for ``torch.as_tensor`` (CUDA array interface)."""

    def __init__(self, ptr: int, shape: Tuple[int, ...], typestr: str = "<f4"):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 2}


def wrap_device_memory(ptr: int, shape: Tuple[int, ...], device: torch.device) -> torch.Tensor:
    return torch.as_tensor(_RawCuda(ptr, shape), device=device)


class SymmetricSlab:
    """A ``[rows, cols]`` fp32 slab on every rank plus the device table of all ranks' base pointers."""

    def __init__(self, rows: int, cols: int, comm: Comm, rows_max: Optional[int] = None):
        lib = _lib.load()
        self.comm, self.rows, self.cols = comm, rows, cols
        device = comm.device
        assert device.type == "cuda"
        rows_alloc = max(rows_max or rows, 1)
        self.nbytes = rows_alloc * cols * 4
        p = C.c_void_p()
        _lib.check(lib.fm_slab_alloc(self.nbytes, C.byref(p)), "slab_alloc")
        self.ptr = int(p.value)
        self.tensor = wrap_device_memory(self.ptr, (rows, cols), device)
        self.peer_ptrs: List[int] = [self.ptr]
        self._opened: List[int] = []
        if comm.is_dist:
            buf = C.create_string_buffer(64)
            _lib.check(lib.fm_ipc_get_handle(self.ptr, buf), "ipc_get_handle")
            handles = comm.all_gather_object(bytes(buf.raw))
            self.peer_ptrs = []
            for r, h in enumerate(handles):
                if r == comm.rank:
                    self.peer_ptrs.append(self.ptr)
                    continue
                q = C.c_void_p()
                _lib.check(lib.fm_ipc_open_handle(C.create_string_buffer(h, 64), C.byref(q)),
                           f"ipc_open_handle(rank {r})")
                self.peer_ptrs.append(int(q.value))
                self._opened.append(int(q.value))
            comm.barrier()
        self.table = torch.tensor(self.peer_ptrs, dtype=torch.int64, device=device)

    def close(self) -> None:
        lib = _lib.load()
        for q in self._opened:
            lib.fm_ipc_close_handle(q)
        self._opened.clear()
        if self.ptr:
            if self.comm.is_dist:
                torch.cuda.synchronize()
                self.comm.barrier()
            lib.fm_slab_free(self.ptr)
            self.ptr = 0
