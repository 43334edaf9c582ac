"""NVLS (multicast) buffers: a symmetric device buffer plus its multicast view.

A multicast address is one virtual address backed by the same offset of a buffer on EVERY GPU of the group:
``multimem.ld_reduce`` on it returns the sum of all GPUs' values, reduced inside the NVSwitch, and ``multimem.st``
writes all copies at once.  The kernels of ``csrc/nvls_kernel.cu`` use it for the elected-mode merge (the reference's
gather + average + broadcast, ``src/Trainer/client_trainer.py:107-130``, ``src/main.py:298-300``) and for the
wide-model all-reduce.

Creating the multicast object needs a handle exchange between the processes (file descriptors over a unix socket);
that plumbing is ``torch.distributed._symmetric_memory``, the kernels are ours.  Hosts whose fabric has no multicast
support (``multicast_ptr == 0``), older torch builds, or ``FEDMSE_NVLS=0`` fall back to the peer-load kernels.
"""
from __future__ import annotations

import logging
import os
from typing import Optional

import torch

from .world import Comm

log = logging.getLogger(__name__)


class NvlsBuffer:
    """``n_floats`` fp32 per rank, zero-initialised; ``tensor`` is the local view, ``mc_ptr`` the multicast address."""

    def __init__(self, tensor: torch.Tensor, handle, mc_ptr: int):
        self.tensor, self.handle, self.mc_ptr = tensor, handle, mc_ptr
        self.ptr = tensor.data_ptr()


def allocate(comm: Comm, n_floats: int) -> Optional[NvlsBuffer]:
    """Collective.  Returns None (on every rank alike) when multicast is unavailable."""
    if not comm.is_dist or comm.device.type != "cuda" or os.environ.get("FEDMSE_NVLS", "1") == "0":
        return None
    buf, why = None, ""
    try:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = comm.group if comm.group is not None else dist.group.WORLD
        enable = getattr(symm_mem, "enable_symm_mem_for_group", None)
        if enable is not None:                      # needed by older torch builds, a deprecated no-op in newer ones
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                try:
                    enable(group.group_name)
                except Exception:
                    pass
        t = symm_mem.empty(n_floats, dtype=torch.float32, device=comm.device)
        t.zero_()
        hdl = symm_mem.rendezvous(t, group)
        mc = int(getattr(hdl, "multicast_ptr", 0) or 0)
        if mc:
            buf = NvlsBuffer(t, hdl, mc)
        else:
            why = "multicast_ptr == 0 (fabric without NVLS multicast)"
    except Exception as e:                          # noqa: BLE001 - any failure means "no NVLS here"
        why = f"{type(e).__name__}: {e}"
    ok = all(comm.all_gather_object(buf is not None))
    if not ok:
        if comm.rank == 0:
            log.info("NVLS multicast unavailable (%s): using peer-load kernels", why or "another rank failed")
        return None
    torch.cuda.synchronize(comm.device)
    comm.barrier()
    return buf
