"""Weighted all-reduce of wide models by in-kernel peer loads (``csrc/allreduce_p2p.cu``).

``P2PReducer.allreduce(rows, w)`` returns ``sum over all ranks' rows of w * row`` on every rank without
NCCL: a local weighted partial, a reduce-scatter by pulling slices of the peers' partials, and an
all-gather by pulling the reduced slices -- ``2 (N-1)/N * P`` bytes per GPU over NVLink.
``allreduce_nvls`` does the same through the NVSwitch's multicast reduction (``multimem.ld_reduce`` /
``multimem.st``, ``csrc/nvls_kernel.cu``) where the fabric exposes it.  ``pull_all``
is the latency-optimal variant the 27 KB merge uses (every rank reads every row).  Buffers are
peer-mapped slabs; ordering uses release/acquire arrival counters in a peer-mapped control block.
"""
from __future__ import annotations

import ctypes as C

import torch

from ..ops import _lib
from .symmetric import SymmetricSlab
from .world import Comm


class P2PReducer:
    def __init__(self, comm: Comm, n_floats: int, max_local_rows: int = 1):
        assert n_floats % 4 == 0
        self.comm, self.P, self.m = comm, n_floats, max_local_rows
        self.lib = _lib.load()
        dev = comm.device
        with torch.cuda.device(dev):
            self.rows = SymmetricSlab(max_local_rows, n_floats, comm)      # this rank's models
            self.partial = SymmetricSlab(1, n_floats, comm)
            self.reduced = SymmetricSlab(1, n_floats, comm)
            self.ctl = SymmetricSlab(1, _lib.CTL_FLOATS, comm)
        self.out = torch.empty(n_floats, dtype=torch.float32, device=dev)
        self.grid_counter = torch.zeros(4, dtype=torch.int32, device=dev)
        self.arrive_ptr = self.ctl.ptr + 4 * 2 * _lib.MAX_SEL
        self.seq = {3: 0, 4: 0, 5: 0}
        self._local_ptrs = torch.tensor([self.rows.ptr + 4 * n_floats * i for i in range(max_local_rows)],
                                        dtype=torch.int64, device=dev)
        w = comm.world
        self._all_ptrs = torch.tensor([p + 4 * n_floats * i for p in self.rows.peer_ptrs for i in range(max_local_rows)],
                                      dtype=torch.int64, device=dev)

    def _stream(self):
        return torch.cuda.current_stream(self.comm.device).cuda_stream

    def allreduce(self, w_local: torch.Tensor) -> torch.Tensor:
        """sum over ranks r, local rows i of w_local_r[i] * rows_r[i]  (bandwidth-optimal)."""
        c, lib, W = self.comm, self.lib, self.comm.world
        s = self._stream()
        ctl_tab = self.ctl.table.data_ptr() if c.is_dist else 0
        if not c.is_dist:
            _lib.check(lib.fm_weighted_rows(self._local_ptrs.data_ptr(), w_local.data_ptr(), self.m, self.out.data_ptr(),
                                            self.P, 0, 0, 0, 1, 0, 0, s), "weighted_rows")
            return self.out
        self.seq[3] += 1
        _lib.check(lib.fm_weighted_rows(self._local_ptrs.data_ptr(), w_local.data_ptr(), self.m, self.partial.ptr, self.P,
                                        0, 0, ctl_tab, W, 3, self.grid_counter.data_ptr(), s), "weighted_rows")
        self.seq[4] += 1
        _lib.check(lib.fm_slice_reduce(self.partial.table.data_ptr(), W, c.rank, self.reduced.ptr, self.P,
                                       self.arrive_ptr + 4 * 3, W * self.seq[3], ctl_tab, 4,
                                       self.grid_counter.data_ptr() + 4, s), "slice_reduce")
        _lib.check(lib.fm_slice_gather(self.reduced.table.data_ptr(), W, self.out.data_ptr(), self.P,
                                       self.arrive_ptr + 4 * 4, W * self.seq[4], s), "slice_gather")
        return self.out

    def enable_nvls(self) -> bool:
        """Collective.  Allocate the multicast-mapped partial buffer; True when the fabric supports NVLS."""
        from . import nvls
        if getattr(self, "nv", None) is None:
            self.nv = nvls.allocate(self.comm, self.P)
            self.seq_nv = {5: 0, 6: 0}
        return self.nv is not None

    def allreduce_nvls(self, w_local: torch.Tensor) -> torch.Tensor:
        """Same result as ``allreduce`` through the NVSwitch: local weighted partial into the multicast-mapped buffer,
        then each rank reduces its slice with ``multimem.ld_reduce`` and re-broadcasts it with ``multimem.st`` -- the
        buffer itself holds the complete sum on every rank afterwards (no gather pass, no second buffer)."""
        c, lib, W, nv = self.comm, self.lib, self.comm.world, self.nv
        s = self._stream()
        ctl_tab = self.ctl.table.data_ptr()
        self.seq_nv[5] += 1
        self.seq_nv[6] += 1
        _lib.check(lib.fm_weighted_rows(self._local_ptrs.data_ptr(), w_local.data_ptr(), self.m, nv.ptr, self.P,
                                        0, 0, ctl_tab, W, 5, self.grid_counter.data_ptr() + 8, s), "weighted_rows")
        _lib.check(lib.fm_nvls_allreduce_slice(nv.mc_ptr, W, c.rank, self.P, self.arrive_ptr + 4 * 5, W * self.seq_nv[5],
                                               ctl_tab, 6, self.grid_counter.data_ptr() + 12, s), "nvls_allreduce_slice")
        _lib.check(lib.fm_flag_wait(self.arrive_ptr + 4 * 6, W * self.seq_nv[6], s), "flag_wait")
        return nv.tensor

    def pull_all(self, w_all: torch.Tensor) -> torch.Tensor:
        """Every rank reads every row of every rank (rows must be complete: caller synchronises ranks first)."""
        n = self.comm.world * self.m
        _lib.check(self.lib.fm_weighted_rows(self._all_ptrs.data_ptr(), w_all.data_ptr(), n, self.out.data_ptr(), self.P,
                                             0, 0, 0, self.comm.world, 0, 0, self._stream()), "weighted_rows")
        return self.out

    def close(self) -> None:
        torch.cuda.synchronize(self.comm.device)
        for slab in (self.rows, self.partial, self.reduced, self.ctl):
            slab.close()
