"""Phase timing and tracing.

The reference has no profiler hooks at all; phase durations can only be
recovered from log timestamps.  ``PhaseTimer`` brackets each
phase of a round with CUDA events on the launching stream (no host sync until
the numbers are read) and an NVTX range; on CPU it falls back to
``perf_counter``.  Multi-GPU numbers are reduced as the max over ranks by the
caller.
"""
from __future__ import annotations

import contextlib
import time
from collections import defaultdict
from typing import Dict, List

import torch


class PhaseTimer:
    def __init__(self, enabled: bool = True, cuda: bool = False):
        self.enabled = enabled
        self.cuda = cuda and torch.cuda.is_available()
        self._pending: List = []
        self.totals_ms: Dict[str, float] = defaultdict(float)
        self.counts: Dict[str, int] = defaultdict(int)

    @contextlib.contextmanager
    def phase(self, name: str):
        if not self.enabled:
            yield
            return
        if self.cuda:
            torch.cuda.nvtx.range_push(name)
            a = torch.cuda.Event(enable_timing=True)
            b = torch.cuda.Event(enable_timing=True)
            a.record()
            try:
                yield
            finally:
                b.record()
                torch.cuda.nvtx.range_pop()
                self._pending.append((name, a, b))
        else:
            t0 = time.perf_counter()
            try:
                yield
            finally:
                self.totals_ms[name] += (time.perf_counter() - t0) * 1e3
                self.counts[name] += 1

    def flush(self) -> None:
        if self.cuda and self._pending:
            torch.cuda.synchronize()
            for name, a, b in self._pending:
                self.totals_ms[name] += a.elapsed_time(b)
                self.counts[name] += 1
            self._pending.clear()

    def summary(self) -> Dict[str, Dict[str, float]]:
        self.flush()
        return {k: {"total_ms": v, "calls": self.counts[k], "mean_ms": v / max(1, self.counts[k])}
                for k, v in self.totals_ms.items()}

    def reset(self) -> None:
        self.flush()
        self.totals_ms.clear()
        self.counts.clear()
