"""Peer topologies for gossip aggregation.

The reference always wires a full mesh (``src/main.py:260-263``) and its gossip
API (``broadcast_model`` / ``request_aggregation``, ``src/Trainer/
client_trainer.py:141-172``) is never called by ``main.py`` (quirk Q21).  The
gossip mode of this framework makes the topology configurable: each client
merges the models of its in-neighbours (plus its own).  Neighbour lists are
stored CSR-style (``off[N+1]``, ``idx[nnz]``) so the gossip kernel walks them on
the device.
"""
from __future__ import annotations

from typing import List, Tuple

import numpy as np
import torch


def full_mesh(n: int) -> List[List[int]]:
    return [[j for j in range(n) if j != i] for i in range(n)]


def ring(n: int, hops: int = 1) -> List[List[int]]:
    """Bidirectional ring: ``hops`` neighbours on each side."""
    out = []
    for i in range(n):
        nb = []
        for h in range(1, hops + 1):
            for j in ((i - h) % n, (i + h) % n):
                if j != i and j not in nb:
                    nb.append(j)
        out.append(nb)
    return out


_M64 = (1 << 64) - 1


def _splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def random_neighbour(seed: int, round_idx: int, g: int, j: int, attempt: int, n: int) -> int:
    """Draw ``attempt`` of neighbour slot ``j`` of client ``g`` in round ``round_idx`` (uniform over the other
    ``n - 1`` clients).  Bit-for-bit the device function of ``csrc/gossip_kernel.cu`` (``fm_random_neighbour``):
    the gossip kernel generates its random-k graph from ``(seed, round)`` alone, nothing is uploaded."""
    h = _splitmix64((seed & _M64) ^ ((0xD1B54A32D192ED03 * (round_idx + 1)) & _M64))
    h = _splitmix64(h ^ ((0x9E3779B97F4A7C15 * (g + 1)) & _M64))
    h = _splitmix64(h ^ ((j << 32) | attempt))
    c = h % (n - 1)
    return c + (1 if c >= g else 0)


def random_k(n: int, k: int, seed: int = 0, round_idx: int = 0) -> List[List[int]]:
    """Each client draws ``k`` distinct random in-neighbours (fresh per round); a slot that repeats an earlier
    pick is redrawn with the next ``attempt``."""
    k = min(k, n - 1)
    out = []
    for i in range(n):
        nb: List[int] = []
        for j in range(k):
            attempt = 0
            while True:
                c = random_neighbour(seed, round_idx, i, j, attempt, n)
                if c not in nb:
                    nb.append(c)
                    break
                attempt += 1
        out.append(nb)
    return out


def build(name: str, n: int, k: int = 4, seed: int = 0, round_idx: int = 0) -> List[List[int]]:
    if name == "full":
        return full_mesh(n)
    if name == "ring":
        return ring(n, max(1, k // 2) if k > 2 else 1)
    if name in ("random-k", "random_k", "random"):
        return random_k(n, k, seed, round_idx)
    raise ValueError(f"unknown topology {name!r}")


def to_csr(neigh: List[List[int]], include_self: bool = True) -> Tuple[torch.Tensor, torch.Tensor]:
    """CSR tensors (int32): self first when ``include_self`` (the aggregating client's own model)."""
    off, idx = [0], []
    for i, nb in enumerate(neigh):
        row = ([i] if include_self else []) + list(nb)
        idx.extend(row)
        off.append(len(idx))
    return torch.tensor(off, dtype=torch.int32), torch.tensor(idx, dtype=torch.int32)
