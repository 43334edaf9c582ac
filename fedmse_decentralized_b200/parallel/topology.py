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


def random_k(n: int, k: int, seed: int = 0, round_idx: int = 0) -> List[List[int]]:
    """Each client draws ``k`` distinct random in-neighbours (fresh per round)."""
    k = min(k, n - 1)
    rng = np.random.default_rng([seed, round_idx])
    out = []
    for i in range(n):
        cand = rng.choice(n - 1, size=k, replace=False)
        out.append([int(c + (c >= i)) for c in cand])
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
