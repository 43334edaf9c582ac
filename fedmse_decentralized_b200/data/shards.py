"""Device-resident client shards.

The reference walks three ``torch.utils.data.DataLoader`` objects per client in
batches of 12 and copies every batch host->device (``src/main.py:180-195``,
``src/Trainer/client_trainer.py:371``).  Here every client's standardised rows
are uploaded once into a few large ``[rows, 128]`` fp32 matrices (115 features,
row pitch padded to 512 B so rows are TMA/bulk-copy legal and K is MMA-padded;
padding is zero) plus CSR-style ``[clients+1]`` offset tables.  Kernels index
``x[off[c] + r]``; nothing is copied per batch.

``prepare_client`` restates the per-client pipeline of ``src/main.py:131-207``
(SURVEY.md A.1): shuffle, 40/10/40/10 split of the normal rows, StandardScaler
fitted on the training split only, test set = [held-out normal ; new-device
normal] (label 0) followed by all abnormal rows (label 1).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, Dict, List, Optional, Sequence

import numpy as np
import torch

from ..layout import D_PAD
from .loader import IoTDataProccessor


@dataclass
class PreparedClient:
    name: str
    train: np.ndarray          # float64 standardised [n_tr, D]
    valid: np.ndarray
    test: np.ndarray           # normal-test ; new-device ; abnormal
    test_label: np.ndarray     # 0/1
    dev_raw: np.ndarray        # raw (unscaled) dev slice, pooled later


def split_sizes(n: int):
    """40 / 10 / 40 / remainder split of ``n`` normal rows (``src/main.py:151-154``)."""
    tr, va, dv = int(0.4 * n), int(0.1 * n), int(0.4 * n)
    return tr, va, dv, n - tr - va - dv


def prepare_client(name: str, normal: np.ndarray, abnormal: np.ndarray,
                   test_normal: Optional[np.ndarray], new_device: bool = True,
                   scaler: str = "standard") -> PreparedClient:
    """Split + standardise one client. ``normal``/``abnormal`` must already be shuffled."""
    n_tr, n_va, n_dv, _ = split_sizes(len(normal))
    tr = normal[:n_tr]
    va = normal[n_tr:n_tr + n_va]
    dv = normal[n_tr + n_va:n_tr + n_va + n_dv]
    te = normal[n_tr + n_va + n_dv:]
    proc = IoTDataProccessor(scaler=scaler)
    p_tr, _ = proc.fit_transform(tr)
    p_va, _ = proc.transform(va)
    p_te, l_te = proc.transform(te)
    p_ab, l_ab = proc.transform(abnormal, type="abnormal")
    if new_device and test_normal is not None:
        p_new, l_new = proc.transform(test_normal)
        p_te = np.concatenate([p_te, p_new], axis=0)
        l_te = np.concatenate([l_te, l_new], axis=0)
    test = np.concatenate([p_te, p_ab], axis=0)
    label = np.concatenate([l_te, l_ab], axis=0)
    return PreparedClient(name, np.asarray(p_tr), np.asarray(p_va), test, label, np.asarray(dv))


def pool_dev_set(clients: Sequence[PreparedClient],
                 sampler: Optional[Callable[[np.ndarray, int], np.ndarray]] = None) -> np.ndarray:
    """Pooled development matrix (``src/main.py:214-223``, quirk Q17).

    ``min_len`` rows are sampled from every client's *raw* dev slice, pooled, and
    standardised with a scaler fitted on the pool itself.
    """
    min_len = min(len(c.dev_raw) for c in clients)
    if sampler is None:
        def sampler(a, n):  # numpy global RNG, like DataFrame.sample(n=...)
            return a[np.random.permutation(len(a))[:n]]
    pool = np.concatenate([sampler(c.dev_raw, min_len) for c in clients], axis=0)
    dev, _ = IoTDataProccessor("standard").fit_transform(pool)
    return np.asarray(dev)


def _pad_rows(mat: np.ndarray) -> torch.Tensor:
    out = torch.zeros(mat.shape[0], D_PAD, dtype=torch.float32)
    out[:, :mat.shape[1]] = torch.from_numpy(np.ascontiguousarray(mat)).to(torch.float32)
    return out


@dataclass
class Segmented:
    """Rows of many clients packed back to back: client ``c`` owns ``x[off[c]:off[c+1]]``."""
    x: torch.Tensor                      # [R, D_PAD] fp32
    off: torch.Tensor                    # [C+1] int32, same device as x
    off_host: List[int]
    y: Optional[torch.Tensor] = None     # [R] fp32 labels (test only)

    @classmethod
    def from_list(cls, mats: Sequence[np.ndarray], labels: Optional[Sequence[np.ndarray]] = None):
        offs = [0]
        for m in mats:
            offs.append(offs[-1] + len(m))
        x = torch.cat([_pad_rows(m) for m in mats], dim=0) if mats else torch.zeros(0, D_PAD)
        y = None
        if labels is not None:
            y = torch.cat([torch.from_numpy(np.asarray(l)).to(torch.float32) for l in labels])
        return cls(x, torch.tensor(offs, dtype=torch.int32), offs, y)

    def rows(self, c: int) -> torch.Tensor:
        return self.x[self.off_host[c]:self.off_host[c + 1]]

    def labels(self, c: int) -> torch.Tensor:
        return self.y[self.off_host[c]:self.off_host[c + 1]]

    def count(self, c: int) -> int:
        return self.off_host[c + 1] - self.off_host[c]

    def max_count(self) -> int:
        return max((self.count(c) for c in range(len(self.off_host) - 1)), default=0)

    def to(self, device, non_blocking: bool = False) -> "Segmented":
        return Segmented(self.x.to(device, non_blocking=non_blocking),
                         self.off.to(device, non_blocking=non_blocking), list(self.off_host),
                         None if self.y is None else self.y.to(device, non_blocking=non_blocking))

    def pin(self) -> "Segmented":
        return Segmented(self.x.pin_memory(), self.off, list(self.off_host),
                         None if self.y is None else self.y.pin_memory())


@dataclass
class ShardStore:
    """All clients' shards of one rank plus the replicated dev matrix."""
    dim: int
    names: List[str]
    train: Segmented
    valid: Segmented
    test: Segmented
    dev: torch.Tensor                    # [n_dev, D_PAD]
    first_client: int = 0                # global index of local client 0 (multi-GPU)
    n_global: int = 0

    def __post_init__(self):
        if not self.n_global:
            self.n_global = len(self.names)

    @property
    def n_clients(self) -> int:
        return len(self.names)

    @property
    def device(self) -> torch.device:
        return self.train.x.device

    @classmethod
    def build(cls, clients: Sequence[PreparedClient], dev: np.ndarray, first_client: int = 0,
              n_global: int = 0) -> "ShardStore":
        dim = clients[0].train.shape[1]
        if dim > D_PAD:
            raise ValueError(f"feature dim {dim} exceeds shard pitch {D_PAD}")
        return cls(
            dim=dim,
            names=[c.name for c in clients],
            train=Segmented.from_list([c.train for c in clients]),
            valid=Segmented.from_list([c.valid for c in clients]),
            test=Segmented.from_list([c.test for c in clients], [c.test_label for c in clients]),
            dev=_pad_rows(dev),
            first_client=first_client,
            n_global=n_global or len(clients),
        )

    def to(self, device, non_blocking: bool = False) -> "ShardStore":
        return ShardStore(self.dim, list(self.names), self.train.to(device, non_blocking),
                          self.valid.to(device, non_blocking), self.test.to(device, non_blocking),
                          self.dev.to(device, non_blocking=non_blocking), self.first_client,
                          self.n_global)

    def nbytes(self) -> int:
        tot = 0
        for seg in (self.train, self.valid, self.test):
            tot += seg.x.numel() * 4
        return tot + self.dev.numel() * 4
