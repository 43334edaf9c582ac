"""Re-partition pooled traffic into N federated clients.

Restates what ``Notebook/N-BaIoT/Data-Examination.ipynb`` (cells 5-32) does with
FedArtML's ``SplitAsFederatedData(method="dirichlet")``: pool the per-device
rows, draw per-client label proportions from ``Dir(alpha)`` (``alpha=1000`` is
the notebook's IID setting, small alpha is non-IID), hold out 40 % of the normal
rows as the "new device" test set and write
``Client-k/{normal,abnormal,test_normal}/data.csv``.  Needed for the 15/20/50-
client configs whose directories are not shipped.
"""
from __future__ import annotations

import os
from typing import Dict, List, Sequence

import numpy as np

from .synthetic import ClientRaw


def dirichlet_partition(labels: np.ndarray, n_clients: int, alpha: float, seed: int = 0,
                        min_rows: int = 10) -> List[np.ndarray]:
    """Row indices per client; class ``k`` is split with proportions ``Dir(alpha)``."""
    rng = np.random.default_rng(seed)
    parts: List[List[int]] = [[] for _ in range(n_clients)]
    for k in np.unique(labels):
        idx = np.flatnonzero(labels == k)
        rng.shuffle(idx)
        prop = rng.dirichlet(np.full(n_clients, alpha))
        cuts = (np.cumsum(prop) * len(idx)).astype(int)[:-1]
        for c, chunk in enumerate(np.split(idx, cuts)):
            if len(chunk) >= min_rows:          # notebook drops classes with < 10 rows
                parts[c].extend(chunk.tolist())
    return [np.asarray(sorted(p), dtype=np.int64) for p in parts]


def repartition(devices: Sequence[ClientRaw], n_clients: int, alpha: float = 1000.0,
                holdout: float = 0.4, seed: int = 0, prefix: str = "Client") -> List[ClientRaw]:
    """Pool ``devices`` and deal them into ``n_clients`` new clients."""
    rng = np.random.default_rng(seed)
    normal = np.concatenate([d.normal for d in devices])
    n_lab = np.concatenate([np.full(len(d.normal), i) for i, d in enumerate(devices)])
    abnormal = np.concatenate([d.abnormal for d in devices])
    a_lab = np.concatenate([np.full(len(d.abnormal), i) for i, d in enumerate(devices)])
    p = rng.permutation(len(normal))
    n_hold = int(holdout * len(normal))
    hold, keep = p[:n_hold], p[n_hold:]
    n_parts = dirichlet_partition(n_lab[keep], n_clients, alpha, seed)
    a_parts = dirichlet_partition(a_lab, n_clients, alpha, seed + 1)
    h_parts = np.array_split(hold, n_clients)
    out = []
    for c in range(n_clients):
        out.append(ClientRaw(f"{prefix}-{c + 1}", normal[keep][n_parts[c]], abnormal[a_parts[c]],
                             normal[h_parts[c]]))
    return out


def write_clients(root: str, clients: Sequence[ClientRaw], name_prefix: str = "Client") -> Dict:
    devices = []
    for k, cl in enumerate(clients, start=1):
        for part, mat in (("normal", cl.normal), ("abnormal", cl.abnormal),
                          ("test_normal", cl.test_normal)):
            d = os.path.join(root, f"Client-{k}", part)
            os.makedirs(d, exist_ok=True)
            np.savetxt(os.path.join(d, "data.csv"), mat, delimiter=",", fmt="%.10g")
        devices.append({"id": k, "name": f"{name_prefix}-{k}",
                        "normal_data_path": f"Client-{k}/normal",
                        "abnormal_data_path": f"Client-{k}/abnormal",
                        "test_normal_data_path": f"Client-{k}/test_normal"})
    return {"data_path": root, "devices_list": devices}
