"""Synthetic 115-feature IoT traffic with the shape of the prepared N-BaIoT / Kitsune splits.

There is no network on the build or GPU boxes, and the 50-/1000-client shards
the reference's configs point at are not in its tree, so the
large scenarios and the benchmarks run on traffic generated here.  The generator
is deterministic in ``seed`` and mimics the properties that matter to the
pipeline: rows are low-rank (a few latent "flow statistics" drive all 115
features) with heavy-tailed per-feature scales, a few constant columns, normal
traffic clustered per device family and attacks shifted off the normal manifold.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List, Optional

import numpy as np


@dataclass
class ClientRaw:
    """Raw (un-standardised) float64 matrices of one client."""
    name: str
    normal: np.ndarray
    abnormal: np.ndarray
    test_normal: np.ndarray


@dataclass(frozen=True)
class SyntheticSpec:
    dim: int = 115
    rank: int = 9
    n_normal: int = 1675      # N-BaIoT IID clients have 1,651-1,700 normal rows
    n_abnormal: int = 3250    # 3,182-3,303
    n_test_normal: int = 552
    jitter_rows: int = 25     # +- uniform jitter on the two first counts
    noniid: float = 0.0       # 0 = IID; >0 shifts each client's latent mean
    constant_cols: int = 3


NBAIOT_SPEC = SyntheticSpec()
KITSUNE_SPEC = SyntheticSpec(n_normal=960, n_abnormal=1900, n_test_normal=632, jitter_rows=35)


def _mixing(rng: np.random.Generator, spec: SyntheticSpec):
    A = rng.normal(size=(spec.rank, spec.dim)) / np.sqrt(spec.rank)
    # heavy-tailed per-feature scales (raw N-BaIoT spans ~20 orders of magnitude)
    scale = np.exp(rng.normal(0.0, 2.5, size=spec.dim))
    offset = rng.normal(0.0, 3.0, size=spec.dim) * scale
    const = rng.choice(spec.dim, size=spec.constant_cols, replace=False)
    return A, scale, offset, const


def make_clients(n_clients: int, spec: SyntheticSpec = NBAIOT_SPEC, seed: int = 0,
                 names: Optional[List[str]] = None, full: Optional[range] = None) -> List[ClientRaw]:
    """Generate ``n_clients`` clients of raw float64 traffic.

    ``full`` restricts the clients for which the abnormal / new-device rows are generated (a
    multi-GPU rank needs every client's normal rows for the replicated validation shards and the
    pooled dev set, but test data only for the clients it owns); the others get empty matrices.
    Every client draws from its own generator keyed by ``(seed, index)``, so the rows of a client
    do not depend on which other clients are generated.
    """
    g = np.random.default_rng(seed)
    A, scale, offset, const = _mixing(g, spec)
    A_att = A + 0.8 * g.normal(size=A.shape) / np.sqrt(spec.rank)
    att_shift = g.normal(0.0, 1.5, size=spec.rank)
    out: List[ClientRaw] = []
    for c in range(n_clients):
        want_full = full is None or c in full
        r = np.random.default_rng([seed, c + 1])
        shift = spec.noniid * r.normal(size=spec.rank)
        j = spec.jitter_rows
        n_nor = spec.n_normal + (int(r.integers(-j, j + 1)) if j else 0)
        n_abn = spec.n_abnormal + (int(r.integers(-2 * j, 2 * j + 1)) if j else 0)

        def normal(n):
            u = r.normal(size=(n, spec.rank)) + shift
            x = u @ A + 0.15 * r.normal(size=(n, spec.dim))
            return x

        def attack(n):
            u = 1.8 * r.normal(size=(n, spec.rank)) + att_shift + shift
            x = u @ A_att + 0.6 * r.standard_t(3, size=(n, spec.dim))
            return x

        def finish(x):
            x = x * scale + offset
            x[:, const] = offset[const]
            return x

        empty = np.zeros((0, spec.dim))
        nor = finish(normal(n_nor))
        out.append(ClientRaw(
            name=(names[c] if names else f"Syn-Client-{c + 1}"),
            normal=nor,
            abnormal=finish(attack(n_abn)) if want_full else empty,
            test_normal=finish(normal(spec.n_test_normal)) if want_full else empty,
        ))
    return out


def write_scenario(root: str, n_clients: int, spec: SyntheticSpec = NBAIOT_SPEC, seed: int = 0,
                   prefix: str = "Syn-Client") -> Dict:
    """Write a scenario directory in the reference's on-disk layout and return its config dict.

    Layout: ``root/Client-k/{normal,abnormal,test_normal}/data.csv`` (headerless CSV), i.e.
    what ``Notebook/N-BaIoT/Data-Examination.ipynb`` produces and what the scenario JSONs
    (``src/Configuration/scen2-nba-iot-10clients.json:1-10``) describe.
    """
    clients = make_clients(n_clients, spec, seed)
    devices = []
    for k, cl in enumerate(clients, start=1):
        for part, mat in (("normal", cl.normal), ("abnormal", cl.abnormal),
                          ("test_normal", cl.test_normal)):
            d = os.path.join(root, f"Client-{k}", part)
            os.makedirs(d, exist_ok=True)
            np.savetxt(os.path.join(d, "data.csv"), mat, delimiter=",", fmt="%.10g")
        devices.append({
            "id": k, "name": f"{prefix}-{k}",
            "normal_data_path": f"Client-{k}/normal",
            "abnormal_data_path": f"Client-{k}/abnormal",
            "test_normal_data_path": f"Client-{k}/test_normal",
        })
    return {"data_path": root, "devices_list": devices}
