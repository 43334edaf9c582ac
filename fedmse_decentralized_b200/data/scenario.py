"""Scenario JSON -> prepared clients.

Schema (unchanged from the reference, ``src/Configuration/scen2-nba-iot-10clients.json:1-10``)::

    {"data_path": "...", "devices_list": [{"id", "name", "normal_data_path",
                                           "abnormal_data_path", "test_normal_data_path"}]}

``prepare_from_scenario`` consumes the python / numpy global RNGs in the same
order as ``src/main.py:126-223`` (device sampling, two ``DataFrame.sample(frac=1)``
per client, dev-set sampling), so a run seeded like the reference sees the same
splits (seeding order).
"""
from __future__ import annotations

import json
import logging
import os
import random
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import pandas as pd

from .loader import load_data
from .shards import PreparedClient, pool_dev_set, prepare_client
from .synthetic import ClientRaw

log = logging.getLogger(__name__)


def _find_config(path: str) -> str:
    """Accept the reference's ``Configuration/<name>.json`` spelling: fall back to ``configs/<name>.json``."""
    if os.path.exists(path):
        return path
    repo = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    cand = os.path.join(repo, "configs", os.path.basename(path))
    return cand if os.path.exists(cand) else path


def load_scenario(path: str) -> Dict:
    path = _find_config(path)
    with open(path, "r") as fh:
        cfg = json.load(fh)
    if "devices_list" not in cfg or "data_path" not in cfg:
        raise ValueError(f"{path}: scenario needs 'data_path' and 'devices_list'")
    return cfg


def resolve_data_path(scenario: Dict, scenario_file: Optional[str], data_root: Optional[str]) -> str:
    """Resolve ``data_path``: explicit override, else relative to CWD (reference
    behaviour: run from ``src/``), else relative to the scenario file's parent."""
    p = scenario["data_path"]
    if data_root:
        return data_root
    if os.path.isabs(p) or os.path.isdir(p):
        return p
    if scenario_file:
        base = os.path.dirname(os.path.dirname(os.path.abspath(scenario_file)))
        cand = os.path.normpath(os.path.join(base, p))
        if os.path.isdir(cand):
            return cand
    return p


def prepare_from_scenario(scenario: Dict, network_size: int, new_device: bool = True,
                          scenario_file: Optional[str] = None, data_root: Optional[str] = None
                          ) -> Tuple[List[PreparedClient], np.ndarray]:
    root = resolve_data_path(scenario, scenario_file, data_root)
    devices = random.sample(scenario["devices_list"], network_size)
    clients: List[PreparedClient] = []
    for dev in devices:
        normal = load_data(os.path.join(root, dev["normal_data_path"]))
        normal = normal.sample(frac=1).reset_index(drop=True)
        abnormal = load_data(os.path.join(root, dev["abnormal_data_path"]))
        abnormal = abnormal.sample(frac=1).reset_index(drop=True)
        new_normal = None
        if new_device and dev.get("test_normal_data_path"):
            new_normal = load_data(os.path.join(root, dev["test_normal_data_path"])).to_numpy()
        log.info("%s has %d normal data and %d abnormal data", dev["name"], len(normal), len(abnormal))
        clients.append(prepare_client(dev["name"], normal.to_numpy(), abnormal.to_numpy(),
                                      new_normal, new_device))
    dev = pool_dev_set(clients, sampler=lambda a, n: pd.DataFrame(a).sample(n=n).to_numpy())
    return clients, dev


def prepare_from_raw(raw: Sequence[ClientRaw], new_device: bool = True, seed: Optional[int] = None
                     ) -> Tuple[List[PreparedClient], np.ndarray]:
    """In-memory variant used for synthetic scenarios (no CSV round trip).

    With a ``seed`` every client shuffles with its own generator keyed by ``(seed, index)`` and the dev pool with
    ``(seed, "dev")``: a multi-GPU rank that only holds the abnormal rows of the clients it owns (``make_clients(...,
    full=owned)``) then still builds the same validation shards and dev matrix as every other rank."""
    if seed is None:
        perm_of = lambda i: np.random.permutation            # noqa: E731 - the reference's global-generator behaviour
        perm_dev = np.random.permutation
    else:
        gens = {}

        def perm_of(i):
            return gens.setdefault(i, np.random.default_rng([seed, 1, i])).permutation
        perm_dev = np.random.default_rng([seed, 2]).permutation
    clients = []
    for i, cl in enumerate(raw):
        perm = perm_of(i)
        normal = cl.normal[perm(len(cl.normal))]
        abnormal = cl.abnormal[perm(len(cl.abnormal))]
        clients.append(prepare_client(cl.name, normal, abnormal, cl.test_normal, new_device))
    dev = pool_dev_set(clients, sampler=lambda a, n: a[perm_dev(len(a))[:n]])
    return clients, dev
