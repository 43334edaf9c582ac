"""Experiment sweep: model_types x update_types x num_runs (the reference's ``__main__``).

Restates ``src/main.py:80-400``: per combination re-seed (``set_seeds(run*10000)``
then python/numpy re-seeded with ``data_seed``), load the scenario, prepare
clients, build the engine, run the round loop with global early stop, track the
best metric, write ``training_summary.json``.  Differences are deliberate and
listed in DESIGN.md: the configuration file is re-read correctly on every
combination (quirk Q1 is *not* reproduced -- it only made later combinations
reuse the first config, which is what happens here anyway), data is uploaded to
the device once per combination, and ``reference_compat=False`` resets the
global early-stop counters per combination (quirk Q2).
"""
from __future__ import annotations

import logging
import random
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..config import ExperimentConfig
from ..data import (KITSUNE_SPEC, NBAIOT_SPEC, ShardStore, load_scenario, make_clients,
                    prepare_from_raw, prepare_from_scenario)
from ..parallel import Comm, block_partition
from ..utils import artefacts
from .engine import FederationEngine, GlobalStop

log = logging.getLogger(__name__)


def set_seeds(seed: int) -> None:
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def pick_device(cfg: ExperimentConfig, comm: Optional[Comm] = None) -> torch.device:
    if comm is not None:
        return comm.device
    if cfg.backend == "oracle" or not torch.cuda.is_available():
        return torch.device("cpu")
    return torch.device("cuda", torch.cuda.current_device())


def prepare_data(cfg: ExperimentConfig):
    """Prepared clients + pooled dev matrix, consuming the RNGs like ``src/main.py:126-223``."""
    if cfg.synthetic_clients > 0:
        spec = KITSUNE_SPEC if cfg.synthetic_spec == "kitsune" else NBAIOT_SPEC
        raw = make_clients(cfg.synthetic_clients, spec, seed=cfg.data_seed)
        if cfg.network_size > len(raw):
            raise ValueError(f"--network_size {cfg.network_size} exceeds --synthetic_clients {cfg.synthetic_clients}: the "
                             "experiment samples network_size devices from the scenario (src/main.py:126-129)")
        raw = random.sample(raw, cfg.network_size)
        return prepare_from_raw(raw, cfg.new_device)
    scen = load_scenario(cfg.config_file)
    return prepare_from_scenario(scen, cfg.network_size, cfg.new_device, cfg.config_file, cfg.data_root)


def build_shards(clients, dev, comm: Comm, device) -> Tuple[ShardStore, List[str]]:
    """Rank-local shard store: train/test of owned clients, validation of *all* clients."""
    from ..data.shards import Segmented
    names = [c.name for c in clients]
    first, cnt = block_partition(len(clients), comm.world, comm.rank)
    local = clients[first:first + cnt]
    sh = ShardStore.build(local, dev, first_client=first, n_global=len(clients))
    sh.valid = Segmented.from_list([c.valid for c in clients])     # replicated (vote / verify data)
    return sh.to(device), names


def run_experiment(cfg: ExperimentConfig, comm: Optional[Comm] = None) -> Dict[str, Dict[str, float]]:
    comm = comm or Comm(device=pick_device(cfg))
    device = comm.device
    best: Dict[str, Dict[str, float]] = {mt: {ut: float("-inf") for ut in cfg.update_types}
                                         for mt in cfg.model_types}
    shared_stop = GlobalStop(cfg.global_patience)
    total = len(cfg.model_types) * len(cfg.update_types) * cfg.num_runs
    k = 0
    for model_type in cfg.model_types:
        for update_type in cfg.update_types:
            for run in range(cfg.num_runs):
                k += 1
                log.info("Starting combination %d/%d: model %s, update %s, run %d/%d", k, total,
                         model_type, update_type, run + 1, cfg.num_runs)
                set_seeds(run * 10000)
                random.seed(cfg.data_seed)
                np.random.seed(cfg.data_seed)
                clients, dev = prepare_data(cfg)
                sh, names = build_shards(clients, dev, comm, device)
                eng = FederationEngine(cfg, sh, model_type, update_type, run=run, comm=comm,
                                       names_global=names)
                stop = shared_stop if cfg.reference_compat else GlobalStop(cfg.global_patience)
                try:
                    eng.fit(stop)
                    best[model_type][update_type] = max(best[model_type][update_type], eng.final_best())
                    log.info("phase timings (ms): %s",
                             {p: round(v["mean_ms"], 3) for p, v in eng.timer.summary().items()})
                finally:
                    eng.close()          # peer-mapped slabs are raw device memory: free them per combination
                del eng, sh
                if device.type == "cuda":
                    torch.cuda.empty_cache()
    if comm.rank == 0:
        for mt in cfg.model_types:
            for ut in cfg.update_types:
                log.info("%s + %s: Best %s = %.10f", mt, ut, cfg.metric, best[mt][ut])
        path = artefacts.write_summary(cfg.output_root, cfg.network_size, cfg.no_Exp, best, cfg.metric,
                                       cfg.num_runs)
        log.info("Saved training summary to %s", path)
    return best
