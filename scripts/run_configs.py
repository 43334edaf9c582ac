#!/usr/bin/env python
"""Runs the scenarios BASELINE.json names and prints one JSON line each.

  python scripts/run_configs.py 1            # scen2-nba-iot-10clients, 2 rounds x 2 epochs, CPU oracle (plumbing)
  python scripts/run_configs.py 2            # scen2-nba-iot-10clients full run: 20 rounds x 100 epochs, lr 1e-5, lambda 10
  python scripts/run_configs.py 3            # Kitsune-shaped, 50 clients, ring and random-k gossip
  python scripts/run_configs.py 4            # 1000 simulated clients (125 per GPU), random-neighbour gossip
  torchrun --nproc-per-node 8 scripts/run_configs.py 4
  (config 5, the aggregation bandwidth sweep, is scripts/bench_bandwidth_sweep.py)

Real data is used when the prepared splits are reachable (--data_root or /root/reference/Data); otherwise
synthetic traffic of the same shape (the 50-/1000-client shards are not shipped by the reference at all).
"""
import argparse
import json
import os
import random
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import numpy as np
import torch

from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.data import KITSUNE_SPEC, NBAIOT_SPEC, make_clients, prepare_from_raw
from fedmse_decentralized_b200.parallel import Comm, block_partition, init_distributed, shutdown
from fedmse_decentralized_b200.trainer import FederationEngine, build_shards, run_experiment

NBAIOT_DIR = "/root/reference/Data/N-BaIoT/IID-10-Client_Data"


def timed_rounds(eng, n_rounds, warmup=1):
    dev = eng.state.device
    for r in range(warmup):
        eng.run_round(r)
    if dev.type == "cuda":
        torch.cuda.synchronize()
    eng.comm.barrier()
    t0 = time.perf_counter()
    samples = 0
    for r in range(warmup, warmup + n_rounds):
        res = eng.run_round(r)
        samples += res.samples
    if dev.type == "cuda":
        torch.cuda.synchronize()
    eng.comm.barrier()
    dt = time.perf_counter() - t0
    tot = eng.comm.sum_(torch.tensor([float(samples)], dtype=torch.float64, device=dev))
    return dt, float(tot[0]), res


def synthetic_engine(n_clients, spec, comm, cfg_kw, model_type="hybrid", update_type="mse_avg", seed=1234):
    first, cnt = block_partition(n_clients, comm.world, comm.rank)
    raw = make_clients(n_clients, spec, seed=seed, full=range(first, first + cnt))
    clients, dev = prepare_from_raw(raw, True, seed=seed)
    sh, names = build_shards(clients, dev, comm, comm.device)
    random.seed(seed); np.random.seed(seed); torch.manual_seed(seed)
    cfg = ExperimentConfig(network_size=n_clients, write_latents=False, save_client_models=False, **cfg_kw)
    return FederationEngine(cfg, sh, model_type, update_type, comm=comm, names_global=names)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("config", type=int, choices=[1, 2, 3, 4])
    ap.add_argument("--data_root", default=None)
    ap.add_argument("--rounds", type=int, default=None)
    ap.add_argument("--epochs", type=int, default=None)
    ap.add_argument("--clients", type=int, default=None)
    a = ap.parse_args()
    dist = int(os.environ.get("WORLD_SIZE", "1")) > 1
    comm = init_distributed() if dist else Comm(device=torch.device("cuda", 0) if torch.cuda.is_available()
                                                else torch.device("cpu"))
    backend = "cuda" if comm.device.type == "cuda" else "oracle"
    out = tempfile.mkdtemp(prefix="fedmse_cfg_")
    data_root = a.data_root or (NBAIOT_DIR if os.path.isdir(NBAIOT_DIR) else None)

    if a.config in (1, 2):
        full = a.config == 2
        kw = dict(epoch=a.epochs or (100 if full else 2), num_rounds=a.rounds or (20 if full else 2),
                  lr_rate=1e-5 if full else 1e-3, shrink_lambda=10 if full else 5, network_size=10,
                  backend="oracle" if not full else backend, output_root=out, model_types=["hybrid"],
                  update_types=["mse_avg"] if full else ["avg", "fedprox", "mse_avg"])
        if data_root:
            cfg = ExperimentConfig(config_file="configs/scen2-nba-iot-10clients.json", data_root=data_root, **kw)
        else:
            cfg = ExperimentConfig(synthetic_clients=10, **kw)
        t0 = time.perf_counter()
        best = run_experiment(cfg, comm if dist else None)
        dt = time.perf_counter() - t0
        if comm.rank == 0:
            print(json.dumps({"config": a.config, "scenario": "scen2-nba-iot-10clients" + ("" if data_root else " (synthetic)"),
                              "backend": cfg.backend, "rounds": cfg.num_rounds, "epochs": cfg.epoch,
                              "wall_s": round(dt, 2), "best_auc": best, "gpus": comm.world}))
    elif a.config == 3:
        n = a.clients or 50
        for topo in ("ring", "random-k"):
            eng = synthetic_engine(n, KITSUNE_SPEC, comm, dict(epoch=a.epochs or 5, backend=backend, aggregation="gossip",
                                                               topology=topo, topology_k=4, num_participants=0.5))
            dt, samples, res = timed_rounds(eng, a.rounds or 5)
            if comm.rank == 0:
                print(json.dumps({"config": 3, "scenario": f"Kitsune-shaped synthetic, {n} clients, gossip {topo}",
                                  "gpus": comm.world, "rounds_per_sec": (a.rounds or 5) / dt, "local_samples_per_sec": samples / dt,
                                  "auc_mean": float(np.mean(res.client_metrics)), "auc_min": float(np.min(res.client_metrics))}))
    else:
        n = a.clients or 125 * comm.world
        eng = synthetic_engine(n, NBAIOT_SPEC, comm, dict(epoch=a.epochs or 5, backend=backend, aggregation="gossip",
                                                          topology="random-k", topology_k=4, num_participants=0.5))
        dt, samples, res = timed_rounds(eng, a.rounds or 5)
        if comm.rank == 0:
            print(json.dumps({"config": 4, "scenario": f"N-BaIoT-shaped synthetic, {n} clients ({n // comm.world}/GPU), "
                                                       "random-neighbour (k=4) gossip", "gpus": comm.world,
                              "rounds_per_sec": (a.rounds or 5) / dt, "local_samples_per_sec": samples / dt,
                              "auc_mean": float(np.mean(res.client_metrics)), "auc_min": float(np.min(res.client_metrics))}))
    shutdown()


if __name__ == "__main__":
    main()
