#!/usr/bin/env python
"""Small full-coverage workload for compute-sanitizer: two federated rounds on six tiny synthetic
clients -- persistent training kernel (tensor-core variant), tcgen05 scoring kernels, vote / elect / merge / verify /
AUC kernels -- plus one gossip round.

  compute-sanitizer --tool memcheck  python scripts/sanitizer_workload.py
  compute-sanitizer --tool racecheck python scripts/sanitizer_workload.py
  compute-sanitizer --tool synccheck python scripts/sanitizer_workload.py
"""
import os
import random
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.data import ShardStore, SyntheticSpec, make_clients, prepare_from_raw
from fedmse_decentralized_b200.trainer import FederationEngine

dev = torch.device("cuda", 0)
random.seed(0); np.random.seed(0); torch.manual_seed(0)
spec = SyntheticSpec(n_normal=160, n_abnormal=60, n_test_normal=30, jitter_rows=10)
clients, devset = prepare_from_raw(make_clients(6, spec, seed=7), True, seed=3)
for aggregation in ("elected", "gossip"):
    sh = ShardStore.build(clients, devset).to(dev)
    kw = dict(topology="ring") if aggregation == "gossip" else {}
    cfg = ExperimentConfig(epoch=1, network_size=6, backend="cuda", write_latents=False, save_client_models=False,
                           score_precision="tf32x3", num_participants=0.7, aggregation=aggregation, **kw)
    eng = FederationEngine(cfg, sh, "hybrid", "mse_avg")
    for r in range(2):
        res = eng.run_round(r, pipeline_next=(r == 0 and aggregation == "elected"))
    torch.cuda.synchronize()
    print(aggregation, "aggregator", res.aggregator, "AUC", [round(x, 3) for x in res.client_metrics])
