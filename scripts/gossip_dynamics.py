#!/usr/bin/env python
"""How does a gossip federation evolve over many rounds?  Per-round AUC statistics plus state diagnostics
(non-finite models, parameter norms, rejection counters) on the bench scenario, one GPU.

  python scripts/gossip_dynamics.py --clients 1000 --rounds 30 [--model hybrid] [--update mse_avg] [--topology random-k]
"""
import argparse
import json
import os
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch

import bench
from fedmse_decentralized_b200.parallel import Comm

ap = argparse.ArgumentParser()
ap.add_argument("--clients", type=int, default=1000)
ap.add_argument("--rounds", type=int, default=30)
ap.add_argument("--model", default="hybrid")
ap.add_argument("--update", default="mse_avg")
ap.add_argument("--aggregation", default="gossip")
ap.add_argument("--topology", default="random-k")
ap.add_argument("--topology-k", type=int, default=4)
ap.add_argument("--participation", type=float, default=0.5)
ap.add_argument("--epochs", type=int, default=5)
ap.add_argument("--compat", type=int, default=1)
a = ap.parse_args()
args = types.SimpleNamespace(seed=1234, participation=a.participation, epochs=a.epochs, score_precision="fp16x3",
                             aggregation=a.aggregation, topology=a.topology, topology_k=a.topology_k, model=a.model,
                             update=a.update, backend="cuda")
comm = Comm(rank=0, world=1, device=torch.device("cuda", 0))
eng, sh, pt, pv = bench.build_engine(args, comm, a.clients, 0, a.clients)
eng.cfg.reference_compat = bool(a.compat)
bench.seed_all(1251)
rows = []
for r in range(a.rounds):
    res = eng.run_round(r)
    st = eng.state
    P = st.params[:, :eng.layout.n_params]
    finite = torch.isfinite(P).all(dim=1)
    m = torch.tensor(res.client_metrics)
    rows.append({"round": r, "auc_mean": round(float(m.mean()), 4), "auc_min": round(float(m.min()), 4),
                 "auc_p10": round(float(m.quantile(0.1)), 4), "nonfinite_models": int((~finite).sum()),
                 "param_norm_mean": round(float(P[finite].norm(dim=1).mean()), 3),
                 "param_norm_max": round(float(P[finite].norm(dim=1).max()), 3),
                 "rejecting_clients": int((st.rejected > 0).sum()), "epochs_mean": round(sum(res.epochs_run.values()) / max(1, len(res.epochs_run)), 2)})
    if r % 3 == 2 or r == a.rounds - 1:
        print(json.dumps(rows[-1]))
