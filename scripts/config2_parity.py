#!/usr/bin/env python
"""BASELINE.json config 2 on the prepared N-BaIoT IID split, CUDA backend against the unmodified reference, same GPU.

  python scripts/config2_parity.py --data_root data_cache/N-BaIoT/IID-10-Client_Data [--rounds 20 --epochs 100]
                                   [--skip-reference]

Both arms get the same raw CSV rows, the same seeds and the same shuffles (the reference's pandas `sample` calls
on numpy's global generator, re-seeded identically), so the per-client AUC of every round can be compared directly;
the bound BASELINE.json names is 0.005.  Prints one JSON line.  (The prepared split is 70 MB of CSV that ships with
the reference checkout; copy it under data_cache/ -- git-ignored -- so that it travels to the GPU box.)
"""
import argparse
import json
import logging
import os
import random
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "baseline"))

import numpy as np
import pandas as pd
import torch


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--data_root", default=os.path.join(ROOT, "data_cache/N-BaIoT/IID-10-Client_Data"))
    ap.add_argument("--rounds", type=int, default=20)
    ap.add_argument("--epochs", type=int, default=100)
    ap.add_argument("--lr", type=float, default=1e-5)
    ap.add_argument("--shrink_lambda", type=float, default=10)
    ap.add_argument("--update", default="mse_avg")
    ap.add_argument("--score-precision", default="fp16x3")
    ap.add_argument("--skip-reference", action="store_true")
    a = ap.parse_args()
    from fedmse_decentralized_b200 import ExperimentConfig
    from fedmse_decentralized_b200.data import ShardStore, load_data, load_scenario, pool_dev_set, prepare_client
    from fedmse_decentralized_b200.data.synthetic import ClientRaw
    from fedmse_decentralized_b200.trainer import FederationEngine
    dev = torch.device("cuda", 0) if torch.cuda.is_available() else torch.device("cpu")
    scen = load_scenario(os.path.join(ROOT, "configs/scen2-nba-iot-10clients.json"))
    raw = [ClientRaw(d["name"], load_data(os.path.join(a.data_root, d["normal_data_path"])).to_numpy(),
                     load_data(os.path.join(a.data_root, d["abnormal_data_path"])).to_numpy(),
                     load_data(os.path.join(a.data_root, d["test_normal_data_path"])).to_numpy())
           for d in scen["devices_list"]]
    n = len(raw)
    out = {"config": 2, "scenario": "scen2-nba-iot-10clients (prepared N-BaIoT IID split, real data)", "clients": n,
           "rounds": a.rounds, "epochs": a.epochs, "lr": a.lr, "shrink_lambda": a.shrink_lambda, "update": a.update,
           "device": str(dev), "score_precision": a.score_precision}
    ref_metrics = None
    import ref_harness
    if not a.skip_reference:
        why = ref_harness.available()
        if why:
            out["reference"] = why
        else:
            logging.disable(logging.CRITICAL)
            cfgd = {"model_type": "hybrid", "update_type": a.update, "epoch": a.epochs, "batch_size": 12, "lr_rate": a.lr,
                    "shrink_lambda": a.shrink_lambda, "patience": 1, "num_participants": 0.5, "seed": 1234}
            random.seed(1234); np.random.seed(1234); torch.manual_seed(0)
            fed = ref_harness.ReferenceFederation(raw, range(n), cfgd, 1, 0, dev)
            random.seed(1234); torch.manual_seed(0)
            t0 = time.perf_counter()
            ref_metrics = [fed.run_round(r) for r in range(a.rounds)]
            if dev.type == "cuda":
                torch.cuda.synchronize()
            out["reference_wall_s"] = round(time.perf_counter() - t0, 2)
            fed.close()
            logging.disable(logging.NOTSET)
    # ours: the same shuffles as the harness (per-client re-seeded numpy generator behind DataFrame.sample)
    pcs = []
    for i, cl in enumerate(raw):
        np.random.seed(ref_harness.client_seed(1234, i))
        normal = pd.DataFrame(cl.normal).sample(frac=1).reset_index(drop=True).to_numpy()
        abnormal = pd.DataFrame(cl.abnormal).sample(frac=1).reset_index(drop=True).to_numpy()
        pcs.append(prepare_client(cl.name, normal, abnormal, cl.test_normal, True))
    np.random.seed(ref_harness.dev_seed(1234))
    devset = pool_dev_set(pcs, sampler=lambda x, k: pd.DataFrame(x).sample(n=k).to_numpy())
    random.seed(1234); torch.manual_seed(0)
    cfg = ExperimentConfig(epoch=a.epochs, num_rounds=a.rounds, lr_rate=a.lr, shrink_lambda=a.shrink_lambda,
                           network_size=n, backend="cuda" if dev.type == "cuda" else "oracle", reference_compat=True,
                           write_latents=False, save_client_models=False, score_precision=a.score_precision)
    eng = FederationEngine(cfg, ShardStore.build(pcs, devset).to(dev), "hybrid", a.update)
    random.seed(1234); torch.manual_seed(0)
    t0 = time.perf_counter()
    ours = [eng.run_round(r) for r in range(a.rounds)]
    if dev.type == "cuda":
        torch.cuda.synchronize()
    out["ours_wall_s"] = round(time.perf_counter() - t0, 3)
    om = [r.client_metrics for r in ours]
    out["ours_auc_mean_per_round"] = [round(float(np.mean(m)), 6) for m in om]
    out["ours_auc_last_round"] = [round(float(x), 6) for x in om[-1]]
    out["ours_epochs_last_round"] = sorted(ours[-1].epochs_run.values())
    if ref_metrics is not None:
        out["reference_auc_mean_per_round"] = [round(float(np.mean(m)), 6) for m in ref_metrics]
        out["reference_auc_last_round"] = [round(float(x), 6) for x in ref_metrics[-1]]
        d = [float(np.max(np.abs(np.array(x) - np.array(y)))) for x, y in zip(om, ref_metrics)]
        out["max_abs_auc_diff_per_round"] = [round(x, 6) for x in d]
        out["max_abs_auc_diff"] = round(max(d), 6)
        out["within_0.005"] = bool(max(d) < 5e-3)
        out["speedup_wall"] = round(out["reference_wall_s"] / out["ours_wall_s"], 1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
