"""On-disk artefacts, byte-compatible with the reference's checkpoint tree.

* client model  ``Checkpoint/{N}/{exp}/{run}/ClientModel/{scen}/{model}/{update}/{device}/model.cpt``
  legacy (non-zip) ``torch.save`` of the 8-key ``state_dict``
  (``src/Trainer/client_trainer.py:337-350``);
* ``training_tracking.pkl`` next to it: pickled ``list[(train_loss, valid_loss)]``
  (``client_trainer.py:405``, ``:419``);
* round metrics JSON-lines, append mode (``src/main.py:342-355``);
* verification JSON-lines, append mode (``src/main.py:319-326``);
* ``training_summary.json`` (``src/main.py:390-399``);
* SAE latent dump pickle ``dict[round -> dict[device -> (latent, label)]]`` -- promised by the
  reference's README (``README.md:11``) and read by ``src/Visualization/
  latent_visualization.ipynb:59`` but never written by its ``main.py`` (quirk Q19).
"""
from __future__ import annotations

import json
import os
import pickle
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch


def client_model_dir(root: str, network_size: int, no_exp: str, run: int, scen_name: str,
                     model_type: str, update_type: str, device_name: str) -> str:
    return os.path.join(root, f"Checkpoint/{network_size}/{no_exp}/{run}/ClientModel", scen_name,
                        model_type, update_type, device_name)


def save_client_model(save_dir: str, state_dict: Dict[str, torch.Tensor]) -> str:
    os.makedirs(save_dir, exist_ok=True)
    path = os.path.join(save_dir, "model.cpt")
    torch.save(state_dict, path, _use_new_zipfile_serialization=False)
    return path


def save_tracking(save_dir: str, tracking: Sequence[Tuple[float, float]]) -> str:
    os.makedirs(save_dir, exist_ok=True)
    path = os.path.join(save_dir, "training_tracking.pkl")
    with open(path, "wb") as fh:
        pickle.dump([(float(a), float(b)) for a, b in tracking], fh)
    return path


def results_file(root: str, network_size: int, no_exp: str, run: int, metric: str, scen_name: str,
                 num_participants: float, model_type: str, update_type: str) -> str:
    d = os.path.join(root, f"Checkpoint/Results/Update/{network_size}/{no_exp}/Run_{run}/{metric}")
    return os.path.join(d, f"{scen_name}_{num_participants}_{model_type}_{update_type}_results.json")


def append_jsonl(path: str, record: Dict) -> None:
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "a") as fh:
        json.dump(record, fh)
        fh.write("\n")


def append_round_metrics(path: str, round_no: int, client_metrics: Sequence[float],
                         update_type: str, model_type: str) -> None:
    metrics = [float(m) for m in client_metrics]
    append_jsonl(path, {
        "round": round_no,
        "client_metrics": metrics,
        "update_type": update_type,
        "model_type": model_type,
        "global_loss": min(metrics) if metrics else float("inf"),
    })


def append_verification(path: str, round_no: int, results: List[Dict]) -> None:
    append_jsonl(path, {"round": round_no, "verification_results": results})


def write_summary(root: str, network_size: int, no_exp: str, best_metrics: Dict, metric: str,
                  num_runs: int) -> str:
    path = os.path.join(root, f"Checkpoint/Results/Update/{network_size}/{no_exp}/training_summary.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "w") as fh:
        json.dump({"best_metrics": best_metrics, "metric_type": metric, "num_runs": num_runs,
                   "network_size": network_size, "experiment_name": no_exp}, fh, indent=4)
    return path


def latent_file(root: str, network_size: int, no_exp: str, run: int, model_type: str,
                update_type: str) -> str:
    return os.path.join(root, f"Checkpoint/LatentData/{network_size}/{no_exp}/Run_{run}",
                        f"latent_{model_type}_{update_type}.pkl")


def write_latents(path: str, latents: Dict[int, Dict[str, Tuple[np.ndarray, np.ndarray]]]) -> None:
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb") as fh:
        pickle.dump(latents, fh)


def read_jsonl(path: str) -> List[Dict]:
    with open(path) as fh:
        return [json.loads(line) for line in fh if line.strip()]
