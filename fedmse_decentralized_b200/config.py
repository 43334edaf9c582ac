"""Experiment configuration.

The reference is configured by editing module-level constants in
``src/main.py:37-66`` (``argparse`` is imported there but unused).  The same
names and defaults are kept here as dataclass fields; in addition every field
can be overridden from the command line (``--epoch 100``) or the environment
(``FEDMSE_EPOCH=100``), and constructor-only knobs of the reference
(``fedprox_mu``, thresholds, quotas, vote chunk) are surfaced.
"""
from __future__ import annotations

import argparse
import dataclasses
import os
from dataclasses import dataclass, field
from typing import List, Optional, Sequence


@dataclass
class ExperimentConfig:
    # ---- reference constants (src/main.py:37-66), checked-in values ----------
    num_participants: float = 0.5
    epoch: int = 5
    num_rounds: int = 3
    lr_rate: float = 1e-3
    shrink_lambda: float = 5
    network_size: int = 10
    data_seed: int = 1234
    no_Exp: Optional[str] = None          # derived when None (f-string of the hyper-parameters)
    checkpoint_dir: Optional[str] = None  # derived when None
    verification_method: str = "val"      # "val" | "dev"
    num_runs: int = 1
    batch_size: int = 12
    new_device: bool = True
    global_patience: int = 1
    metric: str = "AUC"                   # "AUC" | "classification"
    model_types: List[str] = field(default_factory=lambda: ["hybrid", "autoencoder"])
    update_types: List[str] = field(default_factory=lambda: ["avg", "fedprox", "mse_avg"])
    dim_features: int = 115
    scen_name: str = "FL-IoT"
    config_file: str = "Configuration/kitsune-iot-10clients.json"
    # ---- constructor defaults of the reference that never surfaced ---------
    hidden_neus: int = 27
    latent_dim: int = 7
    fedprox_mu: float = 0.001
    local_patience: Optional[int] = None  # reference passes global_patience (src/main.py:246)
    verification_threshold: float = 3.0
    performance_threshold: float = 0.002
    max_aggregation_threshold: int = 3
    max_rejected_updates: int = 3
    vote_chunk: int = 128
    # ---- new in this framework ---------------------------------------------
    backend: str = "auto"                 # "auto" | "cuda" | "oracle"
    reference_compat: bool = True         # reproduce the quirks of SURVEY.md §8 (Q2,Q4,Q5,Q6)
    aggregation: str = "elected"          # "elected" (reference) | "gossip"
    topology: str = "full"                # gossip only: "full" | "ring" | "random-k"
    topology_k: int = 4
    topology_seed: int = 0
    write_latents: bool = True            # SAE latent dumps (README.md:11; quirk Q19)
    save_client_models: bool = True       # model.cpt / training_tracking.pkl per client
    output_root: str = "."                # where Checkpoint/ is rooted (reference: CWD = src/)
    data_root: Optional[str] = None       # overrides the scenario's data_path
    synthetic_clients: int = 0            # >0: ignore config_file, generate this many clients
    synthetic_spec: str = "nbaiot"        # "nbaiot" | "kitsune"
    score_precision: str = "fp32"         # "fp32" (exact FFMA) | "tf32" | "tf32x3" | "bf16x3" (tcgen05 scoring forwards)
    fusion_rows: int = 512                # fusion_avg: dev rows used for the KDE / JS-distance weights
    checkpoint_every: int = 0             # write a resumable engine checkpoint every K rounds (0 = never; not in the reference)
    resume: bool = False                  # continue every (model, update, run) combination from its last checkpoint
    log_level: str = "INFO"

    def __post_init__(self):
        if self.no_Exp is None:
            self.no_Exp = (f"Exp_{self.epoch}epoch_{self.network_size}client_lr{self.lr_rate}"
                           f"_lamda{self.shrink_lambda}_ratio{self.num_participants * 100}")
        if self.checkpoint_dir is None:
            self.checkpoint_dir = f"Checkpoint/Results/Update/{self.network_size}/{self.no_Exp}"
        if self.local_patience is None:
            self.local_patience = self.global_patience
        if self.verification_method not in ("val", "dev"):
            raise ValueError("verification_method must be 'val' or 'dev'")
        for u in self.update_types:
            if u not in ("avg", "fedprox", "mse_avg", "fusion_avg"):
                raise ValueError(f"unknown update type {u!r}")
        for m in self.model_types:
            if m not in ("hybrid", "autoencoder"):
                raise ValueError(f"unknown model type {m!r}")

    # ------------------------------------------------------------------ CLI / env
    @classmethod
    def parser(cls) -> argparse.ArgumentParser:
        p = argparse.ArgumentParser(description="B200-native decentralized FedMSE")
        for f in dataclasses.fields(cls):
            name = "--" + f.name
            default = None
            if f.type in ("bool", bool):
                p.add_argument(name, type=lambda s: s.lower() in ("1", "true", "yes", "on"),
                               default=default, metavar="BOOL")
            elif f.type in ("List[str]",):
                p.add_argument(name, nargs="+", default=default)
            elif f.type in ("int", "Optional[int]"):
                p.add_argument(name, type=int, default=default)
            elif f.type in ("float",):
                p.add_argument(name, type=float, default=default)
            else:
                p.add_argument(name, type=str, default=default)
        return p

    @classmethod
    def from_args(cls, argv: Optional[Sequence[str]] = None, **overrides) -> "ExperimentConfig":
        ns = cls.parser().parse_args(argv)
        kw = {}
        for f in dataclasses.fields(cls):
            env = os.environ.get("FEDMSE_" + f.name.upper())
            if env is not None:
                kw[f.name] = _coerce(f.type, env)
            v = getattr(ns, f.name)
            if v is not None:
                kw[f.name] = v
        kw.update(overrides)
        return cls(**kw)

    def replace(self, **kw) -> "ExperimentConfig":
        d = dataclasses.asdict(self)
        d.update(kw)
        if "no_Exp" not in kw and any(k in kw for k in ("epoch", "network_size", "lr_rate",
                                                         "shrink_lambda", "num_participants")):
            d["no_Exp"] = None
            d["checkpoint_dir"] = None if "checkpoint_dir" not in kw else kw["checkpoint_dir"]
        return ExperimentConfig(**d)


def _coerce(tp, s: str):
    if tp in ("bool", bool):
        return s.lower() in ("1", "true", "yes", "on")
    if tp in ("int", "Optional[int]"):
        return int(s)
    if tp in ("float",):
        return float(s)
    if tp in ("List[str]",):
        return s.split(",")
    return s
