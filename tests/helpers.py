"""Shared test helpers: build matching engines on two backends."""
import copy
import random

import numpy as np
import torch

from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.data import ShardStore
from fedmse_decentralized_b200.data.shards import Segmented
from fedmse_decentralized_b200.trainer import FederationEngine


def make_cfg(tmp, **kw):
    base = dict(epoch=2, num_rounds=2, network_size=6, output_root=str(tmp), write_latents=True,
                save_client_models=True)
    base.update(kw)
    return ExperimentConfig(**base)


def make_shards(tiny_clients, device="cpu"):
    clients, dev = tiny_clients
    sh = ShardStore.build(clients, dev)
    return sh.to(device)


def make_engine(cfg, sh, model_type="hybrid", update_type="mse_avg", backend="oracle", seed=0, comm=None):
    torch.manual_seed(seed)
    random.seed(cfg.data_seed)
    np.random.seed(cfg.data_seed)
    return FederationEngine(cfg, sh, model_type, update_type, backend=backend, comm=comm)


def assert_close(a, b, rtol=1e-4, atol=1e-5, what=""):
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    assert not bool(bad.any()), (f"{what}: {int(bad.sum())}/{a.numel()} mismatched, max abs err "
                                 f"{float(err.max()):.3e}, max rel {float((err / (b.abs() + 1e-30)).max()):.3e}")
