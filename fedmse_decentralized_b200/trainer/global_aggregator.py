"""``GlobalAggregator`` -- the central-server aggregator of the original FedMSE design.

The decentralized fork deleted its source; what the reference still ships is the compiled module
(``src/Trainer/__pycache__/global_aggregator.cpython-313.pyc``, import commented out at ``src/main.py:26`` and
``src/Trainer/__init__.py:10``) whose name table lists ``select_aggregator``, ``update``, ``create_dev_dataset``,
``fusion_avg``, ``fed_mse_avg``, ``fed_avg`` and ``fedprox``.  This class restores that surface on
top of the framework's ops, so code written against the centralized variant (one server holding the dev set, clients
sending their models in, the merged model going back out) keeps working:

* ``fed_avg`` / ``fedprox`` / ``fed_mse_avg`` are the same arithmetic the fork kept inside ``ClientTrainer``
  (``src/Trainer/client_trainer.py:107-130``): sample-count weights, and inverse dev-set MSE weights.
* ``fusion_avg`` is the rule whose only surviving piece is its weight function (``src/Utils/utils.py:10-24``): inverse
  Jensen-Shannon distance between the KDE self-densities of the dev rows and of each client's reconstruction of them
  (``utils.similarity.fusion_distances``; the engine's ``update_type="fusion_avg"`` runs it in ``csrc/similarity_kernel.cu``).
* ``update`` merges the given clients with the configured rule, keeps the result as the global model and hands it to
  every client (no verification gate: that is the decentralized fork's addition, ``ModelVerifier``).
* ``select_aggregator`` applies the fork's election rule (lowest jittered validation MSE among clients under their
  aggregation quota, ``client_trainer.py:249-285``) for callers that rotate the server role.
"""
from __future__ import annotations

import copy
import logging
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..layout import ParamLayout
from ..ops import oracle
from ..utils.similarity import fusion_distances
from .model_verifier import _pad

log = logging.getLogger(__name__)

UPDATE_TYPES = ("avg", "fedprox", "mse_avg", "fusion_avg")


class GlobalAggregator:
    def __init__(self, model=None, update_type: str = "avg", model_type: str = "hybrid", device=None,
                 fusion_rows: int = 512) -> None:
        if model is None:
            raise ValueError("GlobalAggregator needs the global model")
        if update_type not in UPDATE_TYPES:
            raise ValueError(f"unknown update type {update_type!r}; one of {UPDATE_TYPES}")
        self.device = torch.device(device) if device else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.model = model
        self.layout: ParamLayout = model.layout
        self.params = model.pack().to(self.device)
        self.update_type, self.model_type, self.fusion_rows = update_type, model_type, int(fusion_rows)
        self.dev_dataset: Optional[torch.Tensor] = None
        self.current_round = -1
        self.last_weights: Optional[torch.Tensor] = None      # merge weights of the last update, for inspection

    # -- state ---------------------------------------------------------------------------------------------------
    def state_dict(self) -> Dict[str, torch.Tensor]:
        return self.layout.unpack(self.params)

    def _load_state(self, state) -> None:
        self.params.copy_(self.layout.pack(state).to(self.device))
        self.model.load_state_dict(self.state_dict())

    def create_dev_dataset(self, dataset) -> None:
        d = dataset["dataset"] if isinstance(dataset, dict) else dataset
        self.dev_dataset = torch.as_tensor(np.asarray(d), dtype=torch.float32)

    def _dev(self) -> torch.Tensor:
        if self.dev_dataset is None:
            raise RuntimeError("create_dev_dataset() first: mse_avg and fusion_avg weigh the clients on the dev set")
        return _pad(self.dev_dataset).to(self.device)

    # -- the four merge rules (each takes [(state_dict, n_samples), ...] like the reference's) -------------------------
    def _rows(self, local_models: Sequence[Tuple[Dict, float]]) -> torch.Tensor:
        return torch.stack([self.layout.pack(s).to(self.device) for s, _ in local_models])

    def _reduce(self, rows: torch.Tensor, w: torch.Tensor) -> Dict[str, torch.Tensor]:
        self.last_weights = w.detach().cpu()
        return self.layout.unpack(oracle.weighted_reduce(rows, range(rows.shape[0]), w.to(self.device)))

    def fed_avg(self, local_models):
        tot = float(sum(n for _, n in local_models))
        w = torch.tensor([n / tot for _, n in local_models], dtype=torch.float32)
        return self._reduce(self._rows(local_models), w)

    def fedprox(self, local_models):
        # the proximal term acts in the clients' local objective; the server-side merge is FedAvg's
        return self.fed_avg(local_models)

    def fed_mse_avg(self, local_models):
        rows = self._rows(local_models)
        mse = oracle.mse_scores(rows, range(rows.shape[0]), self._dev(), self.layout, 0)
        w = oracle.merge_weights("mse_avg", mse, rows.shape[0], -1, False, self.device)
        return self._reduce(rows, w)

    def fusion_avg(self, local_models):
        rows = self._rows(local_models)
        d = fusion_distances(rows, self._dev(), self.layout, self.fusion_rows)
        inv = 1.0 / d
        return self._reduce(rows, inv / inv.sum())

    # -- server round ------------------------------------------------------------------------------------------------
    def select_aggregator(self, clients, validation_data, current_round: int = 0):
        """The fork's election among ``clients`` (they need ``calculate_mse_score`` / ``aggregation_count``)."""
        self.current_round = current_round
        scored = sorted(((c.calculate_mse_score(validation_data), k) for k, c in enumerate(clients)))
        for _, k in scored:
            c = clients[k]
            if c.aggregation_count < c.max_aggregation_threshold:
                c.votes_received = getattr(c, "votes_received", 0) + 1
                return c
        return None

    def update(self, clients, n_samples: Optional[Sequence[float]] = None, receivers=None, current_round: int = 0):
        """Merge ``clients`` (objects with ``state_dict()``) by ``update_type``; the result becomes the global model and
        is loaded into ``receivers`` (default: the same clients).  Returns the merged ``state_dict``."""
        if not clients:
            return None
        self.current_round = current_round
        n_samples = list(n_samples) if n_samples is not None else [1.0] * len(clients)
        local = [(c.state_dict(), float(n)) for c, n in zip(clients, n_samples)]
        rule = {"avg": self.fed_avg, "fedprox": self.fedprox, "mse_avg": self.fed_mse_avg,
                "fusion_avg": self.fusion_avg}[self.update_type]
        merged = rule(local)
        self._load_state(merged)
        for c in (clients if receivers is None else receivers):
            if hasattr(c, "_load_state"):                         # this framework's ClientTrainer
                c._load_state(merged)
                if hasattr(c, "anchor"):
                    c.anchor.copy_(c.params)
                    c.previous_global_model = copy.deepcopy(c.model)
            else:                                                 # anything holding a torch module
                c.model.load_state_dict(merged)
        return merged
