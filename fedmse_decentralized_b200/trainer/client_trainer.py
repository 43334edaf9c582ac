"""``ClientTrainer`` -- the reference's per-client object API (``src/Trainer/client_trainer.py:26-419``)
on top of this framework's ops.

A user of the reference who builds trainers by hand (``ClientTrainer(model=..., ...)``,
``connect_to_peers``, ``run``, ``vote_for_aggregator``, ``aggregate_models``, ``receive_model``,
``update_from_peers``, ``broadcast_model``, ``request_aggregation``) can keep that code.  Each
trainer keeps its model as one packed fp32 row; ``run`` executes on the GPU through the persistent
training kernel when CUDA is available (one-client launch) and through the torch oracle otherwise.
For throughput use ``FederationEngine``, which drives all clients of a GPU per launch.

Differences from the reference, on purpose: ``receive_model`` stores a *copy* (the reference stores a
live reference, which is what makes its ``fed_mse_avg`` overwrite the aggregator's own slot, SURVEY.md
Q6); pass ``reference_compat=True`` to ``aggregate_models`` to reproduce that effect by weight remap.
"""
from __future__ import annotations

import copy
import logging
import os
import pickle
from typing import Dict, List, Optional

import numpy as np
import torch

from ..layout import D_PAD, ParamLayout
from ..ops import oracle
from ..ops.oracle import TrainHP
from ..utils import artefacts
from .model_verifier import ModelVerifier, _pad

log = logging.getLogger(__name__)


def _rows_of(loader_or_matrix) -> torch.Tensor:
    """Accept a DataLoader over ``IoTDataset`` (reference habit), a dataset, or a matrix."""
    obj = loader_or_matrix
    if hasattr(obj, "dataset"):
        obj = obj.dataset
    if hasattr(obj, "data"):
        obj = obj.data
    return _pad(torch.as_tensor(np.asarray(obj), dtype=torch.float32))


class ClientTrainer:
    def __init__(self, model=None, loss_function=None, optimizer=None, epoch=10, batch_size=100, lr_rate=1e-3,
                 update_type="avg", patience=5, save_dir="Checkpoint/ClientModel/", fedprox_mu=0.001,
                 client_id=None, model_type="hybrid", verification_method="val", verification_threshold=3.0,
                 performance_threshold=0.002, device=None) -> None:
        if model is None:
            raise ValueError(f"[Client {client_id}] Have to indicate the model to train.")
        os.makedirs(save_dir, exist_ok=True)
        self.device = torch.device(device) if device else torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.model = model
        self.layout: ParamLayout = model.layout
        self.params = model.pack().to(self.device)
        self.previous_global_model = copy.deepcopy(model)
        self.anchor = self.params.clone()
        self.adam_m = torch.zeros_like(self.params)
        self.adam_v = torch.zeros_like(self.params)
        self.adam_t = 0
        self.lr_rate, self.epoch, self.batch_size, self.patience = lr_rate, epoch, batch_size, patience
        self.save_dir, self.update_type, self.fedprox_mu, self.client_id = save_dir, update_type, fedprox_mu, client_id
        self.model_type = model_type
        self.peers: List["ClientTrainer"] = []
        self.aggregation_count, self.max_aggregation_threshold = 0, 3
        self.mse_score, self.votes_received, self.current_round = float("inf"), 0, -1
        self.has_aggregated_this_round = False
        self.received_models: Dict = {}
        self.validation_data = None
        self.dev_dataset = None
        self.verifier = ModelVerifier(verification_threshold, performance_threshold, verification_method)
        self.rejected_updates, self.max_rejected_updates = 0, 3
        self.training_tracking: List = []

    # -- state bridging ------------------------------------------------------
    def state_dict(self):
        return self.layout.unpack(self.params)

    def _sync_module(self) -> None:
        self.model.load_state_dict(self.state_dict())

    def _load_state(self, state) -> None:
        self.params.copy_(self.layout.pack(state).to(self.device))
        self._sync_module()

    def create_dev_dataset(self, dataset) -> None:
        d = dataset["dataset"]
        self.dev_dataset = torch.Tensor(d) if isinstance(d, np.ndarray) else d
        self.verifier.set_dev_dataset(self.dev_dataset)

    # -- local training --------------------------------------------------------
    def _hp(self) -> TrainHP:
        lam = float(getattr(self.model, "shrink_lambda", 0.0)) if self.model_type == "hybrid" else 0.0
        return TrainHP(self.epoch, self.batch_size, self.lr_rate, lam,
                       self.fedprox_mu if self.update_type == "fedprox" else 0.0, self.patience)

    def run(self, train_loader, valid_loader=None):
        xt = _rows_of(train_loader).to(self.device)
        xv = _rows_of(valid_loader).to(self.device) if valid_loader is not None else xt[:0]
        hp = self._hp()
        use_kernel = self.device.type == "cuda" and (self.layout.dim, self.layout.hidden, self.layout.latent) == \
            (115, 27, 7) and hp.batch_size <= 32
        if use_kernel:
            best, log_ = self._run_kernel(xt, xv, hp)
        else:
            self.adam_t, best, log_ = oracle.train_client(self.params, self.adam_m, self.adam_v, self.adam_t,
                                                          self.anchor, xt, xv, self.layout, hp)
        if valid_loader is None:
            log_ = []                     # the reference appends to training_tracking only when it validated (:405)
        self.training_tracking = log_
        if valid_loader is not None:      # the reference saves model.cpt only inside its validation block (:387-411)
            artefacts.save_client_model(self.save_dir, self.layout.unpack(best))
        artefacts.save_tracking(self.save_dir, log_)
        self._sync_module()
        return log_

    def _run_kernel(self, xt, xv, hp):
        import ctypes as C
        from ..ops import _lib
        lib = _lib.load()
        dev = self.device
        i32 = dict(dtype=torch.int32, device=dev)
        p, m, v = self.params.view(1, -1), self.adam_m.view(1, -1), self.adam_v.view(1, -1)
        t = torch.tensor([self.adam_t], **i32)
        best = p.clone()
        logbuf = torch.zeros(1, max(1, hp.epochs), 2, device=dev)
        ep, tr = torch.zeros(1, **i32), torch.zeros(1, **i32)
        toff, voff = torch.tensor([0, xt.shape[0]], **i32), torch.tensor([0, xv.shape[0]], **i32)
        sel = torch.zeros(1, **i32)
        a = _lib.FmTrainArgs(p.data_ptr(), m.data_ptr(), v.data_ptr(), t.data_ptr(), self.anchor.data_ptr(),
                             best.data_ptr(), logbuf.data_ptr(), ep.data_ptr(), tr.data_ptr(), xt.data_ptr(),
                             toff.data_ptr(), xv.data_ptr(), voff.data_ptr(), sel.data_ptr(), 1, 0, hp.epochs,
                             hp.batch_size, hp.patience, logbuf.shape[1], hp.lr, hp.shrink_lambda, hp.fedprox_mu, 0)
        _lib.check(lib.fm_train_clients(C.byref(a), torch.cuda.current_stream(dev).cuda_stream), "train_clients")
        self.adam_t = int(t.item())
        n = int(ep.item())
        return best[0], [(float(logbuf[0, e, 0]), float(logbuf[0, e, 1])) for e in range(n)]

    # -- peer protocol -----------------------------------------------------------
    def connect_to_peers(self, peers) -> None:
        self.peers = list(peers)

    def broadcast_model(self) -> None:
        state = self.state_dict()
        for peer in self.peers:
            peer.receive_model(self, state)

    def receive_model(self, sender, model_state) -> None:
        self.received_models[sender] = {k: v.detach().clone() for k, v in model_state.items()}

    def calculate_mse_score(self, validation_data) -> float:
        V = _pad(torch.as_tensor(validation_data, dtype=torch.float32)).to(self.device)
        score = float(oracle.chunked_mse(self.params, oracle.vote_normalise(V, self.layout.dim), self.layout, 128))
        self.mse_score = score * (1.0 + (torch.rand(1).item() - 0.5) * 0.0002)
        return self.mse_score

    def vote_for_aggregator(self, clients, validation_data, current_round):
        for c in clients:
            c.current_round, c.has_aggregated_this_round = current_round, False
        scores = [(c, c.calculate_mse_score(validation_data)) for c in clients if c is not self]
        for c, _ in sorted(scores, key=lambda cs: cs[1]):
            if c.aggregation_count < c.max_aggregation_threshold:
                c.votes_received += 1
                return c
        return None

    def _merge(self, states: List[Dict], agg_pos: int = -1, reference_compat: bool = False) -> Dict:
        rows = torch.stack([self.layout.pack(s).to(self.device) for s in states])
        n = len(states)
        mse = None
        if self.update_type == "mse_avg":
            dev = _pad(torch.as_tensor(self.dev_dataset, dtype=torch.float32)).to(self.device)
            mse = oracle.mse_scores(rows, range(n), dev, self.layout, 0)
        w = oracle.merge_weights(self.update_type, mse, n, agg_pos, reference_compat, self.device)
        return self.layout.unpack(oracle.weighted_reduce(rows, range(n), w))

    def fed_avg(self, local_models):
        tot = sum(w for _, w in local_models)
        rows = torch.stack([self.layout.pack(s).to(self.device) for s, _ in local_models])
        w = torch.tensor([w / tot for _, w in local_models], dtype=torch.float32, device=self.device)
        return self.layout.unpack(oracle.weighted_reduce(rows, range(len(local_models)), w))

    def fedprox(self, local_models):
        return self.fed_avg(local_models)

    def fed_mse_avg(self, local_models):
        keep, self.update_type = self.update_type, "mse_avg"
        try:
            return self._merge([s for s, _ in local_models])
        finally:
            self.update_type = keep

    def aggregate_models(self, clients, validation_data, current_round, reference_compat: bool = False):
        if self.aggregation_count >= self.max_aggregation_threshold or self.has_aggregated_this_round:
            log.warning("[Client %s] This client cannot perform aggregation in this round", self.client_id)
            return None
        if self.update_type == "mse_avg":
            for c in clients:                     # the reference scores every client here too (quirk Q7)
                c.calculate_mse_score(validation_data)
        if self.update_type not in ("avg", "fedprox", "mse_avg"):
            log.error("[Client %s] Unknown update type: %s", self.client_id, self.update_type)
            return None
        pos = next((k for k, c in enumerate(clients) if c is self), -1)
        merged = self._merge([c.state_dict() for c in clients], pos, reference_compat)
        self.aggregation_count += 1
        self.has_aggregated_this_round = True
        self._load_state(merged)
        return merged

    def request_aggregation(self):
        if self.aggregation_count >= self.max_aggregation_threshold or not self.received_models:
            return None
        if self.update_type not in ("avg", "fedprox", "mse_avg"):
            return None
        merged = self._merge(list(self.received_models.values()))
        self.aggregation_count += 1
        self.has_aggregated_this_round = True
        return merged

    def update_from_peers(self) -> None:
        if not self.received_models:
            return
        state = list(self.received_models.values())[0]
        ok, delta = self.verifier.verify_model(self.client_id, state, self.validation_data, self.current_round,
                                               self.model_type)
        if ok:
            self._load_state(state)
            self.previous_global_model = copy.deepcopy(self.model)
            self.anchor.copy_(self.params)
            self.rejected_updates = 0
        else:
            self.rejected_updates += 1
            if self.rejected_updates >= self.max_rejected_updates:
                log.error("[Client %s] Too many rejected updates. Possible attack detected.", self.client_id)
        self.received_models.clear()

    def save_model(self) -> None:
        artefacts.save_client_model(self.save_dir, self.state_dict())

    def save_tracking_information(self) -> None:
        artefacts.save_tracking(self.save_dir, self.training_tracking)
