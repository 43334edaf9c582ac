"""Device-resident state of all clients of one rank.

In the reference every client is a Python object holding an ``nn.Module``, a
deep-copied "previous global model", a ``torch.optim.Adam`` instance, a
``ModelVerifier`` with its history dict, and a handful of counters
(``src/Trainer/client_trainer.py:47-95``, ``src/Trainer/model_verifier.py:13-19``).
Here each of those is one row of a ``[clients, P_PAD]`` slab (or one element of a
``[clients]`` vector), so a kernel can serve all clients of a GPU in one launch
and a peer GPU can read a client's parameters with a plain load from the slab.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional

import torch

from ..layout import ParamLayout


@dataclass
class FederationState:
    layout: ParamLayout
    params: torch.Tensor        # [C, PP]  current model of every client
    adam_m: torch.Tensor        # [C, PP]  Adam first moment (persists across rounds, quirk Q13)
    adam_v: torch.Tensor        # [C, PP]
    adam_t: torch.Tensor        # [C] int32 optimiser step count
    anchor: torch.Tensor        # [C, PP]  FedProx anchor = "previous_global_model"
    best: torch.Tensor          # [C, PP]  best-validation weights of the latest run() (= model.cpt)
    hist: torch.Tensor          # [C, PP]  verifier history: last aggregated model seen
    hist_perf: torch.Tensor     # [C]      1/(1+MSE) of that model
    hist_valid: torch.Tensor    # [C] int32 0 until the first verification
    agg_count: torch.Tensor     # [N_global] int32 aggregations per client (quota 3); replicated on every rank
    rejected: torch.Tensor      # [C] int32 consecutive rejected updates
    loss_log: torch.Tensor      # [C, E, 2] (train, valid) loss per epoch of the latest run()
    epochs_run: torch.Tensor    # [C] int32 epochs executed in the latest run()
    trained_rounds: torch.Tensor  # [C] int32 how many times the client trained

    @classmethod
    def create(cls, init_params: torch.Tensor, layout: ParamLayout, max_epochs: int,
               device=None, params_storage: Optional[torch.Tensor] = None,
               n_global: int = 0) -> "FederationState":
        """Build the state from an initial ``[C, PP]`` slab.

        ``params_storage`` lets the caller place the parameter slab in peer-mapped
        (symmetric) memory; the other slabs are private to the rank.
        """
        device = torch.device(device) if device is not None else init_params.device
        C, PP = init_params.shape
        assert PP == layout.padded
        f32 = dict(dtype=torch.float32, device=device)
        i32 = dict(dtype=torch.int32, device=device)
        if params_storage is not None:
            params = params_storage
            params.copy_(init_params)
        else:
            params = init_params.to(device).clone()
        return cls(
            layout=layout,
            params=params,
            adam_m=torch.zeros(C, PP, **f32),
            adam_v=torch.zeros(C, PP, **f32),
            adam_t=torch.zeros(C, **i32),
            anchor=params.clone(),
            best=params.clone(),
            hist=torch.zeros(C, PP, **f32),
            hist_perf=torch.zeros(C, **f32),
            hist_valid=torch.zeros(C, **i32),
            agg_count=torch.zeros(n_global or C, **i32),
            rejected=torch.zeros(C, **i32),
            loss_log=torch.zeros(C, max(1, max_epochs), 2, **f32),
            epochs_run=torch.zeros(C, **i32),
            trained_rounds=torch.zeros(C, **i32),
        )

    @property
    def n_clients(self) -> int:
        return self.params.shape[0]

    @property
    def device(self) -> torch.device:
        return self.params.device

    _FIELDS = ("params", "adam_m", "adam_v", "adam_t", "anchor", "best", "hist", "hist_perf",
               "hist_valid", "agg_count", "rejected", "loss_log", "epochs_run", "trained_rounds")

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Full resumable snapshot (the reference persists none of this)."""
        return {k: getattr(self, k).detach().cpu().clone() for k in self._FIELDS}

    def load_state_dict(self, sd: Dict[str, torch.Tensor]) -> None:
        for k in self._FIELDS:
            getattr(self, k).copy_(sd[k].to(self.device))

    def client_state_dict(self, c: int, which: str = "params"):
        return self.layout.unpack(getattr(self, which)[c])
