"""``ModelVerifier`` -- object API of the verification gate (reference:
``src/Trainer/model_verifier.py:13-99``).

Accept a received aggregated model iff the summed per-tensor L2 parameter change against the last
model seen is <= ``verification_threshold`` *and* the performance ``1/(1+MSE)`` on the verification
data did not drop by more than ``performance_threshold``; the first model is always accepted; the
history is refreshed even when the model is rejected (quirk Q10).  The math is ``ops.oracle``; the
device path of the same gate is ``verify_kernel`` (``csrc/merge_kernel.cu``).
"""
from __future__ import annotations

import logging
from typing import Dict, Tuple

import numpy as np
import torch

from ..layout import D_PAD, ParamLayout
from ..ops import oracle

log = logging.getLogger(__name__)


def _pad(x: torch.Tensor) -> torch.Tensor:
    x = torch.as_tensor(x, dtype=torch.float32)
    if x.shape[1] == D_PAD:
        return x
    out = torch.zeros(x.shape[0], D_PAD, dtype=torch.float32, device=x.device)
    out[:, :x.shape[1]] = x
    return out


class ModelVerifier:
    def __init__(self, verification_threshold: float = 3.0, performance_threshold: float = 0.002,
                 verification_method: str = "val"):
        self.verification_threshold = verification_threshold
        self.performance_threshold = performance_threshold
        self.verification_method = verification_method
        self.history: Dict = {}
        self.dev_dataset = None

    def set_dev_dataset(self, dev_dataset) -> None:
        self.dev_dataset = torch.Tensor(dev_dataset) if isinstance(dev_dataset, np.ndarray) else dev_dataset

    def _layout(self, state) -> ParamLayout:
        w1 = state["encoder.encoder_network.0.weight"]
        w2 = state["encoder.encoder_network.2.weight"]
        return ParamLayout(w1.shape[1], w1.shape[0], w2.shape[0])

    def _evaluate_model(self, model_state, verification_data, model_type="hybrid"):
        lay = self._layout(model_state)
        return oracle.performance(lay.pack(model_state), _pad(verification_data), lay)

    def _calculate_parameter_changes(self, old_state, new_state):
        lay = self._layout(new_state)
        return oracle.param_change(lay.pack(old_state), lay.pack(new_state), lay)

    def verify_model(self, client_id, new_model_state, validation_data, current_round,
                     model_type: str = "hybrid") -> Tuple[bool, float]:
        data = self.dev_dataset if self.verification_method == "dev" else validation_data
        if self.verification_method == "dev" and self.dev_dataset is None:
            log.error("Development dataset not set for verification")
            return False, 0.0
        snap = {k: v.detach().clone() for k, v in new_model_state.items()}
        perf = self._evaluate_model(snap, data, model_type)
        if client_id not in self.history:
            self.history[client_id] = {"model_state": snap, "performance": perf, "round": current_round}
            return True, 0.0
        prev = self.history[client_id]
        change = self._calculate_parameter_changes(prev["model_state"], snap)
        delta = perf - prev["performance"]
        self.history[client_id] = {"model_state": snap, "performance": perf, "round": current_round}
        log.info("Client %s - Param changes: %.10f, Performance change: %.10f", client_id, float(change), float(delta))
        ok = bool(change <= self.verification_threshold) and bool(delta >= -self.performance_threshold)
        return ok, delta
