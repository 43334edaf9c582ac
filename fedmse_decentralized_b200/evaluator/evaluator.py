"""``Evaluator`` -- object API of the anomaly-scoring stage (reference:
``src/Evaluator/evaluator.py:14-130``).

``evaluate(test_loader, train_loader)`` returns the AUC for ``model_type="autoencoder"`` and
``(AUC, test_latent, test_label)`` for ``"hybrid"`` like the reference; ``metric="classification"``
returns F1 at the fixed 0.5 threshold.  The reference's ``metric == "time"`` branch is broken
(appends to a numpy array); here it times the scoring pass and returns seconds.  Scoring runs on the
device of the model parameters through ``ops.oracle`` (the batched device path for many clients is
``FederationEngine`` -> ``score_rows_kernel`` + ``auc_kernel``).
"""
from __future__ import annotations

import time

import numpy as np
import torch

from ..ops import oracle
from ..trainer.client_trainer import _rows_of


class Evaluator:
    def __init__(self, model, model_type: str = "autoencoder", metric: str = "AUC") -> None:
        self.model, self.model_type, self.metric = model, model_type, metric
        self.device = torch.device("cuda" if torch.cuda.is_available() else "cpu")

    # helpers kept for API parity ---------------------------------------------------
    def calculate_auc(self, y_true, score) -> float:
        return oracle.auc_from_scores(torch.as_tensor(np.asarray(score)), torch.as_tensor(np.asarray(y_true)))

    def score_to_label(self, score, threshold):
        return np.where(np.array(score) > threshold, 1, 0)

    def calculate_classification_metric(self, y_true, score, threshold=0.5):
        return oracle.classification_metrics(torch.as_tensor(np.asarray(score)), torch.as_tensor(np.asarray(y_true)),
                                             threshold)

    def calculate_f1_score(self, y_true, y_pred):
        return oracle.classification_metrics(torch.as_tensor(np.asarray(y_pred, dtype=np.float32)),
                                             torch.as_tensor(np.asarray(y_true)))[0]

    def _labels(self, loader):
        ds = loader.dataset if hasattr(loader, "dataset") else loader
        if hasattr(ds, "datasets"):     # ConcatDataset of IoTDataset
            return np.concatenate([np.asarray(d.label) for d in ds.datasets]).astype(np.float32), \
                np.concatenate([np.asarray(d.data) for d in ds.datasets])
        return np.asarray(ds.label).astype(np.float32), np.asarray(ds.data)

    def evaluate(self, test_loader, train_loader=None):
        lay = self.model.layout
        row = self.model.pack().to(self.device)
        y, xt = self._labels(test_loader)
        x_test = _rows_of(xt).to(self.device)
        x_train = _rows_of(train_loader).to(self.device) if train_loader is not None else None
        t0 = time.time()
        with torch.no_grad():
            score, z = oracle.anomaly_scores(row, x_test, x_train, lay, self.model_type)
        if self.metric == "time":
            if self.device.type == "cuda":
                torch.cuda.synchronize()
            return time.time() - t0
        score = torch.nan_to_num(score)
        yt = torch.from_numpy(y)
        if self.metric == "classification":
            val = oracle.classification_metrics(score, yt)[0]
        else:
            val = oracle.auc_from_scores(score, yt)
        if self.model_type == "hybrid":
            return val, z.detach().cpu().numpy(), y
        return val

    def visualize(self):
        pass
