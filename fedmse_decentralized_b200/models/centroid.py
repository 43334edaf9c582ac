"""Centroid one-class scorer (the "CEN" of SAE-CEN).

Reference behaviour (``src/Model/Centroid.py:6-39``): standardise the training
latents (population std, zero variance -> 1), which moves their centroid to the
origin; the anomaly score of a point is its Euclidean distance to the origin in
that standardised space; ``predict`` thresholds at a percentile of the training
distances.  The device path computes the same score inside the evaluation
kernel; this class is the host-side API and the oracle.
"""
from __future__ import annotations

import numpy as np


class CentroidBasedOneClassClassifier:
    def __init__(self, threshold: float = 0.5, metric: str = "euclidean", scale: str = "standard"):
        self.threshold = threshold
        self.metric = metric
        self.scale = scale
        self.mean_ = None
        self.scale_ = None
        self.centroid = None
        self.abs_threshold = None

    # -- helpers -----------------------------------------------------------
    def _standardise(self, X: np.ndarray) -> np.ndarray:
        return (np.asarray(X, dtype=np.float64) - self.mean_) / self.scale_

    def _dist(self, Xs: np.ndarray) -> np.ndarray:
        diff = Xs - self.centroid
        if self.metric == "euclidean":
            return np.sqrt(np.sum(diff * diff, axis=1))
        if self.metric in ("cityblock", "manhattan"):
            return np.sum(np.abs(diff), axis=1)
        if self.metric == "chebyshev":
            return np.max(np.abs(diff), axis=1)
        raise ValueError(f"unsupported metric {self.metric!r}")

    # -- API ------------------------------------------------------------------
    def fit(self, X) -> "CentroidBasedOneClassClassifier":
        X = np.asarray(X, dtype=np.float64)
        self.mean_ = X.mean(axis=0)
        std = X.std(axis=0)  # population std (ddof=0), as StandardScaler
        std[std == 0.0] = 1.0
        self.scale_ = std
        self.centroid = np.zeros((1, X.shape[1]))
        d = self.get_density(self._standardise(X), scale=False)
        self.abs_threshold = np.percentile(d, 100 * self.threshold)
        return self

    def get_density(self, X, scale: bool = True) -> np.ndarray:
        Xs = self._standardise(X) if scale else np.asarray(X, dtype=np.float64)
        return self._dist(Xs)

    def predict(self, X) -> np.ndarray:
        return self.get_density(X) > self.abs_threshold
