"""CSV ingestion and standardisation (host side, float64, done once at start-up).

API parity with the reference's ``DataLoader`` package
(``src/DataLoader/dataloader.py:22-80``): ``load_data(path)`` concatenates every
``*.csv`` of a directory (headerless), ``IoTDataProccessor`` wraps a scaler and
also emits 0/1 labels, ``IoTDataset`` is a map-style dataset yielding fp32 rows.

Raw N-BaIoT features reach 5.9e17, so standardisation stays in float64 on the
host exactly like sklearn's ``StandardScaler``; only the
standardised matrix is cast to fp32 and shipped to the device shard store.
"""
from __future__ import annotations

import logging
import os
from typing import Dict, Optional, Tuple

import numpy as np
import pandas as pd

log = logging.getLogger(__name__)


def load_data(path: str, header=None) -> pd.DataFrame:
    """Concatenate all CSV files of ``path`` (directory listing order, like the reference)."""
    frames = []
    for name in os.listdir(path):
        if ".csv" in name:
            full = os.path.join(path, name)
            log.debug("loading %s", full)
            frames.append(pd.read_csv(full, header=header))
    if not frames:
        raise FileNotFoundError(f"no .csv file under {path}")
    return pd.concat(frames, ignore_index=True)


class _Standard:
    """float64 standardiser with sklearn's semantics (population std, constant -> 1)."""

    def fit(self, X):
        from sklearn.preprocessing import StandardScaler  # same numerics as the reference

        self._sk = StandardScaler().fit(X)
        self.mean_, self.scale_, self.var_ = self._sk.mean_, self._sk.scale_, self._sk.var_
        return self

    def transform(self, X):
        return self._sk.transform(X)


class _MinMax:
    def fit(self, X):
        from sklearn.preprocessing import MinMaxScaler

        self._sk = MinMaxScaler((0, 1)).fit(X)
        self.mean_, self.scale_ = self._sk.data_min_, self._sk.data_range_
        return self

    def transform(self, X):
        return self._sk.transform(X)


class IoTDataProccessor:
    """Scaler wrapper that also labels rows (0 = normal, 1 = abnormal)."""

    def __init__(self, scaler: str = "standard"):
        if scaler == "standard":
            self.scaler = _Standard()
        elif scaler == "minmax":
            self.scaler = _MinMax()
        else:
            raise ValueError(f"unknown scaler {scaler!r}")

    def fit_transform(self, dataframe) -> Tuple[np.ndarray, np.ndarray]:
        self.scaler.fit(dataframe)
        return self.transform(dataframe, type="normal")

    def transform(self, dataframe, type: str = "normal") -> Tuple[np.ndarray, np.ndarray]:
        if len(dataframe) == 0:      # e.g. test data of a client owned by another rank
            data = np.zeros((0, np.shape(dataframe)[1] if np.ndim(dataframe) == 2 else 0))
        else:
            data = self.scaler.transform(dataframe)
        label = np.zeros(len(dataframe), dtype=np.int64) if type == "normal" \
            else np.ones(len(dataframe), dtype=np.int64)
        return data, label

    def get_metadata(self) -> Dict[str, np.ndarray]:
        return {"mean": self.scaler.mean_, "std": self.scaler.scale_}


class IoTDataset:
    """Map-style dataset over a (data, label) pair; rows are cast to fp32 on access."""

    def __init__(self, data, label):
        self.data = np.asarray(data)
        self.label = np.asarray(label)

    def __len__(self) -> int:
        return len(self.data)

    def __getitem__(self, idx):
        return self.data[idx].astype(np.float32), self.label[idx].astype(np.float32)

    @property
    def input_dim_(self) -> int:
        return self.data.shape[1]
