from . import oracle
from .backend import MergeSpec, OracleBackend, RoundControl
from .oracle import TrainHP

__all__ = ["oracle", "OracleBackend", "RoundControl", "MergeSpec", "TrainHP"]
