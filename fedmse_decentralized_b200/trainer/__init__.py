from .engine import FederationEngine, GlobalStop, RoundResult, make_backend
from .experiment import build_shards, prepare_data, run_experiment, set_seeds
from .state import FederationState

__all__ = ["FederationEngine", "GlobalStop", "RoundResult", "make_backend", "FederationState",
           "run_experiment", "prepare_data", "build_shards", "set_seeds"]
