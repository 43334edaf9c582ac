from .engine import FederationEngine, GlobalStop, RoundResult, make_backend
from .experiment import build_shards, prepare_data, run_experiment, set_seeds
from .client_trainer import ClientTrainer
from .global_aggregator import GlobalAggregator
from .model_verifier import ModelVerifier
from .state import FederationState

__all__ = ["FederationEngine", "GlobalStop", "RoundResult", "make_backend", "FederationState",
           "ClientTrainer", "GlobalAggregator", "ModelVerifier", "run_experiment", "prepare_data", "build_shards", "set_seeds"]
