"""The reference's import surface under one roof.

``from Model import Shrink_Autoencoder`` / ``from Trainer import ClientTrainer`` / ``from Evaluator import
Evaluator`` / ``from DataLoader import load_data, IoTDataProccessor, IoTDataset`` / ``from Utils import
similarity_score`` (reference ``src/*/__init__.py``; plus the ``GlobalAggregator`` whose import the fork commented out) become ``from fedmse_decentralized_b200.compat import ...``.
"""
from ..data.loader import IoTDataProccessor, IoTDataset, load_data
from ..evaluator import Evaluator
from ..models import Autoencoder, CentroidBasedOneClassClassifier, Shrink_Autoencoder
from ..trainer.client_trainer import ClientTrainer
from ..trainer.global_aggregator import GlobalAggregator
from ..trainer.model_verifier import ModelVerifier
from ..utils.similarity import js_divergence, kl_divergence, similarity_score

__all__ = ["Shrink_Autoencoder", "Autoencoder", "CentroidBasedOneClassClassifier", "ClientTrainer", "GlobalAggregator", "ModelVerifier",
           "Evaluator", "load_data", "IoTDataProccessor", "IoTDataset", "similarity_score", "kl_divergence",
           "js_divergence"]
