from .autoencoder import (Autoencoder, Decoder, Encoder, Shrink_Autoencoder, build_model,
                          init_packed_models, state_dict_from_packed)
from .centroid import CentroidBasedOneClassClassifier

__all__ = ["Autoencoder", "Shrink_Autoencoder", "Encoder", "Decoder", "build_model",
           "init_packed_models", "state_dict_from_packed", "CentroidBasedOneClassClassifier"]
