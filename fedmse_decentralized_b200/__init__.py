"""B200-native decentralized federated learning for IoT intrusion detection (FedMSE, peer-to-peer).

Package map (reference layer -> here):

* ``models``    -- SAE / AE modules, centroid scorer            (reference ``src/Model``)
* ``data``      -- CSV ingestion, standardisation, device shards, synthetic traffic, partitioner
                                                                 (reference ``src/DataLoader``, notebook)
* ``ops``       -- fused device ops: sm_100a kernels + torch oracle (new)
* ``parallel``  -- process group, symmetric (peer-mapped) slabs, topologies (new)
* ``trainer``   -- round engine, experiment sweep, ClientTrainer / ModelVerifier API
                                                                 (reference ``src/Trainer``, ``src/main.py``)
* ``evaluator`` -- Evaluator API                                 (reference ``src/Evaluator``)
* ``utils``     -- similarity scores, artefact writers, timers  (reference ``src/Utils``)
"""
from .config import ExperimentConfig
from .layout import D_PAD, DEFAULT_LAYOUT, STATE_KEYS, ParamLayout

__version__ = "0.1.0"
__all__ = ["ExperimentConfig", "ParamLayout", "DEFAULT_LAYOUT", "STATE_KEYS", "D_PAD", "__version__"]
