from .loader import IoTDataProccessor, IoTDataset, load_data
from .partition import dirichlet_partition, repartition, write_clients
from .scenario import load_scenario, prepare_from_raw, prepare_from_scenario
from .shards import PreparedClient, Segmented, ShardStore, pool_dev_set, prepare_client, split_sizes
from .synthetic import (KITSUNE_SPEC, NBAIOT_SPEC, ClientRaw, SyntheticSpec, make_clients,
                        write_scenario)

__all__ = [n for n in dir() if not n.startswith("_")]
