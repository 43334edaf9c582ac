from . import topology
from .world import Comm, block_partition, init_distributed, owner_of, shutdown

__all__ = ["Comm", "block_partition", "owner_of", "init_distributed", "shutdown", "topology"]
