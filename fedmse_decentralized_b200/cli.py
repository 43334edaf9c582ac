"""Command-line entrypoint (``main.py`` at the repository root and the ``fedmse-b200`` console script).

Counterpart of the reference's ``src/main.py``: one invocation sweeps ``model_types x update_types x num_runs``
over the scenario JSON and writes the reference's checkpoint tree.  The reference's module-level constants
(``src/main.py:37-66``) are flags / ``FEDMSE_<NAME>`` environment variables here.
"""
import logging
import os

from . import ExperimentConfig
from .parallel import init_distributed, shutdown
from .trainer import run_experiment


def main(argv=None):
    cfg = ExperimentConfig.from_args(argv)
    logging.basicConfig(level=getattr(logging, cfg.log_level.upper(), logging.INFO),
                        format="%(asctime)s - %(levelname)s - %(message)s")
    log = logging.getLogger("fedmse")
    comm = None
    if int(os.environ.get("WORLD_SIZE", "1")) > 1:
        comm = init_distributed()
    if comm is None or comm.rank == 0:
        log.info("=" * 50)
        log.info("Training Parameters:")
        for k in ("num_runs", "num_rounds", "epoch", "lr_rate", "shrink_lambda", "network_size",
                  "num_participants", "batch_size", "data_seed", "no_Exp", "model_types", "update_types",
                  "backend", "aggregation", "topology", "reference_compat"):
            log.info("%s: %s", k, getattr(cfg, k))
        log.info("Total combinations to run: %d", len(cfg.model_types) * len(cfg.update_types) * cfg.num_runs)
        log.info("=" * 50)
    best = run_experiment(cfg, comm)
    shutdown()
    return best
