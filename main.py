#!/usr/bin/env python
"""Experiment entrypoint -- the counterpart of the reference's ``src/main.py``.

The reference is configured by editing constants at the top of its ``main.py``; the same names are
flags here (``--num_participants 0.5 --epoch 100 --num_rounds 20 --lr_rate 1e-5 --shrink_lambda 10
--config_file configs/scen2-nba-iot-10clients.json``) and ``FEDMSE_<NAME>`` environment variables,
with the reference's checked-in values as defaults.  One invocation sweeps
``model_types x update_types x num_runs`` and writes the reference's checkpoint tree under
``--output_root``.  Multi-GPU: ``torchrun --nproc-per-node N main.py ...`` (clients are
block-partitioned over ranks).
"""
import logging
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from fedmse_decentralized_b200 import ExperimentConfig  # noqa: E402
from fedmse_decentralized_b200.parallel import init_distributed, shutdown  # noqa: E402
from fedmse_decentralized_b200.trainer import run_experiment  # noqa: E402


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


if __name__ == "__main__":
    main()
