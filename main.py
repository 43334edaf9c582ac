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

from fedmse_decentralized_b200.cli import main  # noqa: E402

if __name__ == "__main__":
    main()
