import os

import pytest

from fedmse_decentralized_b200 import ExperimentConfig


def test_defaults_are_the_reference_constants():
    c = ExperimentConfig()
    assert (c.num_participants, c.epoch, c.num_rounds, c.lr_rate, c.shrink_lambda) == (0.5, 5, 3, 1e-3, 5)
    assert (c.network_size, c.data_seed, c.batch_size, c.num_runs, c.global_patience) == (10, 1234, 12, 1, 1)
    assert c.model_types == ["hybrid", "autoencoder"] and c.update_types == ["avg", "fedprox", "mse_avg"]
    assert c.verification_method == "val" and c.metric == "AUC" and c.scen_name == "FL-IoT"
    assert (c.fedprox_mu, c.verification_threshold, c.performance_threshold) == (0.001, 3.0, 0.002)
    assert c.checkpoint_dir == f"Checkpoint/Results/Update/10/{c.no_Exp}" and c.local_patience == 1


def test_cli_and_env_overrides(monkeypatch):
    monkeypatch.setenv("FEDMSE_EPOCH", "100")
    monkeypatch.setenv("FEDMSE_LR_RATE", "1e-5")
    c = ExperimentConfig.from_args(["--num_rounds", "20", "--shrink_lambda", "10", "--update_types", "mse_avg",
                                    "--reference_compat", "false"])
    assert (c.epoch, c.lr_rate, c.num_rounds, c.shrink_lambda) == (100, 1e-5, 20, 10.0)
    assert c.update_types == ["mse_avg"] and c.reference_compat is False
    c2 = ExperimentConfig.from_args(["--epoch", "7"])
    assert c2.epoch == 7          # CLI wins over env


def test_validation():
    with pytest.raises(ValueError):
        ExperimentConfig(update_types=["nope"])
    with pytest.raises(ValueError):
        ExperimentConfig(verification_method="x")
    c = ExperimentConfig().replace(epoch=9)
    assert c.epoch == 9 and "9epoch" in c.no_Exp
