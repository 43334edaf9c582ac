"""Shape-generic kernels (csrc/generic_kernels.cu): autoencoders of any width and batches above 32 rows against the
fp32 torch oracle.  The reference exposes hidden_neus / latent_dim on its model classes
(src/Model/Shrink_Autoencoder.py:125-127) and the thesis trains with batch 64."""
import copy
import random

import pytest
import torch

from fedmse_decentralized_b200.ops.backend import RoundControl

from helpers import assert_close, make_cfg, make_engine, make_shards

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    return torch.device("cuda", 0)


def _pair(tiny_clients, tmp_path, dev, model_type, update_type, **kw):
    cfg = make_cfg(tmp_path, **kw)
    e_ref = make_engine(cfg, make_shards(tiny_clients, "cpu"), model_type, update_type, "oracle")
    e_gpu = make_engine(cfg, make_shards(tiny_clients, dev), model_type, update_type, "cuda")
    assert_close(e_gpu.state.params, e_ref.state.params, 0, 0, "init params")
    return e_ref, e_gpu


@pytest.mark.parametrize("hidden,latent,batch,upd", [(40, 9, 12, "avg"), (64, 16, 64, "fedprox"), (8, 2, 5, "avg")])
def test_generic_training_matches_oracle(tiny_clients, tmp_path, dev, hidden, latent, batch, upd):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", upd, epoch=3, local_patience=5, hidden_neus=hidden,
                         latent_dim=latent, batch_size=batch)
    assert e_gpu.backend.name == "cuda-generic"
    if upd == "fedprox":
        for e in (e_ref, e_gpu):
            e.state.anchor.mul_(0.9)
            e.hp.fedprox_mu = 0.05
    sel = [0, 2, 5]
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, sel, e.hp)
    torch.cuda.synchronize()
    assert e_gpu.state.epochs_run.tolist() == e_ref.state.epochs_run.tolist()
    assert e_gpu.state.adam_t.tolist() == e_ref.state.adam_t.tolist()
    n = int(e_ref.state.epochs_run[0])
    assert_close(e_gpu.state.loss_log[sel, :n], e_ref.state.loss_log[sel, :n], 2e-4, 1e-6, "loss log")
    assert_close(e_gpu.state.params, e_ref.state.params, 2e-3, 2e-5, "params after training")
    assert_close(e_gpu.state.best, e_ref.state.best, 2e-3, 2e-5, "best-epoch snapshot")
    assert torch.equal(e_gpu.state.params[1].cpu(), e_ref.state.params[1])


def test_batch_64_on_the_default_model(tiny_clients, tmp_path, dev):
    """The thesis setting: 115-27-7 with batch 64 -- CudaBackend routes training to the generic kernel."""
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "avg", epoch=2, local_patience=5, batch_size=64)
    assert e_gpu.backend.name == "cuda"
    sel = [1, 4]
    ctl = RoundControl(0, sel, 0, [1.0] * 2, sel[0], -1)
    e_gpu.backend.begin_round(ctl)
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, sel, e.hp)
    torch.cuda.synchronize()
    assert e_gpu.state.epochs_run.tolist() == e_ref.state.epochs_run.tolist()
    assert_close(e_gpu.state.params, e_ref.state.params, 2e-3, 2e-5, "params, batch 64")


@pytest.mark.parametrize("model_type,upd", [("hybrid", "mse_avg"), ("autoencoder", "avg")])
def test_generic_full_round_matches_oracle(tiny_clients, tmp_path, dev, model_type, upd):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, model_type, upd, hidden_neus=48, latent_dim=5,
                         num_participants=0.7)
    for r in range(2):
        random.seed(40 + r); torch.manual_seed(40 + r)
        ctl = e_ref.make_control(r)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert res_gpu.aggregator == res_ref.aggregator
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 3e-5, f"params r{r}")
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")
