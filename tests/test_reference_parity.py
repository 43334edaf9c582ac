"""Differential tests: the torch oracle (and through it the kernels) against the reference's own modules,
imported read-only from /root/reference/src (skipped where the checkout is absent, e.g. on the GPU box)."""
import copy
import os
import random

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200.layout import DEFAULT_LAYOUT as L
from fedmse_decentralized_b200.ops import oracle
from fedmse_decentralized_b200.ops.oracle import TrainHP

from helpers import assert_close, make_cfg, make_engine, make_shards

pytestmark = pytest.mark.reference


def _ref_trainer(RM, RT, tmp, model_type="hybrid", update_type="avg", epoch=3, lam=5, patience=1, lr=1e-3):
    m = RM.Shrink_Autoencoder(input_dim=115, output_dim=115, shrink_lambda=lam) if model_type == "hybrid" \
        else RM.Autoencoder(input_dim=115, output_dim=115)
    return RT.ClientTrainer(model=m, epoch=epoch, batch_size=12, lr_rate=lr, update_type=update_type,
                            patience=patience, save_dir=str(tmp), client_id=0, model_type=model_type)


def _loaders(RDL, pc):
    from torch.utils.data import DataLoader
    mk = lambda x: DataLoader(RDL.IoTDataset(x, np.zeros(len(x))), batch_size=12)
    return mk(pc.train), mk(pc.valid)


@pytest.mark.parametrize("model_type,update_type", [("hybrid", "avg"), ("autoencoder", "avg"), ("hybrid", "fedprox")])
def test_local_training_matches_client_trainer_run(reference_modules, tiny_clients, tmp_path, model_type, update_type):
    import logging
    logging.disable(logging.CRITICAL)
    RM, RT, RDL = reference_modules["Model"], reference_modules["Trainer"], reference_modules["DataLoader"]
    pc = tiny_clients[0][0]
    torch.manual_seed(5)
    tr = _ref_trainer(RM, RT, tmp_path, model_type, update_type, epoch=3, patience=2)
    if update_type == "fedprox":     # move the anchor away so the proximal term matters
        with torch.no_grad():
            for p in tr.previous_global_model.parameters():
                p.mul_(0.8)
        tr.fedprox_mu = 0.05
    row = L.pack(tr.model.state_dict())
    anchor = L.pack(tr.previous_global_model.state_dict())
    tl, vl = _loaders(RDL, pc)
    tr.run(tl, vl)
    import pickle
    tracking = pickle.load(open(tmp_path / "training_tracking.pkl", "rb"))
    hp = TrainHP(epochs=3, batch_size=12, lr=1e-3, shrink_lambda=5.0 if model_type == "hybrid" else 0.0,
                 fedprox_mu=0.05 if update_type == "fedprox" else 0.0, patience=2)
    m, v = torch.zeros_like(row), torch.zeros_like(row)
    xt = torch.zeros(len(pc.train), 128); xt[:, :115] = torch.from_numpy(pc.train).float()
    xv = torch.zeros(len(pc.valid), 128); xv[:, :115] = torch.from_numpy(pc.valid).float()
    t, best, log = oracle.train_client(row, m, v, 0, anchor, xt, xv, L, hp)
    assert len(log) == len(tracking)
    assert_close(torch.tensor(log), torch.tensor(tracking), 1e-4, 1e-6, "loss curve")
    assert_close(row, L.pack(tr.model.state_dict()), 1e-3, 1e-5, "weights after run()")
    saved = torch.load(tmp_path / "model.cpt", weights_only=False)
    assert_close(best, L.pack(saved), 1e-3, 1e-5, "best-epoch checkpoint")
    logging.disable(logging.NOTSET)


def test_run_without_validation_trains_all_epochs(reference_modules, tiny_clients, tmp_path):
    """run(train_loader, valid_loader=None): the reference trains all epochs, never stops early and saves no
    model.cpt (client_trainer.py:387 guards validation, early stop and save_model); oracle and compat class agree."""
    import logging
    logging.disable(logging.CRITICAL)
    RM, RT, RDL = reference_modules["Model"], reference_modules["Trainer"], reference_modules["DataLoader"]
    pc = tiny_clients[0][1]
    torch.manual_seed(6)
    tr = _ref_trainer(RM, RT, tmp_path / "ref", epoch=4, patience=1)
    os.makedirs(tmp_path / "ref", exist_ok=True)
    row = L.pack(tr.model.state_dict())
    tl, _ = _loaders(RDL, pc)
    try:
        tr.run(tl, None)
    except Exception:                 # the reference's tracking append needs valid_loss; tolerate either outcome
        pytest.skip("reference run(valid_loader=None) raises in this version")
    assert not os.path.exists(tmp_path / "ref" / "model.cpt")
    hp = TrainHP(epochs=4, batch_size=12, lr=1e-3, shrink_lambda=5.0, patience=1)
    m, v = torch.zeros_like(row), torch.zeros_like(row)
    xt = torch.zeros(len(pc.train), 128); xt[:, :115] = torch.from_numpy(pc.train).float()
    t, best, log = oracle.train_client(row, m, v, 0, row.clone(), xt, xt[:0], L, hp)
    assert len(log) == 4 and t == 4 * ((len(pc.train) + 11) // 12)
    assert_close(row, L.pack(tr.model.state_dict()), 1e-3, 1e-5, "weights after run(train only)")
    logging.disable(logging.NOTSET)


def test_vote_score_matches_calculate_mse_score(reference_modules, tiny_clients, tmp_path):
    RM, RT = reference_modules["Model"], reference_modules["Trainer"]
    torch.manual_seed(1)
    tr = _ref_trainer(RM, RT, tmp_path)
    V = torch.from_numpy(np.concatenate([c.valid for c in tiny_clients[0]])).float()   # > 128 rows: 2 chunks
    assert len(V) > 128
    torch.manual_seed(9)
    want = tr.calculate_mse_score(V)
    torch.manual_seed(9)
    jit = 1.0 + (torch.rand(1).item() - 0.5) * 0.0002
    Vp = torch.zeros(len(V), 128); Vp[:, :115] = V
    got = oracle.chunked_mse(L.pack(tr.model.state_dict()), oracle.vote_normalise(Vp, 115), L, 128) * jit
    assert abs(float(got) - want) / want < 1e-5


@pytest.mark.parametrize("agg_pos", [0, 1, 3])
def test_fed_mse_avg_aliasing_quirk_reproduced(reference_modules, tiny_clients, tmp_path, agg_pos):
    """The reference's fed_mse_avg overwrites the aggregator's live state; compat mode
    must give the same merged weights, and non-compat the intended MSE-weighted mean."""
    import logging
    logging.disable(logging.CRITICAL)
    RM, RT = reference_modules["Model"], reference_modules["Trainer"]
    torch.manual_seed(2)
    trainers = [_ref_trainer(RM, RT, tmp_path / str(i), update_type="mse_avg") for i in range(4)]
    dev = tiny_clients[1]
    for t in trainers:
        t.create_dev_dataset({"dataset": dev})
    rows = torch.stack([L.pack(t.model.state_dict()) for t in trainers])
    devp = torch.zeros(len(dev), 128); devp[:, :115] = torch.from_numpy(dev).float()
    mse = oracle.mse_scores(rows, range(4), devp, L, 0)
    V = torch.from_numpy(tiny_clients[0][0].valid).float()
    agg = trainers[agg_pos]
    merged = agg.aggregate_models(trainers, V, 0)
    w = oracle.merge_weights("mse_avg", mse, 4, agg_pos, True)
    assert_close(oracle.weighted_reduce(rows, range(4), w), L.pack(merged), 1e-4, 1e-6, "compat merge")
    w_intended = oracle.merge_weights("mse_avg", mse, 4, agg_pos, False)
    assert abs(float(w_intended.sum()) - 1) < 1e-6 and float((w - w_intended).abs().max()) > 1e-3
    logging.disable(logging.NOTSET)


def test_fed_avg_and_verifier(reference_modules, tiny_clients, tmp_path):
    import logging
    logging.disable(logging.CRITICAL)
    RM, RT = reference_modules["Model"], reference_modules["Trainer"]
    torch.manual_seed(3)
    trainers = [_ref_trainer(RM, RT, tmp_path / str(i)) for i in range(3)]
    rows = torch.stack([L.pack(t.model.state_dict()) for t in trainers])
    merged = trainers[0].fed_avg([(t.model.state_dict(), 1.0) for t in trainers])
    assert_close(oracle.weighted_reduce(rows, range(3), oracle.merge_weights("avg", None, 3)), L.pack(merged),
                 1e-5, 1e-7, "fed_avg")
    # verifier: first call accepts; second compares against history
    from Trainer.model_verifier import ModelVerifier
    ver = ModelVerifier(3.0, 0.002, "val")
    V = torch.from_numpy(tiny_clients[0][0].valid).float()
    Vp = torch.zeros(len(V), 128); Vp[:, :115] = V
    s0, s1 = trainers[0].model.state_dict(), trainers[1].model.state_dict()
    assert ver.verify_model(0, s0, V, 0, "hybrid") == (True, 0.0)
    ok, dperf = ver.verify_model(0, s1, V, 1, "hybrid")
    change = oracle.param_change(rows[0], rows[1], L)
    d = oracle.performance(rows[1], Vp, L) - oracle.performance(rows[0], Vp, L)
    assert abs(float(d) - float(dperf)) < 1e-6
    assert ok == (bool(change <= 3.0) and bool(d >= -0.002))
    logging.disable(logging.NOTSET)


@pytest.mark.parametrize("model_type", ["hybrid", "autoencoder"])
def test_evaluator_auc_matches(reference_modules, tiny_clients, model_type):
    import logging
    logging.disable(logging.CRITICAL)
    from torch.utils.data import DataLoader
    RM, REV, RDL = reference_modules["Model"], reference_modules["Evaluator"], reference_modules["DataLoader"]
    import Evaluator.evaluator as _ev
    _ev.tqdm = lambda it, **kw: it
    pc = tiny_clients[0][1]
    torch.manual_seed(4)
    m = RM.Shrink_Autoencoder(input_dim=115, output_dim=115) if model_type == "hybrid" else RM.Autoencoder(115, 115)
    test = DataLoader(RDL.IoTDataset(pc.test, pc.test_label), batch_size=12)
    train = DataLoader(RDL.IoTDataset(pc.train, np.zeros(len(pc.train))), batch_size=12)
    out = REV.Evaluator(m, model_type=model_type, metric="AUC").evaluate(test, train)
    want = out[0] if isinstance(out, tuple) else out
    pad = lambda a: torch.cat([torch.from_numpy(a).float(), torch.zeros(len(a), 13)], 1)
    score, z = oracle.anomaly_scores(L.pack(m.state_dict()), pad(pc.test), pad(pc.train), L, model_type)
    got = oracle.auc_from_scores(score, torch.from_numpy(pc.test_label).float())
    assert abs(got - want) < 1e-4
    if model_type == "hybrid":
        assert np.allclose(out[1], z.numpy(), atol=1e-5) and out[1].shape[1] == 7
    logging.disable(logging.NOTSET)


def test_full_experiment_matches_reference_round_loop(reference_modules, tmp_path):
    """End to end: the engine (oracle backend, reference_compat) against the reference's own round loop on
    the same synthetic scenario, seeds and RNG order.  Per-round, per-client AUC within the parity bound."""
    import logging
    import sys
    logging.disable(logging.CRITICAL)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
    import ref_harness
    ref_harness.REF_DIR = "/root/reference/src"
    from fedmse_decentralized_b200.data import SyntheticSpec, make_clients, prepare_client, pool_dev_set, ShardStore
    import pandas as pd
    spec = SyntheticSpec(n_normal=260, n_abnormal=200, n_test_normal=50, jitter_rows=8)
    raw = make_clients(6, spec, seed=21)
    cfgd = {"model_type": "hybrid", "update_type": "mse_avg", "epoch": 2, "batch_size": 12, "lr_rate": 1e-3,
            "shrink_lambda": 5, "patience": 1, "num_participants": 0.5}
    # --- reference ---
    random.seed(1234); np.random.seed(1234); torch.manual_seed(0)
    fed = ref_harness.ReferenceFederation(raw, range(6), cfgd, 1, 0, torch.device("cpu"))
    ref_metrics = [fed.run_round(r) for r in range(3)]
    fed.close()
    # --- ours, same RNG consumption order ---
    # the harness re-seeds numpy per client (rank-independent shuffles) and shuffles abnormal right after normal
    random.seed(1234); np.random.seed(1234); torch.manual_seed(0)
    pcs = []
    for i, cl in enumerate(raw):
        np.random.seed(ref_harness.client_seed(1234, i))
        normal = pd.DataFrame(cl.normal).sample(frac=1).reset_index(drop=True).to_numpy()
        abnormal = pd.DataFrame(cl.abnormal).sample(frac=1).reset_index(drop=True).to_numpy()
        pcs.append(prepare_client(cl.name, normal, abnormal, cl.test_normal, True))
    np.random.seed(ref_harness.dev_seed(1234))
    dev = pool_dev_set(pcs, sampler=lambda a, n: pd.DataFrame(a).sample(n=n).to_numpy())
    sh = ShardStore.build(pcs, dev)
    from fedmse_decentralized_b200 import ExperimentConfig
    from fedmse_decentralized_b200.trainer import FederationEngine
    cfg = ExperimentConfig(epoch=2, num_rounds=3, network_size=6, backend="oracle", output_root=str(tmp_path),
                           reference_compat=True, write_latents=False, save_client_models=False)
    eng = FederationEngine(cfg, sh, "hybrid", "mse_avg")
    ours = [eng.run_round(r).client_metrics for r in range(3)]
    logging.disable(logging.NOTSET)
    for r in range(3):
        d = np.abs(np.array(ours[r]) - np.array(ref_metrics[r]))
        assert d.max() < 5e-3, (r, d)


@pytest.mark.parametrize("update_type", ["mse_avg", "avg"])
def test_real_nbaiot_scenario_auc_parity(reference_modules, tmp_path, update_type):
    """BASELINE config #1 on the prepared N-BaIoT IID split (real data from the reference checkout): the
    engine must track the reference's own round loop within the 0.005 AUC parity bound, client by client."""
    import logging
    import sys
    data_root = "/root/reference/Data/N-BaIoT/IID-10-Client_Data"
    if not os.path.isdir(data_root):
        pytest.skip("prepared N-BaIoT split not present")
    logging.disable(logging.CRITICAL)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline"))
    import ref_harness
    ref_harness.REF_DIR = "/root/reference/src"
    import pandas as pd
    from fedmse_decentralized_b200 import ExperimentConfig
    from fedmse_decentralized_b200.data import ShardStore, load_data, load_scenario, pool_dev_set, prepare_client
    from fedmse_decentralized_b200.data.synthetic import ClientRaw
    from fedmse_decentralized_b200.trainer import FederationEngine
    scen = load_scenario("configs/scen2-nba-iot-10clients.json")
    raw = []
    for d in scen["devices_list"][:6]:          # 6 clients keep the CPU test short
        raw.append(ClientRaw(d["name"], load_data(os.path.join(data_root, d["normal_data_path"])).to_numpy(),
                             load_data(os.path.join(data_root, d["abnormal_data_path"])).to_numpy(),
                             load_data(os.path.join(data_root, d["test_normal_data_path"])).to_numpy()))
    cfgd = {"model_type": "hybrid", "update_type": update_type, "epoch": 2, "batch_size": 12, "lr_rate": 1e-3,
            "shrink_lambda": 5, "patience": 1, "num_participants": 0.5}
    random.seed(1234); np.random.seed(1234); torch.manual_seed(0)
    fed = ref_harness.ReferenceFederation(raw, range(6), cfgd, 1, 0, torch.device("cpu"))
    ref_metrics = [fed.run_round(r) for r in range(2)]
    fed.close()
    random.seed(1234); np.random.seed(1234); torch.manual_seed(0)
    pcs = []
    for i, cl in enumerate(raw):
        np.random.seed(ref_harness.client_seed(1234, i))
        normal = pd.DataFrame(cl.normal).sample(frac=1).reset_index(drop=True).to_numpy()
        abnormal = pd.DataFrame(cl.abnormal).sample(frac=1).reset_index(drop=True).to_numpy()
        pcs.append(prepare_client(cl.name, normal, abnormal, cl.test_normal, True))
    np.random.seed(ref_harness.dev_seed(1234))
    dev = pool_dev_set(pcs, sampler=lambda a, n: pd.DataFrame(a).sample(n=n).to_numpy())
    cfg = ExperimentConfig(epoch=2, num_rounds=2, network_size=6, backend="oracle", output_root=str(tmp_path),
                           reference_compat=True, write_latents=False, save_client_models=False)
    eng = FederationEngine(cfg, ShardStore.build(pcs, dev), "hybrid", update_type)
    ours = [eng.run_round(r).client_metrics for r in range(2)]
    logging.disable(logging.NOTSET)
    for r in range(2):
        d = np.abs(np.array(ours[r]) - np.array(ref_metrics[r]))
        assert d.max() < 5e-3, (r, ours[r], ref_metrics[r])
    assert min(ours[-1]) > 0.9
