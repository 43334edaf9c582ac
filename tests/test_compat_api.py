"""The reference-named object API (ClientTrainer / ModelVerifier / Evaluator / DataLoader names) behaves like
the reference's classes when driven the way ``src/main.py`` drives them."""
import logging

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200 import compat
from fedmse_decentralized_b200.layout import DEFAULT_LAYOUT as L

from helpers import assert_close


def _loaders(DL, pc, bs=12):
    from torch.utils.data import ConcatDataset, DataLoader
    n_norm = int((pc.test_label == 0).sum())
    test = ConcatDataset([DL.IoTDataset(pc.test[:n_norm], pc.test_label[:n_norm]),
                          DL.IoTDataset(pc.test[n_norm:], pc.test_label[n_norm:])])
    return (DataLoader(DL.IoTDataset(pc.train, np.zeros(len(pc.train))), batch_size=bs),
            DataLoader(DL.IoTDataset(pc.valid, np.zeros(len(pc.valid))), batch_size=bs),
            DataLoader(test, batch_size=bs))


def _federation(ns, tiny_clients, tmp, update_type, device=None):
    clients, dev = tiny_clients
    torch.manual_seed(0)
    trainers, loaders = [], []
    for i, pc in enumerate(clients[:4]):
        m = ns.Shrink_Autoencoder(input_dim=115, output_dim=115, shrink_lambda=5)
        kw = dict(model=m, epoch=2, batch_size=12, lr_rate=1e-3, update_type=update_type, patience=1,
                  save_dir=str(tmp / f"c{i}"), client_id=i, model_type="hybrid")
        if device is not None:
            kw["device"] = device
        t = ns.ClientTrainer(**kw)
        t.create_dev_dataset({"dataset": dev})
        trainers.append(t)
        loaders.append(_loaders(ns, pc))
    for i, t in enumerate(trainers):
        t.connect_to_peers([c for j, c in enumerate(trainers) if j != i])
        t.validation_data = torch.Tensor(clients[3].valid)
    return trainers, loaders


class _RefNS:
    pass


@pytest.mark.parametrize("update_type", ["avg", "mse_avg"])
def test_object_api_round_matches_reference(reference_modules, tiny_clients, tmp_path, update_type):
    logging.disable(logging.CRITICAL)
    import Evaluator.evaluator as _ev
    _ev.tqdm = lambda it, **kw: it
    ref = _RefNS()
    ref.Shrink_Autoencoder = reference_modules["Model"].Shrink_Autoencoder
    ref.ClientTrainer = reference_modules["Trainer"].ClientTrainer
    ref.IoTDataset = reference_modules["DataLoader"].IoTDataset
    ref.Evaluator = reference_modules["Evaluator"].Evaluator
    V = torch.Tensor(tiny_clients[0][0].valid)
    out = {}
    for name, ns, dev in (("ref", ref, None), ("ours", compat, "cpu")):
        trainers, loaders = _federation(ns, tiny_clients, tmp_path / name, update_type, dev)
        for t, (tl, vl, _) in zip(trainers[:3], loaders[:3]):
            t.run(tl, vl)
        torch.manual_seed(42)
        agg = trainers[0].vote_for_aggregator(trainers[:3], V, 0)
        kw = {"reference_compat": True} if name == "ours" else {}
        state = agg.aggregate_models(trainers[:3], V, 0, **kw)
        for t in trainers:
            if t is not agg:
                t.receive_model(agg, state)
        for t in trainers:
            if t is not agg:
                t.update_from_peers()
        aucs = []
        for t, (tl, _, te) in zip(trainers, loaders):
            r = ns.Evaluator(t.model, model_type="hybrid", metric="AUC").evaluate(te, tl)
            aucs.append(r[0])
            assert r[1].shape[1] == 7 and len(r[2]) == len(r[1])
        out[name] = dict(agg=trainers.index(agg), state=L.pack(state), aucs=aucs,
                         rej=[t.rejected_updates for t in trainers],
                         w3=L.pack(trainers[3].model.state_dict()))
    logging.disable(logging.NOTSET)
    assert out["ref"]["agg"] == out["ours"]["agg"] and out["ref"]["rej"] == out["ours"]["rej"]
    assert_close(out["ours"]["state"], out["ref"]["state"], 2e-3, 2e-5, "aggregated state")
    assert_close(out["ours"]["w3"], out["ref"]["w3"], 2e-3, 2e-5, "unselected client after broadcast")
    assert max(abs(a - b) for a, b in zip(out["ours"]["aucs"], out["ref"]["aucs"])) < 2e-3


def test_gossip_api_and_quota(tiny_clients, tmp_path):
    trainers, _ = _federation(compat, tiny_clients, tmp_path, "avg", "cpu")
    for t in trainers:
        t.broadcast_model()
    a = trainers[0]
    assert len(a.received_models) == 3
    merged = a.request_aggregation()
    want = torch.stack([t.params for t in trainers[1:]]).mean(0)
    assert_close(L.pack(merged), want, 1e-5, 1e-7, "gossip mean of received models")
    a.aggregation_count = 3
    assert a.request_aggregation() is None and a.aggregate_models(trainers, None, 0) is None
    v = compat.ModelVerifier(1e-9, 0.002, "val")
    V = torch.Tensor(tiny_clients[0][0].valid)
    assert v.verify_model(0, trainers[0].state_dict(), V, 0)[0]
    assert not v.verify_model(0, trainers[1].state_dict(), V, 1)[0]     # any change exceeds the threshold
    assert not compat.ModelVerifier(3.0, 0.002, "dev").verify_model(0, trainers[0].state_dict(), V, 0)[0]


def test_evaluator_metrics_and_time(tiny_clients):
    pc = tiny_clients[0][2]
    torch.manual_seed(1)
    tl, _, te = _loaders(compat, pc)
    for mt, cls in (("hybrid", compat.Shrink_Autoencoder), ("autoencoder", compat.Autoencoder)):
        m = cls(115, 115)
        auc = compat.Evaluator(m, mt, "AUC").evaluate(te, tl)
        auc = auc[0] if isinstance(auc, tuple) else auc
        assert 0.0 <= auc <= 1.0
        f1 = compat.Evaluator(m, mt, "classification").evaluate(te, tl)
        f1 = f1[0] if isinstance(f1, tuple) else f1
        assert 0.0 <= f1 <= 1.0
        assert compat.Evaluator(m, mt, "time").evaluate(te, tl) >= 0.0


def test_global_aggregator_central_server_rules(tiny_clients, tmp_path):
    """The central-server surface the fork deleted (names from its leftover bytecode): four merge rules + update()."""
    from fedmse_decentralized_b200.ops import oracle
    from fedmse_decentralized_b200.utils.similarity import fusion_distances
    logging.disable(logging.CRITICAL)
    clients, dev = tiny_clients
    trainers, loaders = _federation(compat, tiny_clients, tmp_path, "avg", "cpu")
    for t, (tl, vl, _) in zip(trainers[:3], loaders[:3]):
        t.run(tl, vl)
    rows = torch.stack([t.params for t in trainers[:3]])
    local = [(t.state_dict(), n) for t, n in zip(trainers[:3], (10.0, 30.0, 60.0))]
    mk = lambda rule: compat.GlobalAggregator(model=compat.Shrink_Autoencoder(input_dim=115, output_dim=115, shrink_lambda=5),
                                              update_type=rule, device="cpu", fusion_rows=64)
    # FedAvg / FedProx: sample-count weights
    g = mk("avg")
    want = 0.1 * rows[0] + 0.3 * rows[1] + 0.6 * rows[2]
    assert_close(L.pack(g.fed_avg(local)), want, 1e-6, 1e-7, "fed_avg")
    assert_close(L.pack(mk("fedprox").fedprox(local)), want, 1e-6, 1e-7, "fedprox")
    # MSEAvg: the same state as the decentralized trainer's own fed_mse_avg
    g = mk("mse_avg")
    with pytest.raises(RuntimeError):
        g.fed_mse_avg(local)
    g.create_dev_dataset({"dataset": dev})
    assert_close(L.pack(g.fed_mse_avg(local)), L.pack(trainers[0].fed_mse_avg(local)), 1e-6, 1e-7, "fed_mse_avg")
    assert abs(float(g.last_weights.sum()) - 1) < 1e-6
    # fusion_avg: inverse Jensen-Shannon distance weights
    g = mk("fusion_avg")
    g.create_dev_dataset(dev)
    merged = g.fusion_avg(local)
    from fedmse_decentralized_b200.trainer.model_verifier import _pad
    d = fusion_distances(rows, _pad(torch.as_tensor(dev, dtype=torch.float32)), L, 64)
    w = (1 / d) / (1 / d).sum()
    assert_close(g.last_weights, w, 1e-5, 1e-7, "fusion weights")
    assert_close(L.pack(merged), oracle.weighted_reduce(rows, range(3), w), 1e-6, 1e-7, "fusion_avg")
    # update(): merged model becomes the global model and lands in every receiver
    out = g.update(trainers[:3], receivers=trainers)
    assert_close(g.params, L.pack(out), 0, 0, "global model")
    for t in trainers:
        assert_close(t.params, g.params, 0, 0, "client got the global model")
        assert_close(t.anchor, g.params, 0, 0, "FedProx anchor follows")
    # election helper: lowest validation MSE under quota
    torch.manual_seed(3)
    pick = g.select_aggregator(trainers, torch.Tensor(clients[0].valid), 1)
    assert pick in trainers
    for t in trainers:
        t.aggregation_count = t.max_aggregation_threshold
    assert g.select_aggregator(trainers, torch.Tensor(clients[0].valid), 2) is None
    with pytest.raises(ValueError):
        compat.GlobalAggregator(model=trainers[0].model, update_type="nope")
