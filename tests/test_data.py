import json
import os
import random

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200.data import (IoTDataProccessor, IoTDataset, ShardStore, SyntheticSpec,
                                            dirichlet_partition, load_data, load_scenario, make_clients,
                                            prepare_client, prepare_from_scenario, repartition, split_sizes,
                                            write_scenario)
from fedmse_decentralized_b200.layout import D_PAD


def test_split_sizes_match_reference_arithmetic():
    for n in (1651, 1700, 925, 10, 3):
        tr, va, dv, te = split_sizes(n)
        assert (tr, va, dv) == (int(0.4 * n), int(0.1 * n), int(0.4 * n)) and tr + va + dv + te == n


def test_processor_semantics():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(100, 5)) * [1, 10, 1e9, 1e-3, 0]
    p = IoTDataProccessor("standard")
    Y, lab = p.fit_transform(X)
    assert np.allclose(Y[:, :4].mean(0), 0, atol=1e-9) and np.allclose(Y[:, :4].std(0), 1)
    assert np.all(Y[:, 4] == 0) and lab.sum() == 0           # constant column: scale 1, no NaN
    _, lab2 = p.transform(X, type="abnormal")
    assert lab2.sum() == 100
    meta = p.get_metadata()
    assert set(meta) == {"mean", "std"}
    ds = IoTDataset(Y, lab)
    x, y = ds[3]
    assert x.dtype == np.float32 and y.dtype == np.float32 and ds.input_dim_ == 5 and len(ds) == 100
    mm, _ = IoTDataProccessor("minmax").fit_transform(X[:, :2])
    assert abs(mm.min()) < 1e-12 and abs(mm.max() - 1) < 1e-12


def test_prepare_client_layout_and_labels():
    spec = SyntheticSpec(n_normal=200, n_abnormal=150, n_test_normal=40, jitter_rows=0)
    cl = make_clients(1, spec, seed=1)[0]
    pc = prepare_client("c", cl.normal, cl.abnormal, cl.test_normal, True)
    assert pc.train.shape == (80, 115) and pc.valid.shape == (20, 115) and pc.dev_raw.shape == (80, 115)
    assert pc.test.shape[0] == 20 + 40 + 150
    assert pc.test_label[:60].sum() == 0 and pc.test_label[60:].sum() == 150
    assert np.allclose(pc.train.mean(0), 0, atol=1e-9)       # scaler fitted on the training split only
    pc2 = prepare_client("c", cl.normal, cl.abnormal, cl.test_normal, False)
    assert pc2.test.shape[0] == 20 + 150


def test_shard_store_padding_and_offsets(tiny_clients):
    clients, dev = tiny_clients
    sh = ShardStore.build(clients, dev)
    assert sh.train.x.shape[1] == D_PAD and float(sh.train.x[:, 115:].abs().sum()) == 0
    assert sh.train.off_host[-1] == sum(len(c.train) for c in clients)
    assert torch.equal(sh.test.labels(2), torch.from_numpy(clients[2].test_label).float())
    assert np.allclose(sh.valid.rows(1)[:, :115].numpy(), clients[1].valid.astype(np.float32))
    assert sh.dev.shape == (len(dev), D_PAD) and sh.nbytes() > 0


def test_scenario_roundtrip_and_rng_contract(tmp_path):
    spec = SyntheticSpec(n_normal=120, n_abnormal=90, n_test_normal=30, jitter_rows=5)
    scen = write_scenario(str(tmp_path / "data"), 4, spec, seed=3)
    cfg_file = tmp_path / "scen.json"
    cfg_file.write_text(json.dumps(scen))
    loaded = load_scenario(str(cfg_file))
    assert [d["name"] for d in loaded["devices_list"]] == [f"Syn-Client-{k}" for k in range(1, 5)]
    df = load_data(os.path.join(scen["data_path"], "Client-1/normal"))
    assert df.shape[1] == 115

    def run():
        random.seed(1234); np.random.seed(1234)
        return prepare_from_scenario(loaded, 3, True, str(cfg_file))
    (c1, d1), (c2, d2) = run(), run()
    assert [c.name for c in c1] == [c.name for c in c2] and len(c1) == 3
    assert np.array_equal(c1[0].train, c2[0].train) and np.array_equal(d1, d2)
    assert d1.shape[0] == 3 * min(len(c.dev_raw) for c in c1)
    with pytest.raises(ValueError):
        (tmp_path / "bad.json").write_text("{}")
        load_scenario(str(tmp_path / "bad.json"))


def test_dirichlet_partition_iid_vs_noniid():
    labels = np.repeat(np.arange(5), 400)
    iid = dirichlet_partition(labels, 10, alpha=1000.0, seed=0)
    non = dirichlet_partition(labels, 10, alpha=0.1, seed=0)
    assert sum(len(p) for p in iid) == 2000
    sizes = np.array([len(p) for p in iid])
    assert sizes.std() / sizes.mean() < 0.1
    ent = lambda parts: np.mean([len(np.unique(labels[p])) for p in parts if len(p)])
    assert ent(non) < ent(iid)


def test_repartition_to_more_clients():
    spec = SyntheticSpec(n_normal=300, n_abnormal=200, n_test_normal=20, jitter_rows=0)
    devices = make_clients(3, spec, seed=5)
    out = repartition(devices, 8, alpha=1000.0, seed=1)
    assert len(out) == 8
    assert sum(len(c.normal) + len(c.test_normal) for c in out) == 900
    assert all(c.normal.shape[1] == 115 for c in out)
