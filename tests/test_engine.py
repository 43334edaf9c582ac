import copy
import json
import os
import pickle
import random

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.layout import STATE_KEYS
from fedmse_decentralized_b200.ops.backend import RoundControl
from fedmse_decentralized_b200.trainer import FederationEngine, GlobalStop, run_experiment
from fedmse_decentralized_b200.utils import artefacts

from helpers import assert_close, make_cfg, make_engine, make_shards


def test_round_semantics(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path)
    eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "avg")
    before = eng.state.params.clone()
    res = eng.run_round(0)
    assert len(res.selected) == 3 and len(res.client_metrics) == 6
    assert res.aggregator in res.selected[1:]                 # the first selected client votes, never for itself
    assert len(res.verification) == 5 and all(v["is_verified"] for v in res.verification)
    assert {v["client_id"] for v in res.verification} == set(range(6)) - {res.aggregator}
    # broadcast reaches *all* clients (also unselected ones): everyone now holds the merged model
    for c in range(6):
        assert torch.equal(eng.state.params[c], eng.state.params[res.aggregator])
        assert not torch.equal(eng.state.params[c], before[c])
    # the aggregator's FedProx anchor is not refreshed (quirk Q9), everyone else's is
    assert torch.equal(eng.state.anchor[res.aggregator], before[res.aggregator])
    other = next(c for c in range(6) if c != res.aggregator)
    assert torch.equal(eng.state.anchor[other], eng.state.params[other])
    assert int(eng.state.agg_count[res.aggregator]) == 1 and eng.agg_count_host[res.aggregator] == 1
    assert res.samples == sum(res.epochs_run[g] * (eng.sh.train.count(g) + eng.sh.valid.count(g)) for g in res.selected)


def test_aggregation_quota_and_no_aggregator(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path, epoch=1)
    eng = make_engine(cfg, make_shards(tiny_clients), "autoencoder", "avg")
    eng.agg_count_host = [3] * 6
    eng.state.agg_count.fill_(3)
    before = eng.state.params.clone()
    res = eng.run_round(0)
    assert res.aggregator == -1 and res.verification == []
    unsel = [c for c in range(6) if c not in res.selected]
    assert all(torch.equal(eng.state.params[c], before[c]) for c in unsel)   # nothing merged
    # one candidate with quota left must win regardless of score
    eng.agg_count_host[res.selected[0]] = 0
    eng.state.agg_count[res.selected[0]] = 0
    random.seed(5)
    ctl = eng.make_control(1)
    if ctl.voter_pos >= 0:
        r2 = eng.run_round(1, ctl)
        assert r2.aggregator == res.selected[0] or r2.aggregator == -1


def test_verification_rejects_large_jump(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path, epoch=1, verification_threshold=1e-6)
    eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "avg")
    r0 = eng.run_round(0)                 # first verification always accepts
    assert all(v["is_verified"] for v in r0.verification)
    kept = eng.state.params.clone()
    r1 = eng.run_round(1)                 # any change now exceeds the threshold -> rejected everywhere
    assert r1.aggregator >= 0 and not any(v["is_verified"] for v in r1.verification)
    for c in range(6):
        if c != r1.aggregator and c not in r1.selected:
            assert torch.equal(eng.state.params[c], kept[c])
    assert all(v["rejected_updates"] == 1 for v in r1.verification)
    # history is refreshed even on rejection (quirk Q10)
    c = next(c for c in range(6) if c != r1.aggregator)
    assert torch.equal(eng.state.hist[c], eng.state.params[r1.aggregator])


def test_global_stop_semantics():
    s = GlobalStop(patience=1)
    assert not s.update([0.9, 0.8])       # min improves (lower is "better", quirk Q2)
    assert not s.update([0.9, 0.85])      # worse #1 (not > patience)
    assert s.update([0.9, 0.86])          # worse #2 -> stop


def test_artefacts_match_reference_formats(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path, num_rounds=2)
    eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
    eng.fit()
    res = artefacts.results_file(str(tmp_path), 6, cfg.no_Exp, 0, "AUC", "FL-IoT", 0.5, "hybrid", "mse_avg")
    lines = artefacts.read_jsonl(res)
    assert [l["round"] for l in lines] == list(range(1, len(lines) + 1))
    assert set(lines[0]) == {"round", "client_metrics", "update_type", "model_type", "global_loss"}
    assert lines[0]["global_loss"] == min(lines[0]["client_metrics"]) and len(lines[0]["client_metrics"]) == 6
    ver = artefacts.read_jsonl(f"{tmp_path}/{cfg.checkpoint_dir}/Run_0/verification_results.json")
    assert set(ver[0]["verification_results"][0]) == {"client_id", "rejected_updates", "is_verified"}
    # client checkpoints: legacy (non-zip) torch format with the reference's 8 keys
    sel = eng.history[0].selected
    d = artefacts.client_model_dir(str(tmp_path), 6, cfg.no_Exp, 0, "FL-IoT", "hybrid", "mse_avg",
                                   eng.names_global[sel[0]])
    with open(os.path.join(d, "model.cpt"), "rb") as fh:
        assert fh.read(2) != b"PK"
    sd = torch.load(os.path.join(d, "model.cpt"), weights_only=False)
    assert tuple(sd.keys()) == STATE_KEYS and sd[STATE_KEYS[0]].shape == (27, 115)
    track = pickle.load(open(os.path.join(d, "training_tracking.pkl"), "rb"))
    assert isinstance(track, list) and len(track[0]) == 2
    # latent dump contract of latent_visualization.ipynb: dict[round][device] = (latent [N,7], label [N])
    lat = pickle.load(open(artefacts.latent_file(str(tmp_path), 6, cfg.no_Exp, 0, "hybrid", "mse_avg"), "rb"))
    rounds = list(lat.keys())
    z, y = lat[rounds[0]][eng.names_global[0]]
    assert z.dtype == np.float32 and z.shape == (eng.sh.test.count(0), 7) and y.shape == (eng.sh.test.count(0),)


def test_run_experiment_sweep_and_summary(tmp_path):
    cfg = ExperimentConfig(epoch=1, num_rounds=1, synthetic_clients=4, network_size=4, output_root=str(tmp_path),
                           backend="oracle", model_types=["hybrid", "autoencoder"], update_types=["avg", "mse_avg"],
                           write_latents=False, save_client_models=False)
    best = run_experiment(cfg)
    summ = json.load(open(f"{tmp_path}/Checkpoint/Results/Update/4/{cfg.no_Exp}/training_summary.json"))
    assert set(summ) == {"best_metrics", "metric_type", "num_runs", "network_size", "experiment_name"}
    assert summ["best_metrics"]["hybrid"]["mse_avg"] == best["hybrid"]["mse_avg"] and 0 < best["autoencoder"]["avg"] <= 1


def test_snapshot_restore_resumes_identically(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path, epoch=1)
    a = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
    a.run_round(0)
    snap = a.snapshot()
    ra = a.run_round(1)
    b = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg", seed=99)
    b.restore(copy.deepcopy(snap))
    rb = b.run_round(1)
    assert ra.selected == rb.selected and ra.aggregator == rb.aggregator
    assert ra.client_metrics == rb.client_metrics
    assert torch.equal(a.state.params, b.state.params) and torch.equal(a.state.adam_v, b.state.adam_v)


def test_gossip_round_topologies(tiny_clients, tmp_path):
    from fedmse_decentralized_b200.parallel import topology
    assert topology.ring(5) == [[4, 1], [0, 2], [1, 3], [2, 4], [3, 0]]
    rk = topology.random_k(6, 3, seed=1)
    assert all(len(set(n)) == 3 and i not in n for i, n in enumerate(rk))
    off, idx = topology.to_csr(topology.full_mesh(3))
    assert off.tolist() == [0, 3, 6, 9] and idx.tolist()[:3] == [0, 1, 2]
    for topo in ("ring", "random-k", "full"):
        cfg = make_cfg(tmp_path, epoch=1, aggregation="gossip", topology=topo, topology_k=2, num_participants=1.0)
        eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
        before = eng.state.params.clone()
        res = eng.run_round(0)
        assert res.aggregator == -1 and len(res.client_metrics) == 6
        assert all(not torch.equal(eng.state.params[c], before[c]) for c in range(6))
        assert int(eng.state.hist_valid.sum()) == 6


def test_control_prefetch_keeps_rng_order(tiny_clients, tmp_path):
    """Preparing the next round's selection / jitter draws ahead of time must not change any decision."""
    cfg = make_cfg(tmp_path, epoch=1, num_participants=0.7)
    a = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
    ra = [a.run_round(r) for r in range(3)]
    b = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
    rb = [b.run_round(r, prefetch_next=True) for r in range(3)]
    for x, y in zip(ra, rb):
        assert x.selected == y.selected and x.aggregator == y.aggregator and x.client_metrics == y.client_metrics
    assert torch.equal(a.state.params, b.state.params)


def test_streaming_data_update(tiny_clients, tmp_path):
    cfg = make_cfg(tmp_path, epoch=1)
    eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "avg")
    new_train = eng.sh.train.x * 0.5
    v0, v1 = eng.sh.valid.off_host[0], eng.sh.valid.off_host[6]
    new_valid = eng.sh.valid.x[v0:v1] * 2.0
    eng.stage_next_data(new_train.clone(), new_valid.clone())
    assert not torch.equal(eng.sh.train.x, new_train)          # not visible before the commit
    eng.commit_staged_data()
    assert torch.equal(eng.sh.train.x, new_train) and torch.equal(eng.sh.valid.x[v0:v1], new_valid)
    eng.update_local_data(train_x=new_train * 2.0)
    assert torch.equal(eng.sh.train.x, new_train * 2.0)
    assert len(eng.run_round(0).client_metrics) == 6


def test_next_voter_decidability(tiny_clients, tmp_path):
    """The pipelined round may build the next control block early only when this round's (still unknown)
    aggregator cannot change who votes next."""
    cfg = make_cfg(tmp_path, network_size=6, max_aggregation_threshold=3)
    eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg", "oracle")
    cur, nxt = [0, 1, 2], [3, 1, 4]
    eng.agg_count_host = [0, 0, 0, 0, 0, 0]
    assert eng._voter_is_decided(nxt, cur)                 # plenty of clients under the quota
    eng.agg_count_host = [0, 2, 0, 3, 3, 0]
    assert not eng._voter_is_decided(nxt, cur)             # voter 3 needs client 1, who may reach the quota now
    eng.agg_count_host = [0, 1, 0, 3, 3, 0]
    assert eng._voter_is_decided(nxt, cur)                 # one more aggregation leaves client 1 under it
    eng.agg_count_host = [0, 3, 0, 3, 3, 0]
    assert eng._voter_is_decided(nxt, cur)                 # nobody can vote, whatever happens
    eng.agg_count_host = [0, 2, 0, 3, 3, 0]
    assert eng._voter_is_decided(nxt, [0, 2, 5])           # client 1 is not selected now: its count cannot move


def test_pipeline_flag_is_harmless_without_a_device(tiny_clients, tmp_path):
    """The oracle backend has no asynchronous read-back: pipeline_next must silently run the plain order."""
    runs = []
    for piped in (False, True):
        cfg = make_cfg(tmp_path, epoch=1)
        eng = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg", "oracle")
        random.seed(5); torch.manual_seed(5)
        res = [eng.run_round(r, pipeline_next=piped and r < 2) for r in range(3)]
        assert eng._pre is None
        runs.append((res, eng.state.params.clone()))
    for a, b in zip(runs[0][0], runs[1][0]):
        assert (a.selected, a.aggregator, a.client_metrics) == (b.selected, b.aggregator, b.client_metrics)
    assert torch.equal(runs[0][1], runs[1][1])


def test_fusion_avg_is_a_selectable_merge_rule(tiny_clients, tmp_path):
    """The historical 4th merge rule: weights = normalised inverse JS distance (utils.similarity)."""
    from fedmse_decentralized_b200.ops import oracle
    from fedmse_decentralized_b200.utils.similarity import fusion_distances
    sh = make_shards(tiny_clients)
    cfg = make_cfg(tmp_path, update_types=["fusion_avg"], fusion_rows=64)
    eng = make_engine(cfg, sh, "hybrid", "fusion_avg")
    ctl = eng.make_control(0)
    eng.backend.train(eng.state, sh, eng._sel_local(ctl.sel), eng.hp)
    before = eng.state.params.clone()
    scores = eng.backend.vote_scores(eng.state, sh, ctl, eng.spec)
    agg = eng.backend.elect(eng.state, scores, ctl, eng.spec)
    d = eng.backend.fusion_scores(eng.state, sh, ctl, 64)
    theta = eng.backend.merge(eng.state, sh, ctl, eng.spec, d, agg)
    w = (1.0 / d) / (1.0 / d).sum()
    want = sum(w[k] * before[g] for k, g in enumerate(ctl.sel))
    assert_close(theta, want, rtol=1e-5, atol=1e-6, what="fusion_avg theta")
    assert float(w.min()) > 0 and abs(float(w.sum()) - 1) < 1e-5
    # and it runs as a full round through the engine
    eng2 = make_engine(cfg, sh, "hybrid", "fusion_avg")
    r = eng2.run_round(0)
    assert r.aggregator >= 0 and len(r.client_metrics) == 6


def test_checkpoint_file_resumes_identically(tiny_clients, tmp_path):
    """save_checkpoint / load_checkpoint: a fresh engine continues with exactly the rounds the first one would have run."""
    cfg = make_cfg(tmp_path, epoch=1)
    a = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg")
    a.run_round(0)
    a.run_round(1)
    path = a.save_checkpoint(str(tmp_path / "ckpt" / "engine.pt"))
    assert os.path.exists(path) and not os.path.exists(path + ".tmp")
    ra = [a.run_round(r) for r in (2, 3)]
    b = make_engine(cfg, make_shards(tiny_clients), "hybrid", "mse_avg", seed=99)
    nxt = b.load_checkpoint(path)
    assert nxt == 2
    rb = [b.run_round(r) for r in (nxt, nxt + 1)]
    for x, y in zip(ra, rb):
        assert (x.selected, x.aggregator, x.client_metrics) == (y.selected, y.aggregator, y.client_metrics)
    assert torch.equal(a.state.params, b.state.params) and torch.equal(a.state.adam_m, b.state.adam_m)
    assert b.snapshot()["rounds_done"] == 4
    c = make_engine(cfg, make_shards(tiny_clients), "hybrid", "avg")
    with pytest.raises(ValueError):
        c.load_checkpoint(path)


def test_fit_resumes_from_periodic_checkpoints(tiny_clients, tmp_path):
    """--checkpoint_every / --resume: an interrupted fit() continued by a new process ends where the uninterrupted one does."""
    kw = dict(epoch=1, num_rounds=4, global_patience=10, checkpoint_every=1, save_client_models=False, write_latents=False)
    full = make_engine(make_cfg(tmp_path / "full", **kw), make_shards(tiny_clients), "hybrid", "mse_avg")
    want = full.fit()
    cfg2 = make_cfg(tmp_path / "cut", **dict(kw, num_rounds=2))
    first = make_engine(cfg2, make_shards(tiny_clients), "hybrid", "mse_avg")
    first.fit()                                                     # "crashes" after two of four rounds
    cfg3 = make_cfg(tmp_path / "cut", **dict(kw, resume=True))
    second = make_engine(cfg3, make_shards(tiny_clients), "hybrid", "mse_avg", seed=5)
    got = second.fit()
    assert len(got) == 2 and [r.round for r in got] == [2, 3] if hasattr(got[0], "round") else len(got) == 2
    assert got[-1].client_metrics == want[-1].client_metrics and got[-1].aggregator == want[-1].aggregator
    assert torch.equal(second.state.params, full.state.params)
