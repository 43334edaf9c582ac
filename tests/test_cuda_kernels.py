"""GPU numerics: every sm_100a kernel against the plain-PyTorch fp32 oracle of the same op."""
import copy
import random

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200.layout import DEFAULT_LAYOUT as L
from fedmse_decentralized_b200.ops import oracle
from fedmse_decentralized_b200.ops.backend import MergeSpec, OracleBackend, RoundControl
from fedmse_decentralized_b200.ops.oracle import TrainHP

from helpers import assert_close, make_cfg, make_engine, make_shards

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda", 0)


def _pair(tiny_clients, tmp_path, dev, model_type="hybrid", update_type="mse_avg", **kw):
    cfg = make_cfg(tmp_path, **kw)
    e_ref = make_engine(cfg, make_shards(tiny_clients, "cpu"), model_type, update_type, "oracle")
    e_gpu = make_engine(cfg, make_shards(tiny_clients, dev), model_type, update_type, "cuda")
    assert_close(e_gpu.state.params, e_ref.state.params, 0, 0, "init params")
    return e_ref, e_gpu


@pytest.mark.parametrize("model_type,update_type", [("hybrid", "avg"), ("autoencoder", "avg"),
                                                    ("hybrid", "fedprox")])
def test_train_kernel_matches_oracle(tiny_clients, tmp_path, dev, model_type, update_type):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, model_type, update_type, epoch=3, local_patience=5)
    if update_type == "fedprox":   # make the proximal term non-trivial
        for e in (e_ref, e_gpu):
            e.state.anchor.mul_(0.9)
            e.hp.fedprox_mu = 0.05
    sel = [0, 2, 3, 5]
    ctl = RoundControl(0, sel, 0, [1.0] * 4, sel[0], -1)
    e_gpu.backend.begin_round(ctl)
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, sel, e.hp)
    torch.cuda.synchronize()
    assert e_gpu.state.epochs_run.tolist() == e_ref.state.epochs_run.tolist()
    assert e_gpu.state.adam_t.tolist() == e_ref.state.adam_t.tolist()
    n = int(e_ref.state.epochs_run[0])
    assert_close(e_gpu.state.loss_log[sel, :n], e_ref.state.loss_log[sel, :n], 2e-4, 1e-6, "loss log")
    assert_close(e_gpu.state.params, e_ref.state.params, 2e-3, 2e-5, "params after training")
    assert_close(e_gpu.state.best, e_ref.state.best, 2e-3, 2e-5, "best-epoch snapshot")
    assert_close(e_gpu.state.adam_m, e_ref.state.adam_m, 5e-3, 1e-6, "adam m")
    # untouched clients stay bit-identical
    assert torch.equal(e_gpu.state.params[1].cpu(), e_ref.state.params[1])


@pytest.mark.parametrize("batch", [8, 32])
def test_train_kernel_other_batch_sizes(tiny_clients, tmp_path, dev, batch):
    """Batches below 12 rows run on the tensor-core variant with a partly filled M = 16 tile, batches of 13..32
    rows on the all-FFMA variant; both against the oracle."""
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "avg", epoch=2, local_patience=5, batch_size=batch)
    sel = [0, 3, 4]
    ctl = RoundControl(0, sel, 0, [1.0] * 3, sel[0], -1)
    e_gpu.backend.begin_round(ctl)
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, sel, e.hp)
    torch.cuda.synchronize()
    assert e_gpu.state.epochs_run.tolist() == e_ref.state.epochs_run.tolist()
    assert e_gpu.state.adam_t.tolist() == e_ref.state.adam_t.tolist()
    assert_close(e_gpu.state.params, e_ref.state.params, 2e-3, 2e-5, f"params, batch {batch}")
    assert_close(e_gpu.state.adam_v, e_ref.state.adam_v, 5e-3, 1e-9, f"adam v, batch {batch}")


def test_train_early_stop_and_partial_batch(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "avg", epoch=6, local_patience=1, lr_rate=0.05)
    sel = [1, 4]
    ctl = RoundControl(0, sel, 0, [1.0] * 2, sel[0], -1)
    e_gpu.backend.begin_round(ctl)
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, sel, e.hp)
    assert e_gpu.state.epochs_run.tolist() == e_ref.state.epochs_run.tolist()
    assert_close(e_gpu.state.params[sel], e_ref.state.params[sel], 5e-3, 5e-5, "params")


def test_score_rows_matches_oracle(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev)
    be, st, sh = e_gpu.backend, e_gpu.state, e_gpu.sh
    be._score(st.params, st.n_clients, 0, None, sh.test.x, be.tiles_test, be.test_sse, be.test_lat)
    torch.cuda.synchronize()
    for c in range(st.n_clients):
        x = e_ref.sh.test.rows(c)
        z, _ = oracle.forward(e_ref.state.params[c], x[:, :L.dim], L)
        sse = oracle.row_sse(e_ref.state.params[c], x, L)
        lo, hi = sh.test.off_host[c], sh.test.off_host[c + 1]
        assert_close(be.test_sse[lo:hi], sse, 1e-4, 1e-4, f"sse client {c}")
        assert_close(be.test_lat[lo:hi, :7], z, 1e-4, 1e-5, f"latent client {c}")


@pytest.mark.parametrize("model_type", ["hybrid", "autoencoder"])
def test_auc_kernel_matches_sklearn(tiny_clients, tmp_path, dev, model_type):
    from sklearn.metrics import auc, roc_curve
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, model_type, "avg")
    a_gpu, _, lat = e_gpu.backend.evaluate(e_gpu.state, e_gpu.sh, model_type, True)
    a_ref, a64, _ = e_ref.backend.evaluate(e_ref.state, e_ref.sh, model_type, False)
    assert_close(a_gpu, torch.tensor(a64), 0, 2e-4, "auc vs oracle")
    # exactness of the device rank statistic against sklearn on the device's own scores
    be, sh = e_gpu.backend, e_gpu.sh
    e_gpu.backend.evaluate(e_gpu.state, e_gpu.sh, model_type, False, metric="classification")
    torch.cuda.synchronize()
    for c in range(sh.n_clients):
        lo, hi = sh.test.off_host[c], sh.test.off_host[c + 1]
        s = be.scores_out[lo:hi].cpu().numpy()
        y = sh.test.y[lo:hi].cpu().numpy()
        fpr, tpr, _ = roc_curve(y, s)
        assert abs(auc(fpr, tpr) - float(be.auc[c])) < 1e-12
    assert lat[0].shape == (sh.test.count(0), 7)


def test_auc_kernel_ties_and_nonfinite(dev):
    import ctypes as C
    from fedmse_decentralized_b200.ops import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(1)
    n = 1000
    sse = torch.randint(0, 20, (n,), generator=g).float() * 115.0      # heavy ties
    sse[5] = float("nan"); sse[6] = float("inf"); sse[7] = -0.0
    y = (torch.rand(n, generator=g) > 0.6).float()
    want = oracle.auc_from_scores(sse / 115.0, y)
    d = lambda t: t.to(dev)
    off = d(torch.tensor([0, n], dtype=torch.int32))
    out = torch.zeros(1, dtype=torch.float64, device=dev)
    sse_d, y_d = d(sse), d(y)
    a = _lib.FmAucArgs(0, 1, sse_d.data_ptr(), 0, y_d.data_ptr(), off.data_ptr(), 0, 0, 0, out.data_ptr(), n)
    _lib.check(lib.fm_auc(C.byref(a), torch.cuda.current_stream().cuda_stream))
    assert abs(float(out[0]) - want) < 1e-12


def test_auc_kernel_large_test_sets_use_global_scratch(dev):
    """More than 16,384 test rows per client (raw N-BaIoT devices, Configuration/nba-iot-training.json): the keys
    are sorted in global scratch instead of shared memory; two clients of different sizes, with ties."""
    import ctypes as C
    from fedmse_decentralized_b200.ops import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(3)
    sizes = [40000, 17001]
    per = int(lib.fm_auc_scratch_keys(max(sizes)))
    assert per == 65536 and int(lib.fm_auc_scratch_keys(16384)) == 0
    sse = torch.cat([torch.randn(sizes[0], generator=g).abs() * 115.0,
                     torch.randint(0, 50, (sizes[1],), generator=g).float() * 115.0])
    y = (torch.rand(sum(sizes), generator=g) > 0.7).float()
    off_h = [0, sizes[0], sum(sizes)]
    want = [oracle.auc_from_scores(sse[off_h[c]:off_h[c + 1]] / 115.0, y[off_h[c]:off_h[c + 1]]) for c in range(2)]
    off = torch.tensor(off_h, dtype=torch.int32, device=dev)
    out = torch.zeros(2, dtype=torch.float64, device=dev)
    sse_d, y_d = sse.to(dev), y.to(dev)
    a = _lib.FmAucArgs(0, 2, sse_d.data_ptr(), 0, y_d.data_ptr(), off.data_ptr(), 0, 0, 0, out.data_ptr(), max(sizes), 0)
    assert lib.fm_auc(C.byref(a), torch.cuda.current_stream().cuda_stream) != 0      # scratch is required here
    scratch = torch.empty(2 * per, dtype=torch.int64, device=dev)
    a.keys_scratch = scratch.data_ptr()
    _lib.check(lib.fm_auc(C.byref(a), torch.cuda.current_stream().cuda_stream))
    for c in range(2):
        assert abs(float(out[c]) - want[c]) < 1e-12, (c, float(out[c]), want[c])


@pytest.mark.parametrize("update_type,compat", [("avg", True), ("mse_avg", True), ("mse_avg", False),
                                                ("fedprox", True)])
def test_full_round_matches_oracle(tiny_clients, tmp_path, dev, update_type, compat):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", update_type, reference_compat=compat,
                         num_participants=0.7)
    for r in range(3):
        random.seed(100 + r); torch.manual_seed(100 + r)
        ctl = e_ref.make_control(r)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert res_gpu.aggregator == res_ref.aggregator and res_ref.aggregator >= 0
        assert res_gpu.epochs_run == res_ref.epochs_run
        assert [v["is_verified"] for v in res_gpu.verification] == [v["is_verified"] for v in res_ref.verification]
        assert_close(e_gpu.backend.theta, e_ref.state.params[res_ref.aggregator], 3e-3, 3e-5, f"theta r{r}")
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 3e-5, f"params r{r}")
        assert_close(e_gpu.state.hist_perf, e_ref.state.hist_perf, 1e-4, 1e-6, f"hist perf r{r}")
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")
    assert e_gpu.state.agg_count.tolist() == e_ref.state.agg_count.tolist()
    assert e_gpu.backend.launches > 0


def test_vote_scores_and_election(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev)
    sel = [4, 1, 3, 0]
    ctl = RoundControl(0, sel, 0, [1.0, 1.00005, 0.99995, 1.00002], sel[0], 5)
    e_gpu.backend.begin_round(ctl)
    s_ref = e_ref.backend.vote_scores(e_ref.state, e_ref.sh, ctl, e_ref.spec)
    s_gpu = e_gpu.backend.vote_scores(e_gpu.state, e_gpu.sh, ctl, e_gpu.spec)
    assert_close(s_gpu, s_ref, 1e-4, 1e-6, "vote scores")
    assert float(s_gpu[0]) == 0.0
    # quota: exclude the best candidate
    best = int(torch.argmin(torch.where(s_ref > 0, s_ref, torch.full_like(s_ref, 1e30))))
    for e in (e_ref, e_gpu):
        e.state.agg_count[sel[best]] = 3
    a_ref = e_ref.backend.elect(e_ref.state, s_ref, ctl, e_ref.spec)
    a_gpu = e_gpu.backend.elect(e_gpu.state, s_gpu, ctl, e_gpu.spec)
    assert a_gpu.tolist() == a_ref.tolist() and a_ref[0] != best


@pytest.mark.parametrize("topology,upd", [("ring", "mse_avg"), ("random-k", "mse_avg"), ("full", "mse_avg"),
                                          ("random-k", "avg")])
def test_gossip_matches_oracle(tiny_clients, tmp_path, dev, topology, upd):
    """Fused gossip kernel (device-generated graph, one fetch per neighbour row) against the torch oracle that
    walks the host copy of the same graph."""
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", upd, aggregation="gossip",
                         topology=topology, topology_k=3, topology_seed=11, num_participants=1.0, epoch=1)
    for r in range(2):
        random.seed(7 + r)
        sel = random.sample(range(6), 6)
        ctl = RoundControl(r, sel, -1, [1.0] * 6, sel[0], -1)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 3e-5, f"gossip params r{r}")
        assert e_gpu.state.rejected.tolist() == e_ref.state.rejected.tolist()
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")


def test_gossip_csr_table_equals_device_topology(tiny_clients, tmp_path, dev):
    """An uploaded CSR table of the same ring gives bit-identical parameters to the in-kernel ring."""
    from fedmse_decentralized_b200.parallel import topology
    outs = []
    for use_table in (False, True):
        cfg = make_cfg(tmp_path, aggregation="gossip", topology="ring", num_participants=1.0, epoch=1)
        eng = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda", seed=3)
        if use_table:
            eng.backend.device_topology = False
        eng.run_round(0)
        outs.append(eng.state.params.clone())
    assert torch.equal(outs[0], outs[1])


def test_gossip_pipelined_rounds_are_bit_identical(tiny_clients, tmp_path, dev):
    res = []
    for pipe in (False, True):
        cfg = make_cfg(tmp_path, aggregation="gossip", topology="random-k", topology_k=2, num_participants=0.7, epoch=1)
        eng = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda", seed=3)
        out = [eng.run_round(r, pipeline_next=pipe and r < 3) for r in range(4)]
        res.append(([o.client_metrics for o in out], eng.state.params.clone(), [o.selected for o in out]))
    assert res[0][2] == res[1][2] and res[0][0] == res[1][0] and torch.equal(res[0][1], res[1][1])


@pytest.mark.parametrize("precision,rtol", [("tf32", 2e-2), ("tf32x3", 2e-4)])
def test_tensor_core_scoring_matches_oracle(tiny_clients, tmp_path, dev, precision, rtol):
    """tcgen05 scoring forward (TMA tiles, TMEM accumulators) against the fp32 oracle."""
    cfg = make_cfg(tmp_path, score_precision=precision)
    e_ref = make_engine(cfg, make_shards(tiny_clients, "cpu"), "hybrid", "avg", "oracle")
    e_gpu = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "avg", "cuda")
    be, st, sh = e_gpu.backend, e_gpu.state, e_gpu.sh
    assert be.tc in (1, 3) and be.tile == 128
    be.test_sse.fill_(-1.0); be.test_lat.fill_(-1.0)
    be._score(st.params, st.n_clients, 0, None, sh.test.x, be.tiles_test, be.test_sse, be.test_lat)
    torch.cuda.synchronize()
    for c in range(st.n_clients):
        x = e_ref.sh.test.rows(c)
        z, _ = oracle.forward(e_ref.state.params[c], x[:, :L.dim], L)
        sse = oracle.row_sse(e_ref.state.params[c], x, L)
        lo, hi = sh.test.off_host[c], sh.test.off_host[c + 1]
        assert_close(be.test_sse[lo:hi], sse, rtol, rtol * 10, f"sse client {c}")
        assert_close(be.test_lat[lo:hi, :7], z, rtol, rtol, f"latent client {c}")
    # latents-only mode (hybrid evaluation) and the AUC on top of it
    a_gpu, _, _ = be.evaluate(st, sh, "hybrid", False)
    _, a64, _ = e_ref.backend.evaluate(e_ref.state, e_ref.sh, "hybrid", False)
    assert_close(a_gpu, torch.tensor(a64), 0, 3e-3, "auc")


@pytest.mark.parametrize("precision,rtol", [("tf32", 2e-2), ("tf32x3", 2e-4)])
def test_multi_model_scoring_matches_oracle(tmp_path, dev, precision, rtol):
    """Many models on the pooled dev matrix (X parked in tensor memory, (tile, model-chunk) work units):
    every (model, row) squared error against the fp32 oracle, for full, small and single-model selections
    (the last two leave some model chunks empty)."""
    from fedmse_decentralized_b200.data import SyntheticSpec, make_clients, prepare_from_raw
    spec = SyntheticSpec(n_normal=2400, n_abnormal=40, n_test_normal=20, jitter_rows=10)
    clients = prepare_from_raw(make_clients(12, spec, seed=11), True, seed=5)
    cfg = make_cfg(tmp_path, epoch=1, network_size=12, score_precision=precision, write_latents=False,
                   save_client_models=False)
    eng = make_engine(cfg, make_shards(clients, dev), "hybrid", "mse_avg", "cuda")
    be, st, sh = eng.backend, eng.state, eng.sh
    R = be.dev_rows
    assert R > 6 * 128, "the test wants several tiles"
    st.params.mul_(1.0 + 0.05 * torch.randn_like(st.params))        # different models per client
    host = st.params.cpu()
    X = sh.dev.cpu()[:R]
    for sel in ([3, 0, 11, 7, 5, 1, 9, 2, 10, 4, 8, 6], [4, 9, 2], [6]):
        ctl = RoundControl(round=0, sel=list(sel), voter_pos=0, jitter=[1.0] * len(sel), vote_client=sel[0],
                           verify_client=-1)
        be.begin_round(ctl)
        be.merge_sse = None
        mse = be.merge_scores(st, sh, ctl)
        torch.cuda.synchronize()
        got = be.merge_sse[:len(sel) * R].view(len(sel), R)
        for p, c in enumerate(sel):
            want = oracle.row_sse(host[c], X, L)
            assert_close(got[p], want, rtol, rtol * 10, f"{precision} slot {p} client {c}")
            assert_close(mse[p], want.double().mean() / L.dim, rtol, 0, f"{precision} mse slot {p}")


@pytest.mark.parametrize("n_clients,rows", [(12, 2400), (3, 500), (40, 700)])
def test_fused_f16_scoring_matches_oracle(tmp_path, dev, n_clients, rows):
    """Second-generation MSE-weight scoring (fp16 head + remainder operands on tcgen05 kind::f16, fused 27x27
    bottleneck GEMM, in-kernel MSE reduction): per-model MSE against the fp32 oracle in double, for full, small
    and single-model selections and for unit lengths 1..n (short units exercise the tile hand-over)."""
    from fedmse_decentralized_b200.data import SyntheticSpec, make_clients, prepare_from_raw
    spec = SyntheticSpec(n_normal=rows, n_abnormal=40, n_test_normal=20, jitter_rows=10)
    clients = prepare_from_raw(make_clients(n_clients, spec, seed=11), True, seed=5)
    cfg = make_cfg(tmp_path, epoch=1, network_size=n_clients, score_precision="fp16x3", write_latents=False,
                   save_client_models=False)
    eng = make_engine(cfg, make_shards(clients, dev), "hybrid", "mse_avg", "cuda")
    be, st, sh = eng.backend, eng.state, eng.sh
    assert be.fused_f16
    R = be.dev_rows
    st.params.mul_(1.0 + 0.05 * torch.randn_like(st.params))        # different models per client
    host = st.params.cpu()
    X = sh.dev.cpu()[:R]
    full = list(range(n_clients))
    random.Random(3).shuffle(full)
    worst = 0.0
    for sel in (full, full[:3], full[:1], full[:2]):
        ctl = RoundControl(round=0, sel=list(sel), voter_pos=0, jitter=[1.0] * len(sel), vote_client=sel[0],
                           verify_client=-1)
        be.begin_round(ctl)
        for rep in range(2):                                         # twice: the grid barrier words are reused
            mse = be.merge_scores(st, sh, ctl)
            torch.cuda.synchronize()
            for p, c in enumerate(sel):
                want = float(oracle.row_sse(host[c].double(), X.double(), L).mean() / L.dim)
                rel = abs(float(mse[p]) - want) / want
                worst = max(worst, rel)
                assert rel < 1e-4, f"slot {p} client {c}: {float(mse[p])} vs {want} (rel {rel:.2e})"
    print(f"fused f16 scoring: worst relative MSE error {worst:.2e}")


@pytest.mark.parametrize("rows", [40, 300])
def test_kde_js_kernel_matches_fp64_oracle(tiny_clients, tmp_path, dev, rows):
    """T6 (csrc/similarity_kernel.cu): KDE self log-density + Jensen-Shannon distance of every selected model's
    reconstruction of the dev rows, against the fp64 torch form of the reference's weight function."""
    from fedmse_decentralized_b200.utils.similarity import fusion_distances
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "fusion_avg", update_types=["fusion_avg"],
                         fusion_rows=rows, num_participants=0.7)
    ctl = e_ref.make_control(0)
    e_gpu.backend.begin_round(ctl)
    for e in (e_ref, e_gpu):
        e.backend.train(e.state, e.sh, e._sel_local(ctl.sel), e.hp)
    got = e_gpu.backend.fusion_scores(e_gpu.state, e_gpu.sh, ctl, rows)
    torch.cuda.synchronize()
    n = min(rows, e_gpu.sh.dev.shape[0])
    want = fusion_distances(e_gpu.state.params[ctl.sel].double().float(), e_gpu.sh.dev, L, n)
    assert got.shape == want.shape and bool(torch.isfinite(got).all()) and float(got.min()) > 0
    assert_close(got, want, 2e-3, 1e-6, "JS distances")
    w_got, w_want = (1 / got) / (1 / got).sum(), (1 / want) / (1 / want).sum()
    assert_close(w_got, w_want, 2e-3, 1e-6, "fusion_avg weights")


def test_full_round_with_fusion_avg(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "fusion_avg", update_types=["fusion_avg"],
                         fusion_rows=64, num_participants=0.7)
    for r in range(2):
        random.seed(70 + r); torch.manual_seed(70 + r)
        ctl = e_ref.make_control(r)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert res_gpu.aggregator == res_ref.aggregator
        # the fp32 kernel's JS distances agree with the fp64 oracle to ~1e-3 relative, and so do the merge weights
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 2e-4, f"params r{r}")
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")


def test_full_round_with_fused_f16_scoring(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "hybrid", "mse_avg", score_precision="fp16x3",
                         num_participants=0.7)
    for r in range(2):
        random.seed(60 + r); torch.manual_seed(60 + r)
        ctl = e_ref.make_control(r)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert res_gpu.aggregator == res_ref.aggregator
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 3e-5, f"params r{r}")
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")


def test_full_round_with_tensor_core_scoring(tiny_clients, tmp_path, dev):
    e_ref, e_gpu = _pair(tiny_clients, tmp_path, dev, "autoencoder", "mse_avg", score_precision="tf32x3",
                         num_participants=0.7)
    for r in range(2):
        random.seed(50 + r); torch.manual_seed(50 + r)
        ctl = e_ref.make_control(r)
        res_ref = e_ref.run_round(r, copy.deepcopy(ctl))
        res_gpu = e_gpu.run_round(r, copy.deepcopy(ctl))
        assert res_gpu.aggregator == res_ref.aggregator
        assert_close(e_gpu.state.params, e_ref.state.params, 3e-3, 3e-5, f"params r{r}")
        assert_close(torch.tensor(res_gpu.client_metrics), torch.tensor(res_ref.client_metrics), 0, 2e-3, "auc")


def test_pipelined_rounds_match_plain_rounds(tiny_clients, tmp_path, dev):
    """run_round(..., pipeline_next=True) enqueues the next round's upload + training before it reads its own
    results back; the sequence of results and the final state must be identical to the plain loop."""
    engines = []
    for _ in range(2):
        cfg = make_cfg(tmp_path, epoch=2, score_precision="tf32x3", num_participants=0.7)
        random.seed(11); torch.manual_seed(11)
        engines.append(make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda"))
    plain, piped = engines
    random.seed(21); torch.manual_seed(21)
    res_plain = [plain.run_round(r, prefetch_next=True) for r in range(5)]
    random.seed(21); torch.manual_seed(21)
    res_piped = [piped.run_round(r, pipeline_next=r < 4) for r in range(5)]
    assert piped._pre is None
    for a, b in zip(res_plain, res_piped):
        assert (a.selected, a.aggregator, a.epochs_run, a.samples) == (b.selected, b.aggregator, b.epochs_run, b.samples)
        assert a.client_metrics == b.client_metrics
        assert a.verification == b.verification
    assert torch.equal(plain.state.params, piped.state.params)
    assert torch.equal(plain.state.adam_m, piped.state.adam_m)
    with pytest.raises(RuntimeError):
        piped.run_round(5, pipeline_next=True)
        piped.run_round(7)


def test_staged_upload_swaps_shards_and_refreshes_residuals(tiny_clients, tmp_path, dev):
    cfg = make_cfg(tmp_path, epoch=1, score_precision="tf32x3")
    eng = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda")
    r0 = eng.run_round(0)
    host_train = (eng.sh.train.x * 0.9).cpu().pin_memory()
    v0, v1 = eng.sh.valid.off_host[0], eng.sh.valid.off_host[6]
    host_valid = eng.sh.valid.x[v0:v1].cpu().pin_memory()
    for r in range(1, 3):
        eng.stage_next_data(host_train, host_valid)
        eng.commit_staged_data()
        res = eng.run_round(r)
    torch.cuda.synchronize()
    assert torch.equal(eng.sh.train.x.cpu(), host_train) and len(res.client_metrics) == 6
    lo = eng.backend._resid.get(eng.sh.train.x.data_ptr())
    if lo is not None:       # residual shard belongs to the *current* buffer and matches its contents
        x = eng.sh.train.x
        hi = (x.view(torch.int32) & -8192).view(torch.float32)
        assert torch.equal(lo, x - hi)


def test_checkpoint_file_resumes_identically_on_the_device(tiny_clients, tmp_path, dev):
    """save_checkpoint / load_checkpoint with the CUDA backend: device-resident state round-trips through the file."""
    cfg = make_cfg(tmp_path, epoch=1, score_precision="fp16x3")
    a = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda")
    a.run_round(0)
    path = a.save_checkpoint(str(tmp_path / "engine.pt"))
    ra = [a.run_round(r) for r in (1, 2)]
    b = make_engine(cfg, make_shards(tiny_clients, dev), "hybrid", "mse_avg", "cuda", seed=31)
    assert b.load_checkpoint(path) == 1
    rb = [b.run_round(r) for r in (1, 2)]
    for x, y in zip(ra, rb):
        assert (x.selected, x.aggregator, x.client_metrics) == (y.selected, y.aggregator, y.client_metrics)
    assert torch.equal(a.state.params, b.state.params) and torch.equal(a.state.adam_v, b.state.adam_v)
