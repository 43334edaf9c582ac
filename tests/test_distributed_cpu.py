"""Multi-process host logic on CPU (gloo, world size 2): must reproduce the single-process run."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import json, os, random, sys
import numpy as np, torch
sys.path.insert(0, os.environ["FM_ROOT"])
from fedmse_decentralized_b200 import ExperimentConfig
from fedmse_decentralized_b200.data import SyntheticSpec, make_clients, prepare_from_raw
from fedmse_decentralized_b200.parallel import init_distributed, shutdown
from fedmse_decentralized_b200.trainer import FederationEngine, build_shards
comm = init_distributed("gloo")
spec = SyntheticSpec(n_normal=200, n_abnormal=150, n_test_normal=40, jitter_rows=6)
clients, dev = prepare_from_raw(make_clients(5, spec, seed=4), True, seed=2)
sh, names = build_shards(clients, dev, comm, torch.device("cpu"))
random.seed(1); np.random.seed(1); torch.manual_seed(1)
cfg = ExperimentConfig(epoch=1, num_rounds=2, network_size=5, backend="oracle", output_root=os.environ["FM_OUT"],
                       aggregation=os.environ.get("FM_AGG", "elected"), topology="ring", num_participants=0.8,
                       write_latents=True, save_client_models=False, fusion_rows=48,
                       update_types=[os.environ.get("FM_UPD", "mse_avg")])
eng = FederationEngine(cfg, sh, "hybrid", os.environ.get("FM_UPD", "mse_avg"), comm=comm, names_global=names)
if os.environ.get("FM_CKPT") == "1":       # round 0, checkpoint, a NEW engine resumes from the per-rank files
    first = eng.run_round(0)
    path = eng.save_checkpoint(os.path.join(os.environ["FM_OUT"], "ck", "engine.pt"))
    assert (comm.world == 1) == path.endswith("engine.pt")
    random.seed(77); np.random.seed(77); torch.manual_seed(77)
    eng2 = FederationEngine(cfg, sh, "hybrid", os.environ.get("FM_UPD", "mse_avg"), comm=comm, names_global=names)
    assert eng2.load_checkpoint(os.path.join(os.environ["FM_OUT"], "ck", "engine.pt")) == 1
    out = [first, eng2.run_round(1)]
    eng = eng2
else:
    out = [eng.run_round(r) for r in range(2)]
if comm.rank == 0:
    print("RESULT " + json.dumps({"metrics": [r.client_metrics for r in out], "agg": [r.aggregator for r in out],
                                  "ver": [[v["is_verified"] for v in r.verification] for r in out],
                                  "lat": sorted(eng.latents[min(eng.latents)].keys())}))
shutdown()
'''


def _run(world, tmp_path, agg, upd="mse_avg", ckpt=False):
    env = dict(os.environ, FM_ROOT=ROOT, FM_OUT=str(tmp_path), FM_AGG=agg, FM_UPD=upd, MASTER_ADDR="127.0.0.1",
               FM_CKPT="1" if ckpt else "0")
    tmp_path.mkdir(parents=True, exist_ok=True)
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    if world == 1:
        cmd = [sys.executable, str(script)]
    else:
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
               "--master-addr", "127.0.0.1", "--master-port", "29617", str(script)]
    p = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stderr[-3000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("RESULT ")][-1]
    return json.loads(line[7:])


@pytest.mark.parametrize("agg,upd", [("elected", "mse_avg"), ("gossip", "mse_avg"), ("elected", "fusion_avg")])
def test_two_ranks_match_one(tmp_path, agg, upd):
    one = _run(1, tmp_path, agg, upd)
    two = _run(2, tmp_path, agg, upd)
    assert one["agg"] == two["agg"] and one["ver"] == two["ver"] and one["lat"] == two["lat"]
    for a, b in zip(one["metrics"], two["metrics"]):
        assert len(a) == 5 and max(abs(x - y) for x, y in zip(a, b)) < 1e-9


def test_two_ranks_resume_from_per_rank_checkpoints(tmp_path):
    """Every rank writes / reads its own checkpoint file; the resumed two-rank run equals the uninterrupted one-rank run."""
    one = _run(1, tmp_path / "a", "elected")
    two = _run(2, tmp_path / "b", "elected", ckpt=True)
    assert one["agg"] == two["agg"] and one["ver"] == two["ver"]
    for a, b in zip(one["metrics"], two["metrics"]):
        assert max(abs(x - y) for x, y in zip(a, b)) < 1e-9
