import json
import os
import pickle
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_main_cli_end_to_end_on_synthetic_scenario(tmp_path):
    """`python main.py` with flags: scenario JSON on disk (reference layout), oracle backend, full artefact tree."""
    sys.path.insert(0, ROOT)
    from fedmse_decentralized_b200.data import SyntheticSpec, write_scenario
    spec = SyntheticSpec(n_normal=150, n_abnormal=100, n_test_normal=30, jitter_rows=5)
    scen = write_scenario(str(tmp_path / "Data"), 4, spec, seed=2)
    (tmp_path / "scen.json").write_text(json.dumps(scen))
    cmd = [sys.executable, os.path.join(ROOT, "main.py"), "--config_file", str(tmp_path / "scen.json"),
           "--network_size", "4", "--epoch", "1", "--num_rounds", "2", "--backend", "oracle", "--model_types", "hybrid",
           "--update_types", "avg", "mse_avg", "--output_root", str(tmp_path / "out"), "--log_level", "WARNING"]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=dict(os.environ, FEDMSE_LR_RATE="0.002"))
    assert p.returncode == 0, p.stderr[-2000:]
    summ = [os.path.join(dp, f) for dp, _, fs in os.walk(tmp_path / "out") for f in fs if f == "training_summary.json"]
    assert len(summ) == 1
    s = json.load(open(summ[0]))
    assert set(s["best_metrics"]["hybrid"]) == {"avg", "mse_avg"} and "lr0.002" in s["experiment_name"]
    lat = [os.path.join(dp, f) for dp, _, fs in os.walk(tmp_path / "out") for f in fs if f.startswith("latent_hybrid_")]
    assert len(lat) == 2
    d = pickle.load(open(lat[0], "rb"))
    assert sorted(d.keys())[0] == 0 and len(next(iter(d.values()))) == 4
    q = subprocess.run([sys.executable, os.path.join(ROOT, "scripts/plot_results.py"), "results", str(tmp_path / "out"),
                        "--out", str(tmp_path / "auc.pdf")], capture_output=True, text=True, timeout=300)
    assert q.returncode == 0 and "mse_avg" in q.stdout
