"""bench.py plumbing that can run without a GPU: the --check mode on the oracle backend and the reference arm's
on-demand install."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_check_mode_on_the_oracle_backend():
    """`bench.py --check` builds the scenario twice (N ranks / one rank), runs the pipelined loop with staged uploads
    and compares every discrete outcome; on CPU both arms are the torch oracle, so they must agree exactly."""
    p = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--check", "--backend", "oracle",
                        "--clients-per-gpu", "4", "--epochs", "1", "--check-rounds", "2"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert p.returncode == 0, p.stderr[-2000:]
    line = [l for l in p.stdout.splitlines() if l.startswith("{")][-1]
    v = json.loads(line)
    assert v["check"] == "pass" and v["clients"] == 4 and v["arms"]["p2p"]["mismatches"] == []
    assert v["arms"]["p2p"]["max_abs_auc_diff_strict_rounds"] == 0.0


def test_reference_arm_reports_unavailable_or_runs(tmp_path):
    """--impl reference either finds / installs baseline/_ref or says why not, as one JSON line, exit code 0."""
    sys.path.insert(0, os.path.join(ROOT, "baseline"))
    import ref_harness
    why = ref_harness.ensure_installed()
    if os.path.isdir("/root/reference/src/Trainer"):
        assert why == "" and os.path.isfile(os.path.join(ROOT, "baseline", "_ref", "Trainer", "client_trainer.py"))
    else:
        assert "missing" in why or "failed" in why
