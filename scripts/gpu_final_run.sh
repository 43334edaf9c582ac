#!/usr/bin/env bash
# Final 1-GPU pass: the whole GPU test suite, smoke(), the flagship bench (both aggregation modes), ncu of the training kernel.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 > gpurun_out/f_tests.log 2>&1; tail -4 gpurun_out/f_tests.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/f_bench_elected_n1.json 2> gpurun_out/f_bench_elected_n1.err
timeout 300 python bench.py --steps 20 --warmup 5 --aggregation gossip > gpurun_out/f_bench_gossip_n1.json 2> gpurun_out/f_bench_gossip_n1.err
python - <<'PY'
import json
for f in ("gpurun_out/f_bench_elected_n1.json", "gpurun_out/f_bench_gossip_n1.json"):
    d = json.loads([l for l in open(f) if l.startswith("{")][-1])
    print(f, "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), d["phases_ms"], "auc", round(d["auc_mean_last_round"], 5), "launches", d["gpu_launches"])
PY
python scripts/prof_kernels.py train --iters 2 > gpurun_out/f_plain_train.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:train_clients_kernel -s 1 -c 1 -o gpurun_out/prof_train_final \
    python scripts/prof_kernels.py train --iters 2 > gpurun_out/f_ncu_train.log 2>&1
tail -2 gpurun_out/f_ncu_train.log
