#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q -s --timeout 200 -k "fused" > gpurun_out/t11_tests.log 2>&1; grep -E "worst relative|passed|failed|Error" gpurun_out/t11_tests.log | tail -6
for mult in 1 8; do
  echo "score phase: dev rows x$mult, fp16x3"; timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 4 --dev-rows-mult $mult 2>&1 | tail -1
  timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult $mult --phase-cycles 2>&1 | tail -1
done
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/t11_bench.json 2> gpurun_out/t11_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/t11_bench.json"))
print("bench ms/step", round(d["ms_per_step"], 4), "e2e ms", round(d["e2e"]["ms_per_step"], 4), "phases", d["phases_ms"], "auc", d["auc_mean_last_round"])
PY
