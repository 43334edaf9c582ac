#!/usr/bin/env bash
# Development helper run on the GPU box (1 GPU).
mkdir -p gpurun_out
timeout 300 python -m pytest tests -m gpu -x -q --timeout 200 -k "fused" > gpurun_out/t7_tests.log 2>&1; tail -4 gpurun_out/t7_tests.log
for mult in 1 8; do
  echo "score phase: dev rows x$mult, fp16x3"; timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 4 --dev-rows-mult $mult 2>&1 | tail -2
done
timeout 300 python bench.py --steps 20 --warmup 5 > gpurun_out/t7_bench_fp16x3.json 2> gpurun_out/t7_bench.err
python - <<PY
import json
d = json.load(open("gpurun_out/t7_bench_fp16x3.json"))
print("bench ms/step", round(d["ms_per_step"], 4), "e2e ms", round(d["e2e"]["ms_per_step"], 4), "phases", d["phases_ms"], "auc", d["auc_mean_last_round"])
PY
python scripts/prof_kernels.py score --score-precision fp16x3 --iters 2 --dev-rows-mult 8 > gpurun_out/t7_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_fused -s 1 -c 1 -o gpurun_out/prof_fused_f16_v5 \
    python scripts/prof_kernels.py score --score-precision fp16x3 --iters 2 --dev-rows-mult 8 > gpurun_out/t7_ncu.log 2>&1
tail -2 gpurun_out/t7_ncu.log
