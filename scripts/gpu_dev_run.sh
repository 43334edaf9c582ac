#!/usr/bin/env bash
# Development helper run on the GPU box (1 GPU).
mkdir -p gpurun_out
timeout 400 python -m pytest tests -m gpu -x -q --timeout 200 -k "generic or fused or batch_64" > gpurun_out/t2_tests.log 2>&1; tail -6 gpurun_out/t2_tests.log
for prec in fp16x3 tf32x3; do
  timeout 300 python bench.py --steps 20 --warmup 5 --score-precision $prec > gpurun_out/t2_bench_$prec.json 2> gpurun_out/t2_bench_$prec.err
  python - <<PY
import json
d = json.load(open("gpurun_out/t2_bench_$prec.json"))
print("$prec", "ms/step", round(d["ms_per_step"], 4), "e2e ms", round(d["e2e"]["ms_per_step"], 4), "phases", d["phases_ms"], "auc", d["auc_mean_last_round"])
PY
done
for mult in 1 8; do
  for prec in fp16x3 tf32x3; do
    echo "score phase: dev rows x$mult, $prec"; timeout 200 python scripts/prof_kernels.py score --score-precision $prec --iters 4 --dev-rows-mult $mult 2>&1 | tail -2
  done
done
python scripts/prof_kernels.py score --score-precision fp16x3 --iters 2 --dev-rows-mult 8 > gpurun_out/t2_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:score_fused -s 1 -c 1 -o gpurun_out/prof_fused_f16 \
    python scripts/prof_kernels.py score --score-precision fp16x3 --iters 2 --dev-rows-mult 8 > gpurun_out/t2_ncu.log 2>&1
tail -3 gpurun_out/t2_ncu.log
