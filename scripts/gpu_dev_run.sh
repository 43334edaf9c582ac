#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 200 python -m pytest tests -m gpu -x -q --timeout 200 -k "fused_f16_scoring" > gpurun_out/t8_tests.log 2>&1; tail -3 gpurun_out/t8_tests.log
for mult in 1 8; do
  echo "score phase: dev rows x$mult, fp16x3"; timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 4 --dev-rows-mult $mult 2>&1 | tail -1
  timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult $mult --phase-cycles 2>&1 | tail -2
done
