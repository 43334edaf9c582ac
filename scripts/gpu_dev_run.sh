#!/usr/bin/env bash
mkdir -p gpurun_out
for m in 0 1 7 11 13 3 5 9; do
  echo "probe mode $m"; timeout 100 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult 8 --phase-cycles --probe-mode $m 2>&1 | grep -E "score iter 2|fused f16 kernel" | cut -c1-900
done
