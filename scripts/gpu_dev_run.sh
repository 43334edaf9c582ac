#!/usr/bin/env bash
mkdir -p gpurun_out
for m in 0 7 1; do
echo "probe mode $m"
timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult 8 --phase-cycles --probe-mode $m 2>&1 | grep -E "fused f16|inside" | tail -2 | cut -c1-700
done
