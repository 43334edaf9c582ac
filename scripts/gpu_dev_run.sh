#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult 8 2>&1 | tail -3
timeout 400 ncu --set full --clock-control none --import-source on -k regex:score_fused_f16 -s 2 -c 1 -o gpurun_out/prof_fused_v8 \
    python scripts/prof_kernels.py score --score-precision fp16x3 --iters 2 --dev-rows-mult 8 > gpurun_out/ncu_fused_v8.log 2>&1
tail -2 gpurun_out/ncu_fused_v8.log
