#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 3 --dev-rows-mult 8 --phase-cycles 2>&1 | tail -24 | cut -c1-400
