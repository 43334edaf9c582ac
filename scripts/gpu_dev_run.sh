#!/usr/bin/env bash
# Scratch runner for kernel work: the fused scoring kernel's tests, then its time on the 1-GPU and the 8-GPU-share workload.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_kernels.py -x -q -m gpu -k "fused" 2>&1 | tail -2
timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 4 2>&1 | tail -2
timeout 200 python scripts/prof_kernels.py score --score-precision fp16x3 --iters 4 --dev-rows-mult 8 2>&1 | tail -2
