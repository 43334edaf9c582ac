#!/usr/bin/env bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_cuda_kernels.py -x -q -m gpu -k "kde or fusion" 2>&1 | tail -15
