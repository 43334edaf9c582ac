#!/usr/bin/env bash
# Development helper run on the GPU box: the regular GPU tests (without the fused f16 scoring tests), the gossip
# bench, then the fused f16 tests in both half-packing orders under a short timeout.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q --timeout 300 -k "not fused_f16" > gpurun_out/t1_tests.log 2>&1; tail -8 gpurun_out/t1_tests.log
timeout 300 python bench.py --steps 20 --warmup 5 --aggregation gossip > gpurun_out/t1_gossip_n1.json 2> gpurun_out/t1_gossip_n1.err; tail -c 900 gpurun_out/t1_gossip_n1.json; tail -3 gpurun_out/t1_gossip_n1.err
for sw in 0 1; do
  FM_F16_SWAP=$sw timeout 150 python -m pytest tests/test_cuda_kernels.py -m gpu -x -q -s -k "fused_f16_scoring_matches" > gpurun_out/t1_fused_swap$sw.log 2>&1
  echo "swap=$sw rc=$?"; tail -12 gpurun_out/t1_fused_swap$sw.log
done
