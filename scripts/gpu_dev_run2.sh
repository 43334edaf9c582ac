#!/usr/bin/env bash
# Development helper run on the GPU box with 2 GPUs.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577"
timeout 600 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --timeout 600 -k "quota or pipelined_staged or p2p_weighted" > gpurun_out/t9_multi.log 2>&1; tail -6 gpurun_out/t9_multi.log
timeout 400 $TR bench.py --gpus 2 --check > gpurun_out/t9_check_elected.json 2> gpurun_out/t9_check_elected.err; cat gpurun_out/t9_check_elected.json
timeout 400 $TR bench.py --gpus 2 --check --aggregation gossip > gpurun_out/t9_check_gossip.json 2> gpurun_out/t9_check_gossip.err; cat gpurun_out/t9_check_gossip.json
