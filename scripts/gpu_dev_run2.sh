#!/usr/bin/env bash
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577"
timeout 600 $TR bench.py --gpus 2 --check > gpurun_out/t10_check_elected.json 2> gpurun_out/t10_check_elected.err; cat gpurun_out/t10_check_elected.json; tail -2 gpurun_out/t10_check_elected.err
