#!/usr/bin/env bash
# Development helper run on the GPU box with 2 GPUs: multi-GPU tests, NVLS, gossip, --check, bandwidth sweep.
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577"
timeout 900 python -m pytest tests/test_multi_gpu.py -m gpu -x -q --timeout 600 > gpurun_out/t3_multi.log 2>&1; tail -15 gpurun_out/t3_multi.log
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), d["phases_ms"], "auc", round(d["auc_mean_last_round"], 5), d["config"].get("merge_path"))
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
}
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/t3_elected_n2.json 2> gpurun_out/t3_elected_n2.err; show gpurun_out/t3_elected_n2.json; tail -3 gpurun_out/t3_elected_n2.err
FEDMSE_NVLS=0 timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/t3_elected_n2_p2p.json 2> gpurun_out/t3_elected_n2_p2p.err; show gpurun_out/t3_elected_n2_p2p.json
timeout 300 $TR bench.py --gpus 2 --steps 20 --warmup 5 --aggregation gossip > gpurun_out/t3_gossip_n2.json 2> gpurun_out/t3_gossip_n2.err; show gpurun_out/t3_gossip_n2.json; tail -3 gpurun_out/t3_gossip_n2.err
timeout 400 $TR bench.py --gpus 2 --check > gpurun_out/t3_check_elected.json 2> gpurun_out/t3_check_elected.err; cat gpurun_out/t3_check_elected.json; tail -3 gpurun_out/t3_check_elected.err
timeout 400 $TR bench.py --gpus 2 --check --aggregation gossip > gpurun_out/t3_check_gossip.json 2> gpurun_out/t3_check_gossip.err; cat gpurun_out/t3_check_gossip.json; tail -3 gpurun_out/t3_check_gossip.err
timeout 400 $TR scripts/bench_bandwidth_sweep.py --sizes-mb 1 16 128 512 > gpurun_out/t3_sweep_n2.jsonl 2> gpurun_out/t3_sweep.err; cat gpurun_out/t3_sweep_n2.jsonl; tail -3 gpurun_out/t3_sweep.err
