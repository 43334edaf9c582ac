#!/usr/bin/env bash
# 8-GPU run: scaling bench (elected + gossip), --check at 1000 clients, configs 2-5.
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29588"
show() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    print(sys.argv[1], "ms/step", round(d["ms_per_step"], 4), "e2e", round(d["e2e"]["ms_per_step"], 4), d["phases_ms"], "auc", round(d["auc_mean_last_round"], 5), d["config"].get("merge_path"), "samples/s", round(d["value"] / 1e6, 1), "M")
except Exception as e:
    print(sys.argv[1], "unreadable:", e)
PY
}
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/s_elected_n$N.json 2> gpurun_out/s_elected_n$N.err; show gpurun_out/s_elected_n$N.json; tail -2 gpurun_out/s_elected_n$N.err
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --aggregation gossip > gpurun_out/s_gossip_n$N.json 2> gpurun_out/s_gossip_n$N.err; show gpurun_out/s_gossip_n$N.json; tail -2 gpurun_out/s_gossip_n$N.err
timeout 600 $TR bench.py --gpus $N --check > gpurun_out/s_check_elected_n$N.json 2> gpurun_out/s_check_elected_n$N.err; cat gpurun_out/s_check_elected_n$N.json; tail -2 gpurun_out/s_check_elected_n$N.err
timeout 400 $TR scripts/bench_bandwidth_sweep.py --sizes-mb 1 64 512 > gpurun_out/s_sweep_n$N.jsonl 2> gpurun_out/s_sweep_n$N.err; cat gpurun_out/s_sweep_n$N.jsonl; tail -2 gpurun_out/s_sweep_n$N.err
timeout 300 $TR scripts/run_configs.py 3 > gpurun_out/s_config3_n$N.jsonl 2> gpurun_out/s_config3_n$N.err; cat gpurun_out/s_config3_n$N.jsonl; tail -2 gpurun_out/s_config3_n$N.err
timeout 300 $TR scripts/run_configs.py 2 --data_root data_cache/N-BaIoT/IID-10-Client_Data > gpurun_out/s_config2_n$N.jsonl 2> gpurun_out/s_config2_n$N.err; cat gpurun_out/s_config2_n$N.jsonl; tail -2 gpurun_out/s_config2_n$N.err
