#!/usr/bin/env bash
# N-GPU refresh of the scaling numbers: flagship bench, both aggregation modes (and, with CHECK=1, the gossip --check).
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
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/v8_elected_n$N.json 2> gpurun_out/v8_elected_n$N.err; show gpurun_out/v8_elected_n$N.json; tail -2 gpurun_out/v8_elected_n$N.err
timeout 400 $TR bench.py --gpus $N --steps 20 --warmup 5 --aggregation gossip > gpurun_out/v8_gossip_n$N.json 2> gpurun_out/v8_gossip_n$N.err; show gpurun_out/v8_gossip_n$N.json; tail -2 gpurun_out/v8_gossip_n$N.err
if [ "${CHECK:-0}" = "1" ]; then
  timeout 600 $TR bench.py --gpus $N --check --aggregation gossip > gpurun_out/v8_check_gossip_n$N.json 2> gpurun_out/v8_check_gossip_n$N.err; cat gpurun_out/v8_check_gossip_n$N.json; tail -2 gpurun_out/v8_check_gossip_n$N.err
fi
