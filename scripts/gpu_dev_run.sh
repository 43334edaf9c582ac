#!/usr/bin/env bash
mkdir -p gpurun_out
for cpg in 250 1000; do
timeout 600 python bench.py --steps 10 --warmup 3 --aggregation gossip --clients-per-gpu $cpg > gpurun_out/g_gossip_${cpg}_n1.json 2> gpurun_out/g_gossip_${cpg}_n1.err
python - <<PY
import json
d = json.loads([l for l in open("gpurun_out/g_gossip_${cpg}_n1.json") if l.startswith("{")][-1])
print("gossip 1 GPU, $cpg clients: ms/step", round(d["ms_per_step"], 3), "auc", round(d["auc_mean_last_round"], 5), d["phases_ms"])
PY
done
timeout 600 python bench.py --check --aggregation gossip --clients-per-gpu 1000 --check-rounds 6 > gpurun_out/g_check_gossip_1000_n1.json 2>&1; tail -1 gpurun_out/g_check_gossip_1000_n1.json | cut -c1-900
