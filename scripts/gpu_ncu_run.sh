#!/usr/bin/env bash
# ncu --set full captures of the kernels of one elected round and one gossip round (1 GPU), after a plain run.
mkdir -p gpurun_out
K='regex:train_clients_kernel|merge_kernel|verify_kernel|auc_kernel|score_rows_tc_kernel|score_rows_kernel|elect_kernel|chunk_mean|vote_normalise'
python scripts/prof_kernels.py round --score-precision fp16x3 --iters 2 > gpurun_out/n1_plain_round.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "$K" -s 12 -c 12 -o gpurun_out/prof_round_elected \
    python scripts/prof_kernels.py round --score-precision fp16x3 --iters 2 > gpurun_out/n1_ncu_round.log 2>&1
tail -2 gpurun_out/n1_ncu_round.log
python scripts/prof_kernels.py round --score-precision fp16x3 --iters 2 --aggregation gossip > gpurun_out/n1_plain_gossip.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:gossip_kernel|gossip_publish' -s 2 -c 2 -o gpurun_out/prof_round_gossip \
    python scripts/prof_kernels.py round --score-precision fp16x3 --iters 2 --aggregation gossip > gpurun_out/n1_ncu_gossip.log 2>&1
tail -2 gpurun_out/n1_ncu_gossip.log
python scripts/prof_kernels.py train --iters 2 --phase-cycles 2>&1 | tail -6
