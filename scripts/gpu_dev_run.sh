#!/usr/bin/env bash
mkdir -p gpurun_out
echo "== hybrid mse_avg random-k (baseline)"; timeout 300 python scripts/gossip_dynamics.py --rounds 21 2>&1 | tail -8
echo "== hybrid avg"; timeout 300 python scripts/gossip_dynamics.py --rounds 21 --update avg 2>&1 | tail -4
echo "== autoencoder mse_avg"; timeout 300 python scripts/gossip_dynamics.py --rounds 21 --model autoencoder 2>&1 | tail -4
echo "== hybrid mse_avg, no gossip merging: elected"; timeout 300 python scripts/gossip_dynamics.py --rounds 21 --aggregation elected 2>&1 | tail -4
echo "== hybrid mse_avg participation 1.0"; timeout 300 python scripts/gossip_dynamics.py --rounds 21 --participation 1.0 2>&1 | tail -4
