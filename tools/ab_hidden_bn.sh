#!/bin/bash
# hidden-layer tile width at small batch for the wide-hidden config (cfg A, H = 1536): default heuristic vs forced widths
for b in 64 256; do
  for bn in 0 16 32 64 128; do
    IPOKE_INN_HIDDEN_BN=$bn timeout 100 python tools/small_bench.py --config cfgA --batch $b --reps 100 --tag "hidden_bn=$bn" 2>&1 | tail -1
  done
done
