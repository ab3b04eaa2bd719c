#!/bin/bash
# per-kind event times of the training step, previous vs current library (see ab_wgrad_out.sh)
P=ipoke-main-opticalflow_b200
cp $P/libipoke_inn.so $P/libipoke_inn_new.so
for v in prev new; do
  cp $P/libipoke_inn_$v.so $P/libipoke_inn.so
  for b in 64 512; do
    timeout 100 python tools/train_bench.py --config cfgA --batch $b --reps 20 --tag $v 2>&1 | grep -v Warn | tail -3
  done
done
cp $P/libipoke_inn_new.so $P/libipoke_inn.so
