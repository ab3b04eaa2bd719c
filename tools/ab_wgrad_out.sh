#!/bin/bash
# same-box A/B of the weight-gradient write-out: libipoke_inn_prev.so (scalar, padded staging) vs the current library
# (XOR-swizzled staging, 128-bit shared reads and global stores).  Runs on the GPU box's scratch copy of the repo.
P=ipoke-main-opticalflow_b200
cp $P/libipoke_inn.so $P/libipoke_inn_new.so
for v in prev new prev new; do
  cp $P/libipoke_inn_$v.so $P/libipoke_inn.so
  for b in 64 4096; do
    timeout 100 python tools/train_bench.py --config cfgA --batch $b --reps 30 --tag $v 2>&1 | grep "^$v"
  done
done
cp $P/libipoke_inn_new.so $P/libipoke_inn.so
