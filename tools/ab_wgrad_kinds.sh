#!/bin/bash
# per-kind event times of the training step, previous vs current library (see ab_wgrad_out.sh)
# needs $P/libipoke_inn_prev.so: the library built from the commit to compare against (nvcc command of build.py on
# that commit's csrc/, output name libipoke_inn_prev.so); *.so files are git-ignored but travel with gpurun
P=ipoke-main-opticalflow_b200
[ -f $P/libipoke_inn_prev.so ] || { echo "build $P/libipoke_inn_prev.so first"; exit 1; }
cp $P/libipoke_inn.so $P/libipoke_inn_new.so
for v in prev new; do
  cp $P/libipoke_inn_$v.so $P/libipoke_inn.so
  for b in 64 512; do
    timeout 100 python tools/train_bench.py --config cfgA --batch $b --reps 20 --tag $v 2>&1 | grep -v Warn | tail -3
  done
done
cp $P/libipoke_inn_new.so $P/libipoke_inn.so
