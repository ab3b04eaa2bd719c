#!/bin/bash
# wide-K coverage rule of the hidden GEMMs (cfg A, Hp = 1536): old 80 % rule (IPOKE_INN_WIDEK_MT=0) vs the new default,
# and whether extending it to more row tiles helps (IPOKE_INN_WIDEK_MT=8)
for b in 64 128 256; do
  for mt in 0 2; do
    IPOKE_INN_WIDEK_MT=$mt timeout 100 python tools/small_bench.py --config cfgA --batch $b --reps 100 --tag "widek_mt=$mt" 2>&1 | tail -1
  done
done
for mt in 0 2; do
  IPOKE_INN_WIDEK_MT=$mt IPOKE_INN_STREAMS=2 timeout 100 python tools/small_bench.py --config cfgA --batch 256 --reps 100 --dual --tag "widek_mt=$mt dual" 2>&1 | tail -1
done
for b in 512 1024; do
  for mt in 2 8; do
    IPOKE_INN_WIDEK_MT=$mt timeout 100 python tools/small_bench.py --config cfgA --batch $b --reps 50 --tag "widek_mt=$mt" 2>&1 | tail -1
  done
done
for mt in 0 2; do IPOKE_INN_WIDEK_MT=$mt timeout 100 python tools/train_bench.py --config cfgA --batch 64 --reps 30 --tag "train widek_mt=$mt" 2>&1 | grep "^train"; done
