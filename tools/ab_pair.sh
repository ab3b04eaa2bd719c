for cfg in cfgB cfgA; do for b in 32768 4096; do for o in 0 1; do
IPOKE_INN_PAIR=$o timeout 100 python tools/small_bench.py --config $cfg --batch $b --reps 10 --tag "pair=$o" 2>&1 | tail -1
done; done; done
