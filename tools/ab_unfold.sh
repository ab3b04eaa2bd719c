for cfg in cfgB cfgA; do for b in 32768 4096; do
for opt in "X=0" "IPOKE_INN_FOLD_MAXB=1000" "IPOKE_INN_FOLD_MAXB=1000 IPOKE_INN_FASTB=0"; do
env $opt timeout 100 python tools/small_bench.py --config $cfg --batch $b --reps 10 --tag "$opt" 2>&1 | tail -1
done; done; done
