for b in 1024 2048 8192 16384; do
for opt in "X=0" "IPOKE_INN_FOLD_MAXB=1000"; do
env $opt timeout 100 python tools/small_bench.py --config cfgB --batch $b --reps 20 --tag "$opt" 2>&1 | tail -1
done; done
