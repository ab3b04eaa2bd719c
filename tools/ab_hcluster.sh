for o in 0 1; do
IPOKE_INN_HCLUSTER=$o timeout 100 python tools/small_bench.py --config cfgB --batch 256 --reps 200 --tag "hc=$o" 2>&1 | tail -1
IPOKE_INN_HCLUSTER=$o IPOKE_INN_STREAMS=2 timeout 100 python tools/small_bench.py --config cfgB --batch 256 --reps 200 --dual --tag "hc=$o dual" 2>&1 | tail -1
IPOKE_INN_HCLUSTER=$o timeout 100 python tools/small_bench.py --config cfgB --batch 256 --precision fp32 --reps 100 --tag "hc=$o fp32" 2>&1 | tail -1
IPOKE_INN_HCLUSTER=$o timeout 100 python tools/small_bench.py --config cfgA --batch 256 --reps 100 --tag "hc=$o" 2>&1 | tail -1
IPOKE_INN_HCLUSTER=$o timeout 100 python tools/small_bench.py --config cfgA --batch 64 --reps 100 --tag "hc=$o" 2>&1 | tail -1
IPOKE_INN_HCLUSTER=$o timeout 100 python tools/small_bench.py --config cfgB --batch 1024 --reps 100 --tag "hc=$o" 2>&1 | tail -1
done
