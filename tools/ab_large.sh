for cfg in cfgB cfgA; do
for opt in "X=0" "IPOKE_INN_FOLD=0" "IPOKE_INN_TAIL_BN=64" "IPOKE_INN_TAIL_BN=256" "IPOKE_INN_DIRECT=0" "IPOKE_INN_DIRECT=2" "IPOKE_INN_SLAB4=1" "IPOKE_INN_HIDDEN_BN=128" "IPOKE_INN_MCAST=2" "IPOKE_INN_PDL=0"; do
env $opt timeout 100 python tools/small_bench.py --config $cfg --batch 32768 --reps 5 --tag "$opt" 2>&1 | tail -1
done; done
