for o in 0 1; do IPOKE_INN_WGRAD_WIDE=$o timeout 120 python tools/train_bench.py --config cfgA --batch 4096 2>&1 | grep -v Warn | tail -3; done
IPOKE_INN_WGRAD_WIDE=1 timeout 120 python tools/train_bench.py --config cfgB --batch 4096 2>&1 | grep -v Warn | tail -3
