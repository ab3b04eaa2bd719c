#!/bin/bash
# One GPU call for the MaCow slice: parity tests (tracebacks kept), the bench with baselines, launch-geometry variants.
mkdir -p gpurun_out; rm -f gpurun_out/macow_sweep.jsonl
timeout 400 python -m pytest tests/test_gpu_macow.py -m gpu -q --tb=short > gpurun_out/macow_gputest.log 2>&1; echo "pytest exit $?"
grep -E "^E  |^FAILED|passed|failed|Error" gpurun_out/macow_gputest.log | head -30
timeout 300 python tools/macow_bench.py > gpurun_out/macow_bench.jsonl 2> gpurun_out/macow_bench.err; echo "bench exit $?"
for v in "IPOKE_MACOW_KSPLIT=4" "IPOKE_MACOW_THREADS=256"; do
  echo "## $v" >> gpurun_out/macow_sweep.jsonl
  env $v timeout 120 python tools/macow_bench.py --no-baselines --steps 10 --warmup 3 >> gpurun_out/macow_sweep.jsonl 2>> gpurun_out/macow_bench.err
done
python - <<'P'
import json, os
for f in ("gpurun_out/macow_bench.jsonl", "gpurun_out/macow_sweep.jsonl"):
    if not os.path.exists(f): continue
    for l in open(f):
        if l.startswith("#"): print(l.strip()); continue
        d = json.loads(l)
        print(d["workload"], "fwd", d["forward"]["ms"], d["forward"]["roofline"]["frac"], "rev", d["reverse"]["ms"], d["reverse"]["roofline"]["frac"],
              "eager", d.get("eager_gpu", {}).get("forward_ms"), d.get("eager_gpu", {}).get("reverse_ms"))
P
