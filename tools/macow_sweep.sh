#!/bin/bash
# One GPU call for the MaCow slice and the no-alloc entry: parity tests (full tracebacks kept), then the bench.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/noalloc_alone.log 2>&1; echo "noalloc alone exit $?"
tail -3 gpurun_out/noalloc_alone.log
timeout 400 python -m pytest tests/test_gpu_macow.py tests/test_zz_no_alloc.py -m gpu -q --tb=short > gpurun_out/macow_gputest.log 2>&1; echo "pytest exit $?"
grep -E "^E  |^FAILED|passed|failed|Error" gpurun_out/macow_gputest.log | head -40
timeout 300 python tools/macow_bench.py > gpurun_out/macow_bench.jsonl 2> gpurun_out/macow_bench.err; echo "bench exit $?"
python - <<'P'
import json
for l in open("gpurun_out/macow_bench.jsonl"):
    d = json.loads(l)
    print(d["workload"], "fwd", d["forward"]["ms"], d["forward"]["roofline"]["frac"], "rev", d["reverse"]["ms"], d["reverse"]["roofline"]["frac"],
          "eager", d.get("eager_gpu", {}).get("forward_ms"), d.get("eager_gpu", {}).get("reverse_ms"), d.get("cpu"))
P
