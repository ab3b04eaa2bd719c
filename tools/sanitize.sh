#!/bin/bash
# compute-sanitizer passes over the hand-written kernels (tools/sanitize_run.py); logs under gpurun_out/ (copy the
# summaries to profiles/).  Usage: tools/sanitize.sh [tool ...]   tools: memcheck racecheck synccheck initcheck
OUT=${OUT:-gpurun_out}
mkdir -p $OUT
TOOLS=${@:-memcheck racecheck synccheck initcheck}
for t in $TOOLS; do
  echo "== compute-sanitizer --tool $t"
  timeout 1500 compute-sanitizer --tool $t --print-limit 40 --log-file $OUT/r02_sanitizer_$t.log \
      python tools/sanitize_run.py > $OUT/r02_sanitizer_$t.out 2>&1
  echo "exit $?"; tail -2 $OUT/r02_sanitizer_$t.out; grep -E "ERROR SUMMARY|RACECHECK SUMMARY" $OUT/r02_sanitizer_$t.log | tail -2
done
