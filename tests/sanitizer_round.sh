#!/bin/bash
# compute-sanitizer over the product's kernels (GPU box): bash tests/sanitizer_round.sh <tag>
# One log per tool under gpurun_out/<tag>/; the last lines of each hold the tool's error summary.
tag=${1:-sanitizer}; out=gpurun_out/$tag; mkdir -p $out
timeout 300 python tests/sanitizer_workload.py > $out/plain.log 2>&1; echo "plain exit $?"; tail -2 $out/plain.log
for tool in memcheck synccheck racecheck initcheck; do
  timeout ${SAN_TIMEOUT:-1500} compute-sanitizer --tool $tool --print-limit 40 python tests/sanitizer_workload.py > $out/$tool.log 2>&1
  echo "$tool exit $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SANITIZER_WORKLOAD_OK|^done" $out/$tool.log | tail -8
done
