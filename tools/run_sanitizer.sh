#!/bin/bash
# compute-sanitizer over small shapes of every kernel family (tools/sanitize_small.py); logs go to gpurun_out/ and the
# summaries are copied to profiles/ by hand. racecheck covers the shared-memory protocols (SASO ring + mbarriers, tcgen05 rings).
set -x
cd "$(dirname "$0")/.."
for tool in memcheck racecheck; do
  timeout 600 compute-sanitizer --tool $tool --log-file gpurun_out/r02_sanitizer_$tool.log python tools/sanitize_small.py > gpurun_out/r02_sanitizer_$tool.out 2>&1
  echo "$tool rc=$?"
  tail -n 5 gpurun_out/r02_sanitizer_$tool.log
done
