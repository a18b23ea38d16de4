#!/bin/bash
# compute-sanitizer over one small call of every kernel family.  usage (GPU box): scripts/sanitize.sh <tag>
TAG=${1:-r1}
O=gpurun_out
mkdir -p $O
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python scripts/sanitize_case.py > $O/${TAG}_sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|^ok" $O/${TAG}_sanitizer_$tool.log
done
