#!/bin/bash
# compute-sanitizer over tools/gpu_sanitize.py: one log per tool under gpurun_out/ (copy the summaries to profiles/)
O=gpurun_out; T=${1:-r02}
mkdir -p $O
for tool in memcheck racecheck initcheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file $O/${T}_sanitizer_$tool.txt python tools/gpu_sanitize.py > $O/${T}_sanitizer_$tool.out 2>&1
  echo "$tool rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $O/${T}_sanitizer_$tool.txt | tail -1)"
done
