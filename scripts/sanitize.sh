#!/bin/bash
# compute-sanitizer over scripts/sanitize_driver.py, one log per tool under gpurun_out/ (copied to profiles/ by hand).
# Usage (through gpurun): bash scripts/sanitize.sh [tag]
tag=${1:-r2}
mkdir -p gpurun_out
for tool in memcheck racecheck synccheck initcheck; do
  timeout 1200 compute-sanitizer --tool $tool --print-limit 20 --log-file gpurun_out/${tag}_sanitizer_$tool.log \
    python scripts/sanitize_driver.py > gpurun_out/${tag}_sanitizer_${tool}_stdout.log 2>&1
  echo "$tool rc=$? $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' gpurun_out/${tag}_sanitizer_$tool.log | tail -1) $(tail -1 gpurun_out/${tag}_sanitizer_${tool}_stdout.log)"
done
