#!/bin/bash
# One A/B sweep of an experiment knob (DGB_* environment variable, read once per process by csrc/knobs.h) over the
# bench workload or the per-config bench. Replaces the one-off sweep scripts of round 1 (results: DESIGN.md section 8).
# Usage: bash scripts/knob_sweep.sh <KNOB> "<v1> <v2> ..." [bench|configs] [extra args of the chosen driver]
knob=$1; vals=$2; what=${3:-bench}; shift 3 2>/dev/null
for v in $vals; do
  echo "== $knob=$v"
  if [ "$what" = configs ]; then
    env $knob=$v timeout 600 python scripts/bench_configs.py --steps 20 "$@" 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('%-40s c %.4f ms  d %.4f ms' % (d['config'][:40], d['ours']['compress_ms'], d['ours']['decompress_ms']))"
  else
    env $knob=$v timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e "$@" 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['phases']['kernels_ms_per_step'])"
  fi
done
