#!/bin/bash
# A/B sweep over sets of experiment knobs: every line of <file> is a list of VAR=value assignments (DGB_* knobs, read
# once per process by csrc/knobs.h). Per line: bench.py's device-timed kernels, and (with "ncu" as 2nd argument) the DRAM
# bytes / issue activity of one launch of the kernel matching the 3rd argument.
# Usage: bash scripts/env_sweep.sh <file> [ncu] [kernel-regex]
file=$1; withncu=${2:-}; kre=${3:-k_compressFused}
while IFS= read -r line || [ -n "$line" ]; do
  case "$line" in \#*) continue;; esac
  echo "== ${line:-<defaults>}"
  env $line timeout 150 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.read())
    print('   value', d['value'], 'ms/step', d['ms_per_step'], d['phases']['kernels_ms_per_step'])
except Exception as e:
    print('   bench failed', e)"
  if [ "$withncu" = ncu ]; then
    env $line timeout 240 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum,lts__t_sector_hit_rate.pct \
      --clock-control none -k regex:$kre --launch-skip 3 -c 1 --csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null \
      | python -c "
import sys, csv
rows = [r for r in csv.reader(sys.stdin) if len(r) > 10]
print('   ncu', {r[-3]: r[-1] + ' ' + r[-2] for r in rows[1:]})"
  fi
done < "$file"
