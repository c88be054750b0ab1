#!/bin/bash
# ncu --set full capture of one kernel of the bench (after the same command ran clean without ncu).
# Usage: bash scripts/ncu_kernel.sh <kernel-name (exact base name, or regex:...)> <out-tag> [launch-skip]
k=$1; tag=$2; skip=${3:-3}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k $k --launch-skip $skip -c 1 -f -o gpurun_out/$tag \
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_ncu.log 2>&1
echo "ncu rc=$?"; ls -la gpurun_out/$tag.ncu-rep
