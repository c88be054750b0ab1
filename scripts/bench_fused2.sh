#!/bin/bash
run() { echo "fused=$1 split_ctas=$2"; DGB_FUSED=$1 DGB_FUSED_SPLIT_CTAS=$2 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases']['compress_GBps_uncompressed'], d['phases']['kernels_ms_per_step'])"; }
for k in 5 6 7 8 9; do run 1 $k; done
