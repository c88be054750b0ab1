#!/bin/bash
# experiment: element-group pipelining with occupancy-limited kernels so split and encode co-reside
run() { echo "groups=$1 enc=$2 split=$3"; DGB_OVERLAP_GROUPS=$1 DGB_ENC_CTAS=$2 DGB_SPLIT_CTAS=$3 python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['phases']['compress_GBps_uncompressed'], d['phases']['kernels_ms_per_step'])"; }
run 1 10 6
run 1 7 6
run 1 10 2
run 1 10 3
run 4 7 2
run 8 7 2
run 8 6 3
run 8 10 6
run 16 7 2
