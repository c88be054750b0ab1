#!/bin/bash
# Shrinking split claims at the end of the fused compress kernel (straggler control).
run() { echo "== $*"; env "$@" python scripts/bench_configs.py --only ${ONLY:-bf16,f32_64,f16} --steps 20 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'][:30], 'compress ms', d['ours']['compress_ms'], 'GB/s', d['ours']['compress_GBps'])
"; }
run DGB_SPLIT_TAIL_ROUNDS=0
run DGB_SPLIT_TAIL_ROUNDS=1 DGB_SPLIT_TAIL_CLAIM=2
run DGB_SPLIT_TAIL_ROUNDS=2 DGB_SPLIT_TAIL_CLAIM=2
run DGB_SPLIT_TAIL_ROUNDS=4 DGB_SPLIT_TAIL_CLAIM=2
run DGB_SPLIT_TAIL_ROUNDS=2 DGB_SPLIT_TAIL_CLAIM=4
run DGB_SPLIT_TAIL_ROUNDS=2 DGB_SPLIT_TAIL_CLAIM=8
