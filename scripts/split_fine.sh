#!/bin/bash
# Fine-prefix experiment for the fused compress kernel: the first T tiles are claimed in 4096-word slices.
run() { echo "== $*"; env "$@" python scripts/bench_configs.py --only ${ONLY:-bf16,f32_64,f16} --steps 20 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'][:30], 'compress ms', d['ours']['compress_ms'], 'GB/s', d['ours']['compress_GBps'])
"; }
for t in 0 128 256 512 1024 2048; do run DGB_SPLIT_FINE_TILES=$t; done
ONLY=f32_1g,ans_big run DGB_X=1
