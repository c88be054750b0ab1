#!/bin/bash
# Split-claim granularity experiments for the fused compress kernel (few huge elements vs many small ones).
run() { echo "== $*"; env "$@" python scripts/bench_configs.py --only ${ONLY:-f32_1g} --steps 10 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'][:30], 'compress ms', d['ours']['compress_ms'], 'GB/s', d['ours']['compress_GBps'], 'ref', d.get('reference', {}).get('compress_GBps'))
"; }
run DGB_X=1
run DGB_SPLIT_SLICE_SHIFT=3 DGB_FUSED_SPLIT_CTAS=3
run DGB_SPLIT_SLICE_SHIFT=3 DGB_FUSED_SPLIT_CTAS=6
run DGB_SPLIT_SLICE_SHIFT=3 DGB_FUSED_SPLIT_CTAS=8
run DGB_SPLIT_SLICE_SHIFT=2 DGB_FUSED_SPLIT_CTAS=9
run DGB_SPLIT_SLICE_SHIFT=1 DGB_FUSED_SPLIT_CTAS=9
run DGB_SPLIT_SLICE_SHIFT=0 DGB_FUSED_SPLIT_CTAS=9
run DGB_SPLIT_SLICE_SHIFT=3 DGB_SPLIT_CLAIM=2 DGB_FUSED_SPLIT_CTAS=9
ONLY=bf16,f32_64,f16,ans run DGB_X=1
ONLY=bf16,f32_64 run DGB_SPLIT_SLICE_SHIFT=2 DGB_SPLIT_CLAIM=4
