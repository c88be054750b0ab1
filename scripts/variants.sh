#!/bin/bash
# A/B of the compress path variants on the small / large extras: default, separate k_coalesce, unfused launches
for v in "" "DGB_FUSED_COALESCE=0" "DGB_FUSED=0"; do
  echo "== variant: ${v:-default}"
  env $v python scripts/bench_configs.py --steps 30 ${1:---only ans,bf16_small,bf16_tiny,ans_big} 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    r = d.get('reference', {})
    print('%-34s ours c %.4f ms d %.4f ms | ref c %.4f d %.4f' % (d['config'], d['ours']['compress_ms'], d['ours']['decompress_ms'], r.get('compress_ms', 0), r.get('decompress_ms', 0)))
"
done
