timeout 600 python -m pytest tests -x -q -m gpu -k "sparse" 2>&1 | tail -2
timeout 300 python scripts/bench_configs.py --steps 30 --only sparse 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    o = d['ours']; print(d['config'], 'c', o['compress_GBps'], 'd', o['decompress_GBps'], d.get('kernels_ms_per_call'))
"
