timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
timeout 900 python scripts/bench_configs.py --steps 10 2>&1 | tee gpurun_out/r1_configs.jsonl | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: print(l.strip()[:200]); continue
    o, r, s = d['ours'], d.get('reference', {}), d.get('speedup', {})
    print(f\"{d['config'][:36]:36s} ratio {d['ratio']:.3f} | ours c {o['compress_GBps']:7.1f} d {o['decompress_GBps']:7.1f} | ref c {r.get('compress_GBps', 0):7.1f} d {r.get('decompress_GBps', 0):7.1f} | x {s.get('compress')} {s.get('decompress')} {s.get('round_trip')}\")
"
