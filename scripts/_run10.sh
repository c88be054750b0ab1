timeout 600 python scripts/bench_configs.py --steps 30 --only ans --custom bf16:1:65536,bf16:8:16384,bf16:16:8192,bf16:32:2048 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    o, r, s = d['ours'], d.get('reference', {}), d.get('speedup', {})
    print(f\"{d['config'][:36]:36s} ours c {o['compress_ms']:.4f} d {o['decompress_ms']:.4f} ms | ref c {r.get('compress_ms', 0):.4f} d {r.get('decompress_ms', 0):.4f} | x {s.get('compress')} {s.get('decompress')} | {d.get('kernels_ms_per_call')}\")
"
