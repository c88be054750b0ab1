cd /root/repo
python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-e2e --no-collective 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read())
print(d['value'], d['ms_per_step'], d['phases']['kernels_ms_per_step'])"
python scripts/bench_configs.py --steps 20 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('%-40s c %.4f ms  d %.4f ms' % (d['config'][:40], d['ours']['compress_ms'], d['ours']['decompress_ms']))"
