cd /root/repo
timeout 600 python scripts/stress_fused.py 400 2>&1 | tail -6
python -m pytest tests/test_gpu_stress.py tests/test_gpu_large.py -x -q -m gpu 2>&1 | tail -3
bash scripts/knob_sweep.sh DGB_FUSED "1 1" bench --no-collective
python scripts/bench_configs.py --steps 20 --only f32_1g,f32_64 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    print('%-40s c %.4f ms  d %.4f ms' % (d['config'][:40], d['ours']['compress_ms'], d['ours']['decompress_ms']))"
