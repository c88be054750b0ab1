timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for c in 1 2 3 4; do echo "claim=$c"; DGB_DEC_CLAIM=$c python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['phases']['decompress_GBps_uncompressed'], d['phases']['kernels_ms_per_step'])"; done
