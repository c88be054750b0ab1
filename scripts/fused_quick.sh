timeout 600 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -1
for k in ${@:-3}; do echo "split_ctas=$k"; DGB_FUSED_SPLIT_CTAS=$k python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>&1 | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['phases']['compress_GBps_uncompressed'], d['phases']['kernels_ms_per_step'])"; done
DGB_FUSED_TRACE=/tmp/trace.txt python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python scripts/trace_fused.py /tmp/trace.txt
