DGB_FUSED_TRACE=/tmp/trace.txt python bench.py --steps 2 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
python scripts/trace_fused.py /tmp/trace.txt
