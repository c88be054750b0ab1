mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r2e_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -4 gpurun_out/r2e_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r2e_pytest.log 2>&1; echo "pytest all rc=$?"; tail -4 gpurun_out/r2e_pytest.log
timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>&1 | tail -1 | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print('bench', d['value'], d['ms_per_step'], d['phases']['kernels_ms_per_step'])"
