timeout 600 python -m pytest tests/test_gpu_parity.py -x -q 2>&1 | tail -2
DGB_SPLIT_PIPE=1 python scripts/prof_small.py 2>&1 | grep -v wall
DGB_SPLIT_PIPE=0 python scripts/prof_small.py 2>&1 | grep "f32\|f64"
