for e in "A=1" "DGB_FUSED_PREFETCH_NEXT=1" "DGB_DEC_PDL=1" "DGB_FUSED_PREFETCH_NEXT=1 DGB_DEC_PDL=1" "A=1"; do echo "== $e: $(env $e python scripts/time_phases.py 40)"; done
DGB_FUSED_PREFETCH_NEXT=1 DGB_DEC_PDL=1 timeout 900 python -m pytest tests/test_gpu_baseline_configs.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -2
