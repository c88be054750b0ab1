#!/bin/bash
# One GPU box visit: parity tests, smoke, both bench arms, then the ncu launch list of the bench command.
# Usage (from the repo root, through gpurun): bash scripts/gpu_round.sh <tag> [skip-tests]
set -u
tag=${1:-run}
mkdir -p gpurun_out
if [ "${2:-}" != "skip-tests" ]; then
  timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/${tag}_pytest.log 2>&1; echo "pytest rc=$?" | tee gpurun_out/${tag}_rc.log
  tail -5 gpurun_out/${tag}_pytest.log
fi
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${tag}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a gpurun_out/${tag}_rc.log
tail -2 gpurun_out/${tag}_smoke.log
timeout 600 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/${tag}_bench_ref.json 2> gpurun_out/${tag}_bench_ref.err; echo "bench ref rc=$?" | tee -a gpurun_out/${tag}_rc.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/${tag}_bench.json 2> gpurun_out/${tag}_bench.err; rc=$?; echo "bench rc=$rc" | tee -a gpurun_out/${tag}_rc.log
tail -1 gpurun_out/${tag}_bench_ref.json | cut -c1-1800
tail -3 gpurun_out/${tag}_bench_ref.err
tail -1 gpurun_out/${tag}_bench.json
tail -3 gpurun_out/${tag}_bench.err
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/${tag}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_ncu.log 2>&1
  echo "ncu rc=$?" | tee -a gpurun_out/${tag}_rc.log
fi
timeout 600 python scripts/bench_configs.py --steps 20 --only none --custom bf16:1:65536,bf16:1:262144,bf16:8:16384,bf16:16:8192,bf16:32:2048 2>/dev/null | tee gpurun_out/${tag}_configs_medium.jsonl | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    o, r, s = d['ours'], d.get('reference', {}), d.get('speedup', {})
    print(f\"{d['config'][:36]:36s} ours c {o['compress_GBps']:7.1f} d {o['decompress_GBps']:7.1f} | ref c {r.get('compress_GBps', 0):7.1f} d {r.get('decompress_GBps', 0):7.1f} | x {s.get('compress')} {s.get('decompress')} {s.get('round_trip')}\")
"
