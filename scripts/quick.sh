#!/bin/bash
# quick GPU check: parity tests then two short bench runs (no CPU baseline)
python -m pytest tests -x -q -m gpu 2>&1 | tail -3
bash scripts/bench_twice.sh
