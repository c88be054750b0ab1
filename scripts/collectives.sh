#!/bin/bash
# Compressed collectives (BASELINE config 5) on N GPUs of one box: NCCL-moved archives and the peer-memory variant.
N=${1:-2}; MIB=${2:-64}
export NCCL_DEBUG=WARN
for c in 1 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$c scripts/bench_collectives.py --mib $MIB --steps 10 --chunks $c 2>&1 | grep -v Warning | tail -6
done
