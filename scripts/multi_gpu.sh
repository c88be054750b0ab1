#!/bin/bash
# N-GPU checks: bench.py under torchrun (sharded batch, weak scaling), then the compressed collectives bench.
N=${1:-2}
export NCCL_DEBUG=WARN
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 10 --warmup 3 2>&1 | tail -2 | cut -c1-600
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 scripts/bench_collectives.py --mib 64 --steps 10 --chunks 4 2>&1 | tail -4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 scripts/bench_collectives.py --mib 64 --steps 10 --chunks 1 2>&1 | tail -4
