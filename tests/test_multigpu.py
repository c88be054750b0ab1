"""`-m gpu` multi-GPU test (skipped with fewer than 2 GPUs): tests/multigpu_worker.py under torch.distributed.run
with one rank per GPU - compressed all-gather / all-to-all in both forms over real NVLink reads, five consecutive
calls with different data per call and rank (buffer-half reuse under skew), and the sharded 64-element batch of
BASELINE configs[2](ii)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        return torch.cuda.device_count()
    except Exception:
        return 0


@pytest.mark.skipif(_gpus() < 2, reason="needs at least 2 GPUs on one node")
def test_compressed_collectives_and_sharded_batch_multi_gpu():
    world = min(_gpus(), 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29631", os.path.join(ROOT, "tests", "multigpu_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and f"MULTIGPU-OK world={world}" in r.stdout, (r.stdout[-2000:], r.stderr[-4000:])
