"""world_size-2 `gloo` tests (CPU) of the multi-GPU host logic in dietgpu_fork_b200/dist.py: element
partitioning, the sizes exchange, row packing and the receive-side addressing of the compressed all-gather /
all-to-all. The CUDA codec cannot run here, so the tests inject a stand-in codec built on the CPU oracle (test
infrastructure only); the exchange code under test is exactly what runs over NCCL on the GPUs."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from dietgpu_fork_b200 import dist as dgd

_NP_VIEW = {torch.bfloat16: (torch.int16, np.uint16, oracle.BFLOAT16),
            torch.float16: (torch.int16, np.uint16, oracle.FLOAT16),
            torch.float32: (torch.int32, np.uint32, oracle.FLOAT32)}


class OracleCodec:
    """Same interface as dist.CudaFloatCodec, CPU tensors, archives made by the oracle."""

    def compress(self, ts, out=None, out_sizes=None):
        tv, nv, ft = _NP_VIEW[ts[0].dtype]
        arch = [oracle.float_compress(ft, t.contiguous().view(tv).numpy().view(nv)) for t in ts]
        width = max(oracle.float_max_compressed_size(ft, t.numel()) for t in ts)
        comp = torch.zeros((len(ts), width), dtype=torch.uint8)
        for i, a in enumerate(arch):
            comp[i, :a.size] = torch.from_numpy(a)
        return comp, torch.tensor([a.size for a in arch], dtype=torch.int32)

    def decompress(self, rows, outs):
        for r, o in zip(rows, outs):
            tv, nv, _ = _NP_VIEW[o.dtype]
            ok, _, words = oracle.float_decompress(r.numpy())
            assert ok
            o.view(tv).copy_(torch.from_numpy(np.ascontiguousarray(words).view(
                {np.uint16: np.int16, np.uint32: np.int32}[nv])))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, fn, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ret[rank] = fn(rank, world)
    finally:
        dist.destroy_process_group()


def _run(fn, world=2):
    ret = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), fn, ret), nprocs=world, join=True)
    return [ret[r] for r in range(world)]


def _shard(rank, n, dtype, seed=5000):
    g = torch.Generator().manual_seed(seed + rank)
    return (torch.randn(n, generator=g) * 0.02).to(dtype)


def _all_gather_case(rank, world):
    ok = True
    for dtype, n, chunks in [(torch.bfloat16, 50001, 4), (torch.float32, 4096, 1), (torch.float16, 7, 3)]:
        mine = _shard(rank, n, dtype)
        out = dgd.all_gather_compressed(mine, num_chunks=chunks, codec=OracleCodec())
        for r in range(world):
            ok &= torch.equal(out[r].view(torch.uint8), _shard(r, n, dtype).view(torch.uint8))
    return bool(ok)


def _all_to_all_case(rank, world):
    ok = True
    for dtype, n in [(torch.bfloat16, 30011), (torch.float32, 513)]:
        send = torch.stack([_shard(rank * 16 + j, n, dtype) for j in range(world)])
        out = dgd.all_to_all_compressed(send, codec=OracleCodec())
        for i in range(world):
            ok &= torch.equal(out[i].view(torch.uint8), _shard(i * 16 + rank, n, dtype).view(torch.uint8))
    return bool(ok)


def _sharded_case(rank, world):
    batch = [_shard(100 + i, 1000 + 37 * i, torch.bfloat16) for i in range(5)]
    lo, hi, comp, sizes = dgd.compress_sharded(batch, codec=OracleCodec())
    # gather every rank's (lo, hi, archives) and check that together they are the single-process result
    mine = [(lo + k, bytes(comp[k, :int(sizes[k])].numpy())) for k in range(hi - lo)]
    gathered = [None] * world
    dist.all_gather_object(gathered, mine)
    flat = sorted(x for part in gathered for x in part)
    whole, wsz = OracleCodec().compress(batch)
    return [i for i, _ in flat] == list(range(5)) and all(
        b == bytes(whole[i, :int(wsz[i])].numpy()) for i, b in flat)


def test_partition_covers_every_element_once():
    for num in (0, 1, 5, 8, 128, 1001):
        for world in (1, 2, 3, 8):
            got = []
            for r in range(world):
                lo, hi = dgd.partition(num, world, r)
                assert 0 <= hi - lo <= -(-num // world)
                got += list(range(lo, hi))
            assert got == list(range(num))


def test_all_gather_compressed_world2():
    assert _run(_all_gather_case) == [True, True]


def test_all_to_all_compressed_world2():
    assert _run(_all_to_all_case) == [True, True]


def test_sharded_batch_equals_single_process_world2():
    assert _run(_sharded_case) == [True, True]
