"""`-m gpu` test of the peer-memory collectives' code path on ONE GPU (a group of one rank): peer buffer allocation
and IPC export, compress into the raw buffer, the stream-ordered barrier, decode from the (own) buffer, alternating
halves. The multi-GPU runs (reads over NVLink) are scripts/bench_collectives.py, which asserts bit equality with
the uncompressed NCCL collectives before it times anything."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def group_of_one():
    import torch.distributed as dist
    if dist.is_initialized():
        pytest.skip("a process group already exists in this process")
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:29577", world_size=1, rank=0,
                            device_id=torch.device("cuda", 0))
    yield dist
    dist.destroy_process_group()


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16, torch.float32])
def test_peer_all_gather_and_all_to_all_world1(group_of_one, dtype):
    from dietgpu_fork_b200 import dist as dgd
    n = 300001 if dtype != torch.float32 else 150001
    peers = dgd.PeerBuffers(dgd.peer_bytes_needed(dtype, n, 3))
    try:
        for it in range(3):  # both halves, and the first one again
            t = (torch.randn(n, device="cuda") * 0.02).to(dtype)
            out = dgd.all_gather_compressed_peer(t, peers, num_chunks=3, mode="inplace" if it == 1 else "pull")
            torch.cuda.synchronize()
            assert out.shape == (1, n)
            assert torch.equal(out[0].view(torch.uint8), t.view(torch.uint8)), f"all-gather round {it}"
            t2 = (torch.randn(1, n // 2, device="cuda") * 0.02).to(dtype)
            out2 = dgd.all_to_all_compressed_peer(t2, peers)
            torch.cuda.synchronize()
            assert torch.equal(out2.view(torch.uint8), t2.view(torch.uint8)), f"all-to-all round {it}"
    finally:
        peers.close()


def test_peer_buffer_too_small_is_refused(group_of_one):
    from dietgpu_fork_b200 import dist as dgd
    peers = dgd.PeerBuffers(4096)
    try:
        t = torch.zeros(100000, dtype=torch.bfloat16, device="cuda")
        with pytest.raises(AssertionError):
            dgd.all_gather_compressed_peer(t, peers)
    finally:
        peers.close()
