"""Worker of tests/test_multigpu.py (launched by torch.distributed.run, one rank per GPU): consecutive compressed
collectives with different data per call and per rank against the uncompressed NCCL collectives, bit for bit, in
both forms (archives moved by NCCL / peer memory read in place); then BASELINE configs[2](ii), a 64-element fp32 batch
cut across the ranks by dist.compress_sharded and checked against a single-GPU compress of the same elements."""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import codec, dist as dgd  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
    temp = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    cdc = dgd.CudaFloatCodec(temp_mem=temp)
    n = 3 * 1024 * 1024 + 77 * 8  # words per rank, a multiple of 8
    per = n // world // 8 * 8
    peers = dgd.PeerBuffers(max(dgd.peer_bytes_needed(torch.bfloat16, n, 2),
                                dgd.peer_bytes_needed(torch.bfloat16, per, 1) * world))
    g = torch.Generator(device="cuda")
    # at least 3 consecutive calls of each peer collective: both buffer halves are reused while slower ranks may
    # still be decoding the previous call (the one-barrier-per-call argument of dist.PeerBuffers)
    for it in range(5):
        g.manual_seed(1000 * it + rank)
        shard = (torch.randn(n, device="cuda", generator=g) * (0.02 + 0.01 * rank)).to(torch.bfloat16)
        raw = torch.empty((world, n), dtype=torch.bfloat16, device="cuda")
        dist.all_gather_into_tensor(raw.view(-1), shard)
        mode = "inplace" if it == 3 else "pull"  # both forms, interleaved on the same buffers
        out = dgd.all_gather_compressed_peer(shard, peers, num_chunks=1 + it % 2, temp=temp, mode=mode,
                                             stages=2 if it == 4 else None)
        assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), f"peer all-gather ({mode}), call {it}"
        out = dgd.all_gather_compressed(shard, num_chunks=2, codec=cdc)
        assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), f"NCCL-form all-gather, call {it}"
        send = shard[:per * world].view(world, per)
        raw2 = torch.empty_like(send)
        dist.all_to_all_single(raw2.view(-1), send.reshape(-1))
        out2 = dgd.all_to_all_compressed_peer(send, peers, temp=temp, mode=mode)
        assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), f"peer all-to-all ({mode}), call {it}"
        out2 = dgd.all_to_all_compressed(send, codec=cdc)
        assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), f"NCCL-form all-to-all, call {it}"
        if it == 2 and rank == world - 1:
            torch.cuda._sleep(200_000_000)  # one rank falls behind: the others run ahead into the next call
    torch.cuda.synchronize()
    peers.close()

    # configs[2](ii): every rank holds the description of the same 64 x 1 Mi-word fp32 batch, compresses its slice
    g.manual_seed(77)
    batch = [torch.randn(1 << 20, device="cuda", generator=g) for _ in range(64)]
    lo, hi, comp, sizes = dgd.compress_sharded(batch)
    assert (lo, hi) == dgd.partition(64, world, rank)
    mx = codec.get_max_float_compressed_size(codec.FLOAT32, 1 << 20)
    ref = torch.zeros((hi - lo, mx), dtype=torch.uint8, device="cuda")
    rsz = torch.zeros(hi - lo, dtype=torch.int32, device="cuda")
    codec.float_compress(codec.FLOAT32, batch[lo:hi], [ref[i] for i in range(hi - lo)], rsz)
    torch.cuda.synchronize()
    assert torch.equal(sizes[:hi - lo].cpu(), rsz.cpu())
    for i in range(hi - lo):
        s = int(rsz[i])
        assert torch.equal(comp[i, :s], ref[i, :s]), f"sharded archive {lo + i} differs from the single-GPU one"
    outs = [torch.empty_like(batch[0]) for _ in range(hi - lo)]
    cdc.decompress([comp[i, :int(sizes[i])] for i in range(hi - lo)], outs)
    torch.cuda.synchronize()
    for i in range(hi - lo):
        assert torch.equal(outs[i].view(torch.int32), batch[lo + i].view(torch.int32))
    dist.barrier()
    if rank == 0:
        print(f"MULTIGPU-OK world={world}", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
