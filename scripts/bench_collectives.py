"""BASELINE config 5: compressed all-gather / all-to-all of bf16 gradient shards vs the uncompressed NCCL
collectives on the same box. Launch: python -m torch.distributed.run --nnodes=1 --nproc-per-node N
--master-addr 127.0.0.1 --master-port P scripts/bench_collectives.py [--mib 64] [--steps 10] [--chunks 4].
Effective GB/s = uncompressed bytes delivered per rank / time (CUDA events, max over ranks). Checks bit
equality of the compressed result against the uncompressed collective's before timing."""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import dist as dgd  # noqa: E402


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return ms.item()


def measure(dist_mod, rank, world, mib=64, steps=10, warmup=3, chunks=1, peer=True):
    """Runs inside an initialised NCCL process group (one rank per GPU, device already selected). Returns a dict:
    raw NCCL, archives moved by NCCL, and the peer-memory form, for all-gather and all-to-all; every compressed
    result is compared bit for bit with the uncompressed collective's before and after timing."""
    n = mib * (1 << 20) // 2
    g = torch.Generator(device="cuda").manual_seed(5000 + rank)
    shard = (torch.randn(n, device="cuda", generator=g) * 0.02).to(torch.bfloat16)
    temp = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    codec = dgd.CudaFloatCodec(temp_mem=temp)
    res = {"workload": f"bf16 N(0, 0.02), {mib} MiB per rank, {world} GPUs", "n_gpus": world, "shard_mib": mib,
           "chunks": chunks, "unit": "GB/s of uncompressed bytes delivered per rank", "bit_exact": True}

    # --- all-gather ---
    raw = torch.empty((world, n), dtype=torch.bfloat16, device="cuda")
    out = torch.empty_like(raw)
    dist.all_gather_into_tensor(raw.view(-1), shard)
    dgd.all_gather_compressed(shard, out=out, num_chunks=chunks, codec=codec)
    assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), "compressed all-gather is not bit exact"
    t_raw = timed(lambda: dist.all_gather_into_tensor(raw.view(-1), shard), steps, warmup)
    t_cmp = timed(lambda: dgd.all_gather_compressed(shard, out=out, num_chunks=chunks, codec=codec), steps, warmup)
    delivered = world * n * 2
    ag = {"nccl_raw_ms": round(t_raw, 4), "nccl_raw": round(delivered / t_raw / 1e6, 1),
          "compressed_nccl_ms": round(t_cmp, 4), "compressed_nccl": round(delivered / t_cmp / 1e6, 1)}
    peers = None
    if peer:
        peers = dgd.PeerBuffers(max(dgd.peer_bytes_needed(torch.bfloat16, n, chunks),
                                    dgd.peer_bytes_needed(torch.bfloat16, n // world, 1) * world))
        out.zero_()
        dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=chunks, temp=temp)
        torch.cuda.synchronize()
        assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), "peer all-gather is not bit exact"
        t_peer = timed(lambda: dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=chunks, temp=temp),
                       steps, warmup)
        assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), "peer all-gather is not bit exact (timed)"
        ag.update({"compressed_peer_ms": round(t_peer, 4), "compressed_peer": round(delivered / t_peer / 1e6, 1),
                   "peer_vs_raw": round(t_raw / t_peer, 3), "peer_mode": "decoder reads the peers' archives in place"})
        out.zero_()
        dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=chunks, temp=temp, mode="pull")
        torch.cuda.synchronize()
        assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), "pull-mode peer all-gather is not bit exact"
        t_inp = timed(lambda: dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=chunks, temp=temp,
                                                             mode="pull"), steps, warmup)
        ag.update({"compressed_peer_pull_ms": round(t_inp, 4),
                   "compressed_peer_pull": round(delivered / t_inp / 1e6, 1)})
        if world > 3:  # fewer, larger pull / decode steps
            t_s2 = timed(lambda: dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=chunks, temp=temp,
                                                                mode="pull", stages=2), steps, warmup)
            assert torch.equal(raw.view(torch.int16), out.view(torch.int16)), "peer all-gather (2 stages) is not bit exact"
            ag.update({"compressed_peer_pull_2stages_ms": round(t_s2, 4)})
    res["all_gather"] = ag

    # --- all-to-all: mib/world MiB to every peer ---
    per = n // world
    send = shard[:per * world].view(world, per)
    raw2 = torch.empty_like(send)
    out2 = torch.empty_like(send)
    dist.all_to_all_single(raw2.view(-1), send.reshape(-1))
    dgd.all_to_all_compressed(send, out=out2, codec=codec)
    assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), "compressed all-to-all is not bit exact"
    t_raw = timed(lambda: dist.all_to_all_single(raw2.view(-1), send.reshape(-1)), steps, warmup)
    t_cmp = timed(lambda: dgd.all_to_all_compressed(send, out=out2, codec=codec), steps, warmup)
    delivered = world * per * 2
    aa = {"chunk_mib": per * 2 / (1 << 20), "nccl_raw_ms": round(t_raw, 4),
          "nccl_raw": round(delivered / t_raw / 1e6, 1), "compressed_nccl_ms": round(t_cmp, 4),
          "compressed_nccl": round(delivered / t_cmp / 1e6, 1)}
    if peers is not None:
        out2.zero_()
        dgd.all_to_all_compressed_peer(send, peers, out=out2, temp=temp)
        torch.cuda.synchronize()
        assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), "peer all-to-all is not bit exact"
        t_peer = timed(lambda: dgd.all_to_all_compressed_peer(send, peers, out=out2, temp=temp), steps, warmup)
        assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), "peer all-to-all is not bit exact (timed)"
        aa.update({"compressed_peer_ms": round(t_peer, 4), "compressed_peer": round(delivered / t_peer / 1e6, 1),
                   "peer_vs_raw": round(t_raw / t_peer, 3)})
        out2.zero_()
        dgd.all_to_all_compressed_peer(send, peers, out=out2, temp=temp, mode="pull")
        torch.cuda.synchronize()
        assert torch.equal(raw2.view(torch.int16), out2.view(torch.int16)), "pull-mode peer all-to-all is not bit exact"
        t_inp = timed(lambda: dgd.all_to_all_compressed_peer(send, peers, out=out2, temp=temp, mode="pull"), steps,
                      warmup)
        aa.update({"compressed_peer_pull_ms": round(t_inp, 4),
                   "compressed_peer_pull": round(delivered / t_inp / 1e6, 1)})
        peers.close()
    res["all_to_all"] = aa
    return res


def measure_sharded(dist_mod, rank, world, batch=64, words=4 << 20, steps=10, warmup=3):
    """BASELINE configs[2](ii): one fp32 batch of 64 x 16 MiB whose elements are partitioned over the ranks by index
    (dist.partition / compress_sharded; no data-path collective). Every rank compresses and decompresses its slice;
    the job's throughput is the whole batch's uncompressed bytes over the slowest rank's time (strong scaling: the
    total work is fixed). The round trip of every element is checked."""
    from dietgpu_fork_b200 import codec
    lo, hi = dgd.partition(batch, world, rank)
    g = torch.Generator(device="cuda").manual_seed(3000)
    ins = []
    for i in range(batch):  # the same batch on every rank; only the own slice is kept
        t = torch.randn(words, device="cuda", generator=g)
        if lo <= i < hi:
            ins.append(t)
    n = hi - lo
    mx = (codec.get_max_float_compressed_size(codec.FLOAT32, words) + 15) // 16 * 16
    comp = torch.empty((max(n, 1), mx), dtype=torch.uint8, device="cuda")
    rows = [comp[i] for i in range(n)]
    sizes = torch.zeros(max(n, 1), dtype=torch.int32, device="cuda")
    outs = [torch.empty_like(t) for t in ins]
    ok = torch.zeros(max(n, 1), dtype=torch.uint8, device="cuda")
    temp = torch.empty(int(2.5 * n * words * 4) + (64 << 20), dtype=torch.uint8, device="cuda")

    def comp_fn():
        if n:
            codec.float_compress(codec.FLOAT32, ins, rows, sizes, 10, False, temp=temp)

    def dec_fn():
        if n:
            codec.float_decompress(codec.FLOAT32, rows, outs, ok, None, 10, False, temp=temp)

    t_c = timed(comp_fn, steps, warmup)
    t_d = timed(dec_fn, steps, warmup)
    for a, b in zip(ins, outs):
        assert torch.equal(a.view(torch.int32), b.view(torch.int32)), "sharded round trip mismatch"
    U = batch * words * 4
    return {"workload": f"fp32 N(0, 1), one batch of {batch} x {words * 4 >> 20} MiB partitioned over {world} GPUs by element",
            "scaling": "strong", "elements_per_rank": n, "compress_ms": round(t_c, 4), "decompress_ms": round(t_d, 4),
            "compress_GBps": round(U / t_c / 1e6, 1), "decompress_GBps": round(U / t_d / 1e6, 1),
            "round_trip_GBps": round(2 * U / (t_c + t_d) / 1e6, 1)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mib", type=int, default=64)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--chunks", type=int, default=1)
    ap.add_argument("--peer", type=int, default=1, help="also time the peer-memory collectives (CUDA IPC)")
    a = ap.parse_args()
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    res = measure(dist, rank, world, a.mib, a.steps, a.warmup, a.chunks, bool(a.peer))
    res["sharded_fp32"] = measure_sharded(dist, rank, world, steps=a.steps, warmup=a.warmup)
    if rank == 0:
        print(json.dumps(res), flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
