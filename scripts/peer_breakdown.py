"""Where the time of all_gather_compressed_peer goes (N GPUs, one box): compress alone, barrier alone, decode of
local archives only, decode of the peers' archives. torchrun launch like bench_collectives.py."""
import argparse, json, os, sys
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import codec, dist as dgd

def timed(fn, steps=10, warmup=3):
    for _ in range(warmup): fn()
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps): fn()
    b.record(); torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / steps], device="cuda")
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return round(ms.item(), 4)

ap = argparse.ArgumentParser(); ap.add_argument("--mib", type=int, default=64); ap.add_argument("--chunks", type=int, default=4)
a = ap.parse_args()
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
dist.init_process_group("nccl")
n = a.mib * (1 << 20) // 2
shard = (torch.randn(n, device="cuda") * 0.02).to(torch.bfloat16)
temp = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
peers = dgd.PeerBuffers(dgd.peer_bytes_needed(torch.bfloat16, n, a.chunks))
out = torch.empty((world, n), dtype=torch.bfloat16, device="cuda")
chunks = dgd._chunks(shard.view(-1), a.chunks)
stride = dgd._round_up(codec.get_max_float_compressed_size(2, chunks[0].numel()), 256)
rows = [dgd._Raw(peers.local + c * stride) for c in range(len(chunks))]
sizes = torch.zeros(len(chunks), dtype=torch.int32, device="cuda")
res = {"n_gpus": world, "shard_mib": a.mib, "chunks": a.chunks}
res["compress_ms"] = timed(lambda: codec.float_compress(2, chunks, rows, sizes, 10, False, temp=temp))
res["barrier_ms"] = timed(lambda: peers.barrier())
def dec(srcs):
    ins, outs = [], []
    for k, r in enumerate(srcs):
        off = 0
        for c, ch in enumerate(chunks):
            ins.append(dgd._Raw(peers.ptrs[r] + c * stride)); outs.append(out[k].narrow(0, off, ch.numel())); off += ch.numel()
    codec.float_decompress(2, ins, outs, None, None, 10, False, temp=temp)
res["decode_local_x_world_ms"] = timed(lambda: dec([rank] * world))
res["decode_remote_staggered_ms"] = timed(lambda: dec([(rank + k) % world for k in range(1, world + 1)]))
res["decode_one_remote_ms"] = timed(lambda: dec([(rank + 1) % world]))
res["decode_one_local_ms"] = timed(lambda: dec([rank]))
res["whole_call_ms"] = timed(lambda: dgd.all_gather_compressed_peer(shard, peers, out=out, num_chunks=a.chunks, temp=temp))
csz = int(sizes.sum().item())
res["remote_GBps_one"] = round(csz / res["decode_one_remote_ms"] / 1e6, 1)
if rank == 0: print(json.dumps(res), flush=True)
peers.close(); dist.destroy_process_group()
