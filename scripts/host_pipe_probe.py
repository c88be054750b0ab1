"""Times dgb_float_compress_host / dgb_float_decompress_host separately on bench.py's workload (bf16, 128 x 4 MiB,
pinned host memory). DGB_HOST_GROUP_MB sets the group size."""
import os, sys, time, json
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from dietgpu_fork_b200 import codec
B, W = 128, 2 * 1024 * 1024
torch.manual_seed(1)
host_in = (torch.randn(B, W) * 0.02).to(torch.bfloat16).pin_memory()
mx = codec.get_max_float_compressed_size(2, W)
host_comp = torch.empty((B, mx), dtype=torch.uint8).pin_memory()
host_out = torch.empty((B, W), dtype=torch.bfloat16).pin_memory()
ins = [host_in[i] for i in range(B)]; comps = [host_comp[i] for i in range(B)]; outs = [host_out[i] for i in range(B)]
torch.cuda.init()
res = {"group_mb": os.environ.get("DGB_HOST_GROUP_MB", "32")}
for it in range(4):
    t0 = time.perf_counter(); csz = codec.float_compress_host(2, ins, comps); t1 = time.perf_counter()
    codec.float_decompress_host(2, comps, csz, outs); t2 = time.perf_counter()
    res[f"it{it}"] = {"compress_ms": round((t1 - t0) * 1e3, 2), "decompress_ms": round((t2 - t1) * 1e3, 2)}
assert torch.equal(host_out.view(torch.int16), host_in.view(torch.int16))
U = B * W * 2
res["ideal_ms_each_at_48GBps"] = round(U / 48e9 * 1e3, 2)
print(json.dumps(res))
