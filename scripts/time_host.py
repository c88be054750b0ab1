"""Wall time of the host-buffer entry points on BASELINE configs[1] (bf16 128 x 4 MiB), compress and decompress
separately (bench.py's e2e times the pair): `python scripts/time_host.py [steps]`. For A/B runs of the pipeline's
knobs (DGB_HOST_LANES, DGB_HOST_GROUP_MB)."""
import sys
import time

import torch

sys.path.insert(0, ".")
from dietgpu_fork_b200 import codec  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 20
N, W = 128, 1 << 21
g = torch.Generator().manual_seed(2000)
host_in = (torch.randn((N, W), generator=g) * 0.02).to(torch.bfloat16).pin_memory()
host_out = torch.empty((N, W), dtype=torch.bfloat16).pin_memory()
mx = codec.get_max_float_compressed_size(2, W)
host_comp = torch.empty((N, mx), dtype=torch.uint8).pin_memory()
hin, hout, hcomp = [host_in[i] for i in range(N)], [host_out[i] for i in range(N)], [host_comp[i] for i in range(N)]


def timed(fn):
    fn()
    fn()
    torch.cuda.synchronize()
    t = []
    for _ in range(steps):
        t0 = time.perf_counter()
        fn()
        t.append(time.perf_counter() - t0)
    t.sort()
    return 1e3 * t[len(t) // 2], 1e3 * t[0]


csz = codec.float_compress_host(2, hin, hcomp, 10, False)
tc = timed(lambda: codec.float_compress_host(2, hin, hcomp, 10, False))
td = timed(lambda: codec.float_decompress_host(2, hcomp, csz, hout, 10, False))
assert torch.equal(host_out.view(torch.int16), host_in.view(torch.int16))
U, C = N * W * 2, sum(csz)
print(f"compress median {tc[0]:.2f} ms (min {tc[1]:.2f}; H2D {U / tc[0] / 1e6:.1f} GB/s)  "
      f"decompress median {td[0]:.2f} ms (min {td[1]:.2f}; D2H {U / td[0] / 1e6:.1f} GB/s)  "
      f"pair {2 * U / (tc[0] + td[0]) / 1e6:.1f} GB/s  ratio {C / U:.3f}")
