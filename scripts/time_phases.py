"""Device time of compress and decompress of BASELINE configs[1] WITHOUT the per-launch profiling events (bench.py keeps
them on inside its timed region): `python scripts/time_phases.py [steps]`. For A/B runs of launch-level knobs."""
import sys

import torch

sys.path.insert(0, ".")
from dietgpu_fork_b200 import _capi, codec  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 30
g = torch.Generator(device="cuda").manual_seed(2000)
ins = [(torch.randn(1 << 21, device="cuda", generator=g) * 0.02).to(torch.bfloat16) for _ in range(128)]
mx = codec.get_max_float_compressed_size(2, 1 << 21)
comp = torch.empty((128, mx), dtype=torch.uint8, device="cuda")
rows = [comp[i] for i in range(128)]
sizes = torch.zeros(128, dtype=torch.int32, device="cuda")
outs = [torch.empty_like(t) for t in ins]
ok = torch.zeros(128, dtype=torch.uint8, device="cuda")
temp = torch.empty(_capi.lib.dgb_temp_bytes_required(2, 2, 128, 1 << 21), dtype=torch.uint8, device="cuda")


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / steps


tc = timed(lambda: codec.float_compress(2, ins, rows, sizes, 10, False, temp=temp))
td = timed(lambda: codec.float_decompress(2, rows, outs, ok, None, 10, False, temp=temp))
assert all(torch.equal(o.view(torch.int16), t.view(torch.int16)) for o, t in zip(outs, ins))
print(f"compress {tc:.4f} ms  decompress {td:.4f} ms  round trip {2 * 128 * (1 << 22) / (tc + td) / 1e6:.1f} GB/s")
