"""Repeat the fused compress of tests/test_gpu_large.py's cases many times and compare every archive with the
separate-launch schedule's, byte for byte: a race in the persistent kernel shows as an occasional difference.
`python scripts/stress_fused.py [iterations]` on the GPU; prints where the first difference of each bad run lies."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import gpu_harness as H  # noqa: E402
from dietgpu_fork_b200 import _capi  # noqa: E402

F16, BF16, F32 = 1, 2, 3
iters = int(sys.argv[1]) if len(sys.argv) > 1 else 300


def words(ft, n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, device="cuda", generator=g) * (0.02 if ft != F32 else 1.0)
    if ft == F32:
        return x.view(torch.int32)
    return (x.to(torch.bfloat16) if ft == BF16 else x.to(torch.float16)).view(torch.int16)


CASES = [
    (F16, [9 * 1024 * 1024, 1, 0, 4097, 3 * 1024 * 1024 + 5, 33]),
    (BF16, [5 * 1024 * 1024 + 77, 6 * 1024 * 1024, 4 * 1024 * 1024 + 4095, 7 * 1024 * 1024 + 1]),
    (BF16, [20 * 1024 * 1024 + 12345]),
    (BF16, [300000 + 1000 * i for i in range(40)]),
]


def describe(ft, n, a, b):
    d = np.nonzero(a != b)[0]
    off = int(d[0])
    raw = n if ft != F32 else 3 * n
    raw = (raw + 15) // 16 * 16 if ft != F32 else ((2 * n + 15) // 16 * 16 + (n + 15) // 16 * 16)
    ans0 = 32 + raw
    nb = (n + 4095) // 4096
    where = "float header" if off < 32 else "raw section" if off < ans0 else None
    if where is None:
        r = off - ans0
        where = ("ANS header" if r < 32 else "pdf table" if r < 32 + 512 else "lane states" if r < 32 + 512 + 128 * nb
                 else "block index" if r < 32 + 512 + 128 * nb + 8 * ((nb + 1) // 2 * 2) else "payload")
    return f"{len(d)} bytes differ, first at {off} ({where}), last at {int(d[-1])}"


bad = 0
for ft, counts in CASES:
    ins = [words(ft, n, 700 + i) if n else torch.zeros(0, dtype=torch.int16 if ft != F32 else torch.int32, device="cuda")
           for i, n in enumerate(counts)]
    _capi.lib.dgb_set_fused_min_tiles(1 << 40)
    ms, ss = H.ours_float_compress(ft, ins, counts, 10)
    ms = ms.clone()
    _capi.lib.dgb_set_fused_min_tiles(0)
    for it in range(iters):
        mf, sf = H.ours_float_compress(ft, ins, counts, 10)
        if list(sf) != list(ss):
            bad += 1
            print(f"case {counts[:3]}... iteration {it}: sizes differ {list(sf)} vs {list(ss)}", flush=True)
            continue
        for i in range(len(ins)):
            if not torch.equal(mf[i, :sf[i]], ms[i, :ss[i]]):
                bad += 1
                print(f"case {counts[:3]}... iteration {it} element {i} (n={counts[i]}): "
                      + describe(ft, counts[i], mf[i, :sf[i]].cpu().numpy(), ms[i, :ss[i]].cpu().numpy()), flush=True)
    _capi.lib.dgb_set_fused_min_tiles(-1)
    print(f"case ft={ft} {counts[:4]}...: done", flush=True)
print("BAD RUNS:", bad)
