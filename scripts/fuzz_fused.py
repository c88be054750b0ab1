"""Randomised search for timing-dependent differences in the fused compress kernel: random small / ragged batch
shapes (including empty and one-word elements, which leave most warps of the persistent kernel without an encode
run), ANS bytes and f16 / bf16 / f32 words, checksum on and off; each shape compressed `reps` times under the fused
schedule and compared byte for byte with the separate-launch schedule's archives, then decoded reps / 3 times; bf16 /
f32 shapes without checksum also go through the sparse codec (90 % zeros), whose compaction has its own look-back.
`python scripts/fuzz_fused.py [shapes] [reps] [seed]` on the GPU. Prints every difference; exit code 1 if any."""
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import gpu_harness as H  # noqa: E402
from dietgpu_fork_b200 import _capi  # noqa: E402

SIZES = [0, 0, 1, 2, 31, 33, 4095, 4096, 4097, 8191, 32767, 32768, 32769, 70001, 262144, 300007, (1 << 20) + 5,
         3 * (1 << 20) + 77]


def make(ft, n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    if ft == 0:
        x = torch.clamp(torch.floor(-torch.log(torch.rand(max(n, 1), device="cuda", generator=g)) * 12.0), max=255)
        return x.to(torch.uint8)[:n]
    x = torch.randn(max(n, 1), device="cuda", generator=g) * (0.02 if ft != 3 else 1.0)
    if ft == 3:
        return x.view(torch.int32)[:n]
    return (x.to(torch.bfloat16) if ft == 2 else x.to(torch.float16)).view(torch.int16)[:n]


def compress(ft, ins, counts, cs):
    if ft == 0:
        return H.ours_ans_encode(ins, 10, checksum=cs, fill=0xA5)
    return H.ours_float_compress(ft, ins, counts, 10, checksum=cs, fill=0xA5)


def run(shapes, reps, seed):
    rng = np.random.default_rng(seed)
    bad = 0
    for s in range(shapes):
        ft = int(rng.integers(0, 4))
        k = int(rng.integers(1, 13))
        counts = [int(SIZES[i]) for i in rng.integers(0, len(SIZES), k)]
        if ft == 0:
            counts = [max(c, 1) for c in counts]  # the raw codec's harness sizes its rows from numel()
        cs = bool(rng.integers(0, 2))
        ins = [make(ft, n, 1000 * s + i) for i, n in enumerate(counts)]
        try:
            _capi.lib.dgb_set_fused_min_tiles(1 << 40)
            ms, ss = compress(ft, ins, counts, cs)
            _capi.lib.dgb_set_fused_min_tiles(0)
            for r in range(reps):
                mf, sf = compress(ft, ins, counts, cs)
                if list(sf) != list(ss):
                    bad += 1
                    print(f"shape {s} ft {ft} cs {cs} {counts} rep {r}: sizes {list(sf)} vs {list(ss)}", flush=True)
                    continue
                for i in range(k):
                    if not torch.equal(mf[i, :sf[i]], ms[i, :ss[i]]):
                        bad += 1
                        d = torch.nonzero(mf[i, :sf[i]] != ms[i, :ss[i]]).flatten()
                        print(f"shape {s} ft {ft} cs {cs} {counts} rep {r} element {i} (n={counts[i]}): {d.numel()} bytes "
                              f"differ, first at {int(d[0])}, last at {int(d[-1])}", flush=True)
        finally:
            _capi.lib.dgb_set_fused_min_tiles(-1)
        rows = [mf[i] for i in range(k)]
        wb = {0: 1, 1: 2, 2: 2, 3: 4}[ft]
        for r in range(max(1, reps // 3)):  # the decoder claims its tiles dynamically too
            if ft == 0:
                outs, ok, dsz = H.ours_ans_decode(rows, counts, 10, checksum=cs)
            else:
                outs, ok, dsz = H.ours_float_decompress(ft, rows, counts, 10, checksum=cs)
            for i, t in enumerate(ins):
                if not (ok[i] and torch.equal(outs[i][:counts[i] * wb], t.view(torch.uint8).reshape(-1))):
                    bad += 1
                    print(f"shape {s} ft {ft} cs {cs} {counts} decode {r} element {i}: round trip differs (ok {ok[i]})",
                          flush=True)
        if ft in (2, 3) and not cs and min(counts) >= 2:
            bad += sparse_round(s, ft, counts, reps)
    return bad


def sparse_round(s, ft, counts, reps):
    """The sparse codec's compaction uses a decoupled look-back between CTAs: same shape with 90 % zeros, `reps`
    compressions compared with the first, one decode."""
    bad = 0
    g = torch.Generator(device="cuda").manual_seed(5000 + s)
    ins = []
    for i, n in enumerate(counts):
        t = make(ft, n, 9000 * s + i).clone()
        t[torch.rand(n, device="cuda", generator=g) < 0.9] = 0
        ins.append(t)
    first, fs = H.ours_float_compress(ft, ins, counts, 10, sparse=True, fill=0xA5)
    for r in range(reps):
        m, sz = H.ours_float_compress(ft, ins, counts, 10, sparse=True, fill=0xA5)
        if list(sz) != list(fs) or any(not torch.equal(m[i, :sz[i]], first[i, :fs[i]]) for i in range(len(counts))):
            bad += 1
            print(f"shape {s} sparse ft {ft} {counts} rep {r}: archives differ from the first run", flush=True)
    outs, ok, dsz = H.ours_float_decompress(ft, [first[i] for i in range(len(counts))], counts, 10, sparse=True)
    wb = 2 if ft == 2 else 4
    for i, t in enumerate(ins):
        if not torch.equal(outs[i][:counts[i] * wb], t.view(torch.uint8).reshape(-1)):
            bad += 1
            print(f"shape {s} sparse ft {ft} {counts} element {i}: round trip differs", flush=True)
    return bad


if __name__ == "__main__":
    shapes = int(sys.argv[1]) if len(sys.argv) > 1 else 150
    reps = int(sys.argv[2]) if len(sys.argv) > 2 else 30
    seed = int(sys.argv[3]) if len(sys.argv) > 3 else 1
    n_bad = run(shapes, reps, seed)
    print("BAD:", n_bad)
    sys.exit(1 if n_bad else 0)
