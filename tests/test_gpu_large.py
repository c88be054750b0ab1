"""`-m gpu` tests at sizes the small parity cases do not reach: the schedules and modes that only large inputs
select (fused kernel in its few-huge-elements mode: sliced split claims, 9 split CTAs per SM; stand-alone split
dealing work round robin over the list or inside every element; in-kernel compaction of many tiles and groups, in
the fused kernel). Every case: both compress schedules
byte-identical, the CPU oracle byte for byte where it finishes in seconds, the reference's CUDA build (when built)
under its defined-byte mask, round trip bit exact."""
import numpy as np
import pytest
import torch

import oracle
import gpu_harness as H
from dietgpu_fork_b200 import _capi

pytestmark = pytest.mark.gpu

F16, BF16, F32 = 1, 2, 3
NP_WORD = {F16: np.uint16, BF16: np.uint16, F32: np.uint32}


def _words(ft, n, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, device="cuda", generator=g) * (0.02 if ft != F32 else 1.0)
    if ft == F32:
        return x.view(torch.int32)
    return (x.to(torch.bfloat16) if ft == BF16 else x.to(torch.float16)).view(torch.int16)


def _compress_both(ft, ins, counts, pb=10):
    out = {}
    for name, tiles in (("fused", 0), ("separate", 1 << 40)):
        _capi.lib.dgb_set_fused_min_tiles(tiles)
        try:
            m, sz = H.ours_float_compress(ft, ins, counts, pb)
        finally:
            _capi.lib.dgb_set_fused_min_tiles(-1)
        out[name] = (m, sz)
    (mf, sf), (ms, ss) = out["fused"], out["separate"]
    assert list(sf) == list(ss), "compressed sizes differ between the schedules"
    for i in range(len(ins)):
        assert torch.equal(mf[i, :sf[i]], ms[i, :ss[i]]), f"archive {i}: fused != separate"
    return mf, sf


CASES = [
    # (float type, word counts): one huge element (sliced fused mode / round robin over the list) ...
    (BF16, [20 * 1024 * 1024 + 12345]),
    # ... one of more than 2048 tiles (more blocks than two rounds of resident warps in the separate schedule)
    (BF16, [66 * 1024 * 1024 + 4321]),
    (F32, [17 * 1024 * 1024 + 3]),
    # ... several long elements (round robin inside every element; fused with few elements) ...
    (BF16, [5 * 1024 * 1024 + 77, 6 * 1024 * 1024, 4 * 1024 * 1024 + 4095, 7 * 1024 * 1024 + 1]),
    # ... and a ragged mix of long and tiny ones
    (F16, [9 * 1024 * 1024, 1, 0, 4097, 3 * 1024 * 1024 + 5, 33]),
]


@pytest.mark.parametrize("ft,counts", CASES, ids=lambda v: str(v) if isinstance(v, int) else "x".join(f"{c >> 20}M" for c in v))
def test_large_inputs_both_schedules(ft, counts):
    ins = [_words(ft, n, 700 + i) if n else torch.zeros(0, dtype=torch.int16 if ft != F32 else torch.int32, device="cuda")
           for i, n in enumerate(counts)]
    m, sz = _compress_both(ft, ins, counts)
    # oracle on the smallest non-trivial element and on the first one if it is not too long for the CPU
    check = {int(np.argmin([c if c > 4096 else 1 << 62 for c in counts]))}
    if counts[0] <= 24 * 1024 * 1024:
        check.add(0)
    for i in sorted(check):
        host = ins[i].cpu().numpy().view(NP_WORD[ft])
        exp = oracle.float_compress(ft, host, 10)
        got = H.row_bytes(m, i, sz[i])
        assert got.size == exp.size and np.array_equal(got, exp), f"archive {i} differs from the oracle"
    if H.ref_lib() is not None:
        rm, rsz = H.ref_float_compress(ft, ins, counts, 10)
        assert list(rsz) == list(sz)
        for i in range(len(ins)):
            a, b = H.row_bytes(m, i, sz[i]), H.row_bytes(rm, i, rsz[i])
            mask = oracle.float_defined_mask(a)
            assert np.array_equal(a[mask], b[mask]), f"archive {i} differs from the reference build"
    outs, ok, dsz = H.ours_float_decompress(ft, [m[i] for i in range(len(ins))], counts)
    assert ok.all() and list(dsz) == list(counts)
    for i, t in enumerate(ins):
        assert torch.equal(outs[i].view(t.dtype)[:t.numel()], t), f"round trip {i}"


def test_large_ans_both_schedules():
    rng = np.random.default_rng(11)
    sizes = [40 * 1024 * 1024 + 9, 3 * 1024 * 1024 + 1]
    host = [np.minimum(rng.exponential(1 / 20.0, n), 1.0) for n in sizes]
    host = [(h * 255.99).astype(np.uint8) for h in host]
    dev = [torch.from_numpy(h).cuda() for h in host]
    res = {}
    for name, tiles in (("fused", 0), ("separate", 1 << 40)):
        _capi.lib.dgb_set_fused_min_tiles(tiles)
        try:
            res[name] = H.ours_ans_encode(dev, 10)
        finally:
            _capi.lib.dgb_set_fused_min_tiles(-1)
    (mf, sf), (ms, ss) = res["fused"], res["separate"]
    assert list(sf) == list(ss)
    for i in range(len(dev)):
        assert torch.equal(mf[i, :sf[i]], ms[i, :ss[i]]), f"ANS archive {i}: fused != separate"
    exp = oracle.ans_encode(host[1], 10)
    assert np.array_equal(H.row_bytes(mf, 1, sf[1]), exp)
    outs, ok, dsz = H.ours_ans_decode([mf[i] for i in range(len(dev))], sizes, 10)
    assert ok.all() and list(dsz) == sizes
    for i, d in enumerate(dev):
        assert torch.equal(outs[i], d), f"ANS round trip {i}"
