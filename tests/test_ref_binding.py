"""Drop-in check at the reference's own boundary: the reference's torch binding translation unit
(/root/reference/dietgpu/DietGpu.cpp) compiled UNMODIFIED against include/compat + include/dietgpu_b200.hpp and
linked with libdietgpu_b200.so (oracle/build_ref_torch_binding.sh -> oracle/_ref/libdietgpu_torch_b200.so).
The test loads it with torch.ops.load_library, exactly as the reference's ans_test.py / float_test.py do, in a
fresh interpreter (the `dietgpu` operator namespace can be defined once per process), and checks the archives
against the CPU oracle byte for byte."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "libdietgpu_torch_b200.so")

SCRIPT = r'''
import sys, numpy as np, torch
sys.path.insert(0, %(root)r)
import oracle
torch.ops.load_library(%(lib)r)
ops = torch.ops.dietgpu
dev = torch.device("cuda:0")
temp = torch.empty([64 * 1024 * 1024], dtype=torch.uint8, device=dev)
torch.manual_seed(1)
FT = {torch.float16: oracle.FLOAT16, torch.bfloat16: oracle.BFLOAT16, torch.float32: oracle.FLOAT32}
NP = {torch.float16: (torch.int16, np.uint16), torch.bfloat16: (torch.int16, np.uint16), torch.float32: (torch.int32, np.uint32)}
for dt in (torch.bfloat16, torch.float16, torch.float32):
    ts = [torch.normal(0, 1.0, [n], dtype=dt, device=dev) for n in (10000, 100000, 1000000)]
    for tm in (None, temp):
        comp, sizes, used = ops.compress_data(True, ts, True, tm)
        trunc = [t.narrow(0, 0, s.item()).clone() for s, t in zip(sizes, [*comp])]
        for t, c in zip(ts, trunc):
            tv, nv = NP[dt]
            want = oracle.float_compress(FT[dt], t.view(tv).cpu().numpy().view(nv), 10, True)
            got = c.cpu().numpy()
            assert got.size == want.size and (got == want).all(), "archive differs from the oracle"
        outs = [torch.empty_like(t) for t in ts]
        st = torch.empty([len(ts)], dtype=torch.uint8, device=dev)
        sz = torch.empty([len(ts)], dtype=torch.int32, device=dev)
        ops.decompress_data(True, trunc, outs, True, tm, st, sz)
        assert st.cpu().all() and sz.cpu().tolist() == [t.numel() for t in ts]
        for a, b in zip(ts, outs):
            assert torch.equal(a, b)
    cts = ops.compress_data_simple(True, ts, True)
    dts = ops.decompress_data_simple(True, cts, True)
    assert all(torch.equal(a, b) for a, b in zip(ts, dts))
    # split-size variants
    sizes = [1000, 8, 4096, 77777]
    t = torch.normal(0, 1.0, [sum(sizes)], dtype=dt, device=dev)
    comp_ts, _, _ = ops.compress_data_split_size(True, t, torch.IntTensor(sizes), True, temp)
    back = ops.decompress_data_simple(True, comp_ts, True)
    assert all(torch.equal(a, b) for a, b in zip(torch.split(t, sizes), back))
    out = torch.empty_like(t)
    ops.decompress_data_split_size(True, comp_ts, out, torch.IntTensor(sizes), True, temp)
    assert torch.equal(t, out)
# raw ANS
ts = [torch.randint(0, 65, [n], dtype=torch.uint8, device=dev) for n in (4, 4096, 123457)]
comp, sizes, _ = ops.compress_data(False, ts, True, temp)
for t, c, s in zip(ts, [*comp], sizes):
    want = oracle.ans_encode(t.cpu().numpy(), 10, True)
    got = c[: s.item()].cpu().numpy()
    assert got.size == want.size and (got == want).all()
outs = [torch.empty_like(t) for t in ts]
ops.decompress_data(False, [c[: s.item()].clone() for c, s in zip([*comp], sizes)], outs, True, temp)
assert all(torch.equal(a, b) for a, b in zip(ts, outs))
e = [torch.empty([0], dtype=torch.uint8, device=dev)]
ce = ops.compress_data_simple(False, e, True)
assert ce[0].numel() > 0 and torch.equal(ops.decompress_data_simple(False, ce, True)[0], e[0])
assert ops.max_any_compressed_size(1 << 20) == oracle.ans_max_compressed_size(1 << 20)
print("reference binding over libdietgpu_b200: OK")
'''


@pytest.mark.gpu
def test_reference_torch_binding_over_our_library():
    if not os.path.exists(LIB):
        pytest.skip("oracle/_ref/libdietgpu_torch_b200.so not built (needs /root/reference at build time)")
    r = subprocess.run([sys.executable, "-c", SCRIPT % {"root": ROOT, "lib": LIB}], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "OK" in r.stdout
