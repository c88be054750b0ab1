"""`-m gpu`: dgb_decode_batch_noalloc (include/dietgpu_b200.h; no counterpart in the reference)
against the existing decode entry points, through the C ABI, on two calls in a row.

The new entry point keeps the decoder's scratch in a fixed `static __device__` array and allocates nothing, so its
output must be bit-identical to dgb_float_decompress / dgb_ans_decode_batch_pointer on the same archives (integer /
byte work: the tolerance is zero, as everywhere in this suite). The second call follows the first on the same
stream with different data, a different batch size and different probability tables, so anything stale left in the
shared array by call one would show. Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

import datagen
import gpu_harness as H
from dietgpu_fork_b200 import _capi, codec

pytestmark = pytest.mark.gpu

BF16, F32, F64 = 2, 3, 4
WB = {0: 1, 1: 2, 2: 2, 3: 4, 4: 8}


def _noalloc_decode(ft, archives, caps, pb=10):
    """-> (outputs, success, sizes, rc). Outputs are allocated BEFORE the call; the call itself gets raw pointers."""
    n = len(archives)
    outs = [torch.zeros(max(c * WB[ft], 1), dtype=torch.uint8, device="cuda") for c in caps]
    ok = torch.zeros(max(n, 1), dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(max(n, 1), dtype=torch.int32, device="cuda")
    ins_p = (C.c_void_p * max(n, 1))(*[a.data_ptr() for a in archives])
    outs_p = (C.c_void_p * max(n, 1))(*[o.data_ptr() for o in outs])
    caps_p = (C.c_uint32 * max(n, 1))(*caps)
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    rc = _capi.lib.dgb_decode_batch_noalloc(ft, pb, n, ins_p, outs_p, caps_p, C.c_void_p(ok.data_ptr()),
                                            C.c_void_p(sizes.data_ptr()), stream)
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return outs, ok.cpu().numpy()[:n], sizes.cpu().numpy()[:n], rc, free0 - free1


def _check_float(ft, counts, seed, sigma, second_call):
    host = [datagen.float_words(ft, c, seed=seed + i, scale=sigma) for i, c in enumerate(counts)]
    dev = [H.to_dev(h) for h in host]
    m, sz = H.ours_float_compress(ft, dev, counts, 10)
    arch = [m[i] for i in range(len(counts))]
    ref_outs, ref_ok, ref_sz = H.ours_float_decompress(ft, arch, counts, 10)  # the existing entry point
    outs, ok, dsz, rc, grew = _noalloc_decode(ft, arch, counts)
    assert rc == 0
    assert list(ok) == list(ref_ok) and ok.all()
    assert list(dsz) == list(ref_sz) == list(counts)
    for i, h in enumerate(host):
        nb = counts[i] * WB[ft]
        assert torch.equal(outs[i][:nb], ref_outs[i][:nb]), f"element {i}: no-alloc decode differs from dgb_float_decompress"
        assert np.array_equal(outs[i][:nb].cpu().numpy(), h.view(np.uint8)), f"element {i}: round trip"
    if second_call:  # kernels are loaded by now: device memory in use must not have grown across the call
        assert grew <= 0, f"free device memory shrank by {grew} bytes across dgb_decode_batch_noalloc"


def test_noalloc_float_decode_two_calls_in_a_row():
    # call 1: a full batch (the maximum), ragged sizes incl. empty, tile tails and a multi-tile element
    counts1 = [0, 1, 33, 4096, 4097, 32768, 32769, 100001, 300007] + [20000 + 977 * i for i in range(23)]
    assert len(counts1) == _capi.NOALLOC_MAX_BATCH
    _check_float(BF16, counts1, seed=11, sigma=0.02, second_call=False)
    # call 2, straight after: fewer elements, other sizes, other exponent statistics (other tables)
    _check_float(BF16, [70001, 5, 1 << 20], seed=900, sigma=3.0, second_call=True)


def test_noalloc_other_types_and_raw_ans():
    _check_float(F32, [12345, 1, 65536 + 3], seed=40, sigma=1.0, second_call=False)
    _check_float(F64, [8083, 40000], seed=50, sigma=1.0, second_call=False)  # two exponent planes
    # raw ANS bytes (float_type 0), twice
    for rep, (sizes, lam) in enumerate([([1, 4096, 70001, 1 << 20], 20.0), ([10013, 333], 2.0)]):
        host = [datagen.exp_bytes(n, lam, seed=60 + 10 * rep + i) for i, n in enumerate(sizes)]
        dev = [H.to_dev(h) for h in host]
        m, sz = H.ours_ans_encode(dev, 10)
        arch = [m[i] for i in range(len(sizes))]
        ref_outs, ref_ok, ref_sz = H.ours_ans_decode(arch, sizes, 10)
        outs, ok, dsz, rc, _ = _noalloc_decode(0, arch, sizes)
        assert rc == 0 and ok.all() and list(dsz) == list(ref_sz) == sizes
        for i in range(len(sizes)):
            assert torch.equal(outs[i][:sizes[i]], ref_outs[i]), f"ANS element {i} (call {rep})"


def test_noalloc_capacity_too_small_is_reported_like_the_existing_call():
    host = [datagen.float_words(BF16, 5000, seed=1), datagen.float_words(BF16, 6000, seed=2)]
    m, sz = H.ours_float_compress(BF16, [H.to_dev(h) for h in host], [5000, 6000], 10)
    arch = [m[0], m[1]]
    _, ref_ok, ref_sz = H.ours_float_decompress(BF16, arch, [5000, 100], 10)
    outs, ok, dsz, rc, _ = _noalloc_decode(BF16, arch, [5000, 100])
    assert rc == 0 and list(ok) == list(ref_ok) == [1, 0] and list(dsz) == list(ref_sz) == [5000, 6000]
    assert np.array_equal(outs[0].cpu().numpy(), host[0].view(np.uint8))


def test_noalloc_rejects_batches_beyond_the_maximum():
    n = _capi.NOALLOC_MAX_BATCH + 1
    host = datagen.float_words(BF16, 1000, seed=3)
    m, sz = H.ours_float_compress(BF16, [H.to_dev(host)], [1000], 10)
    launches0 = _capi.lib.dgb_profile_launches()
    outs, ok, dsz, rc, _ = _noalloc_decode(BF16, [m[0]] * n, [1000] * n)
    assert rc == -5  # DGB_ERR_TOO_LARGE
    assert _capi.lib.dgb_profile_launches() == launches0, "a rejected call must launch nothing"
    assert not any(o.any() for o in outs)
    with pytest.raises(_capi.DietGpuError, match="too large"):
        _capi.check(rc, "dgb_decode_batch_noalloc")
