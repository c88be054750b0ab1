"""`-m gpu` byte parity at the BASELINE.json shapes themselves, in the DEFAULT compress schedule (no
dgb_set_fused_min_tiles: what bench.py and a user run), over EVERY batch element: archives equal the
reference's own CUDA build (oracle/_ref) under the defined-byte mask, sizes equal, both libraries decode
each other's archives, the round trip is bit exact; the CPU oracle byte for byte where it finishes in
seconds. Plus the entry points the small parity cases never reached: Stride addressing
(dietgpu/ans/ANSTest.cu:171-241, :277-282), caller-supplied histograms (dietgpu/ans/GpuANSEncode.cuh:688-711),
the info calls with their checksum output, and the size bounds against the reference build's.

Sizes follow BASELINE.json `configs` and the reference's large-batch test (dietgpu/float/FloatTest.cu:316-328:
256 x 512 Ki words, all four float types)."""
import numpy as np
import pytest
import torch

import datagen
import gpu_harness as H
import oracle
from dietgpu_fork_b200 import _capi, codec

pytestmark = pytest.mark.gpu

F16, BF16, F32, F64 = 1, 2, 3, 4
NP_WORD = {F16: np.uint16, BF16: np.uint16, F32: np.uint32, F64: np.uint64}
T_WORD = {F16: torch.int16, BF16: torch.int16, F32: torch.int32, F64: torch.int64}
needs_ref = pytest.mark.skipif(H.ref_lib() is None, reason="oracle/_ref/libdietgpu_ref.so not built")


@pytest.fixture(autouse=True)
def default_schedule():
    _capi.lib.dgb_set_fused_min_tiles(-1)  # the library's own choice, as in bench.py
    _capi.lib.dgb_profile_enable(1)
    yield
    _capi.lib.dgb_profile_enable(0)
    torch.cuda.empty_cache()


def _kernels_run():
    import ctypes as C
    buf = C.create_string_buffer(1 << 16)
    _capi.lib.dgb_profile_report(buf, len(buf))
    return {ln.split()[0] for ln in buf.value.decode().splitlines() if ln.strip()}


def _gauss_words(ft, n, seed, sigma):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(n, device="cuda", generator=g, dtype=torch.float64 if ft == F64 else torch.float32) * sigma
    if ft == F64:
        return x.view(torch.int64)
    if ft == F32:
        return x.view(torch.int32)
    return (x.to(torch.bfloat16) if ft == BF16 else x.to(torch.float16)).view(torch.int16)


def _float_case(ft, counts, sigma, oracle_elems, expect_kernel=None):
    H.ref_temp(3 << 30)
    ins = [_gauss_words(ft, n, 9000 + i, sigma) for i, n in enumerate(counts)]
    m, sz = H.ours_float_compress(ft, ins, counts, 10)
    if expect_kernel:
        ran = _kernels_run()
        assert expect_kernel in ran, f"{expect_kernel} did not run: {ran}"
    rows = [m[i] for i in range(len(ins))]
    # ---- the reference build: every archive, under the defined-byte mask ----
    if H.ref_lib() is not None:
        rm, rsz = H.ref_float_compress(ft, ins, counts, 10)
        H.assert_rows_match_ref(m, sz, rm, rsz, oracle.float_defined_mask, f"ft {ft}")
        rrows = [rm[i] for i in range(len(ins))]
        outs, ok, dsz = H.ref_float_decompress(ft, rows, counts, 10)  # the reference decodes ours
        assert ok.all() and list(dsz) == list(counts)
        for i, t in enumerate(ins):
            assert torch.equal(outs[i].view(t.dtype), t), f"reference decode of our archive {i}"
        del outs
        outs, ok, dsz = H.ours_float_decompress(ft, rrows, counts, 10)  # we decode the reference's
        assert ok.all() and list(dsz) == list(counts)
        for i, t in enumerate(ins):
            assert torch.equal(outs[i].view(t.dtype), t), f"our decode of the reference's archive {i}"
        del outs, rm, rrows
    # ---- the CPU oracle, byte for byte ----
    for i in oracle_elems:
        exp = oracle.float_compress(ft, ins[i].cpu().numpy().view(NP_WORD[ft]), 10)
        got = H.row_bytes(m, i, sz[i])
        assert got.size == exp.size and np.array_equal(got, exp), f"archive {i} differs from the oracle"
    # ---- full round trip of every element ----
    outs, ok, dsz = H.ours_float_decompress(ft, rows, counts, 10)
    assert ok.all() and list(dsz) == list(counts)
    for i, t in enumerate(ins):
        assert torch.equal(outs[i].view(t.dtype), t), f"round trip {i}"


def test_config1_ans_16x1MiB():
    """BASELINE configs[0]: ANS byte codec, 16 x 1 MiB, exponential bytes (lambda = 20), probBits 10."""
    n = 1 << 20
    host = [datagen.exp_bytes(n, 20.0, seed=500 + i) for i in range(16)]
    dev = [H.to_dev(h) for h in host]
    m, sz = H.ours_ans_encode(dev, 10)
    for i, h in enumerate(host):
        exp = oracle.ans_encode(h, 10)
        got = H.row_bytes(m, i, sz[i])
        assert got.size == exp.size and np.array_equal(got, exp), f"ANS archive {i} differs from the oracle"
    arch = [m[i] for i in range(16)]
    if H.ref_lib() is not None:
        rm, rsz = H.ref_ans_encode(dev, 10)
        H.assert_rows_match_ref(m, sz, rm, rsz, oracle.ans_defined_mask, "ANS 16 x 1 MiB")
        outs, ok, _ = H.ref_ans_decode(arch, [n] * 16, 10)
        assert ok.all() and all(torch.equal(outs[i], dev[i]) for i in range(16))
        outs, ok, _ = H.ours_ans_decode([rm[i] for i in range(16)], [n] * 16, 10)
        assert ok.all() and all(torch.equal(outs[i], dev[i]) for i in range(16))
    outs, ok, dsz = H.ours_ans_decode(arch, [n] * 16, 10)
    assert ok.all() and list(dsz) == [n] * 16
    assert all(torch.equal(outs[i], dev[i]) for i in range(16))


def test_config2_bf16_128x4MiB():
    """BASELINE configs[1], the benchmarked shape: the fused persistent kernel under full load."""
    _float_case(BF16, [2 << 20] * 128, 0.02, oracle_elems=range(0, 128, 9), expect_kernel="k_compressFused")


def test_config2_ragged_bf16_batch():
    """The same regime with ragged element sizes (tile and block tails inside a loaded fused launch)."""
    rng = np.random.default_rng(77)
    counts = [int(x) for x in rng.integers(1, 3 << 20, 96)] + [0, 1, 4096, 32768, 32769]
    _float_case(BF16, counts, 0.02, oracle_elems=[0, 5, 50, 96, 97, 98, 99, 100], expect_kernel="k_compressFused")


def test_config3_fp32_single_1GiB():
    """BASELINE configs[2](i): one 1 GiB fp32 element (sliced split claims, compaction between runs)."""
    _float_case(F32, [256 << 20], 1.0, oracle_elems=[0], expect_kernel="k_compressFused")


def test_config3_fp32_64x16MiB():
    """BASELINE configs[2](ii): the 64 x 16 MiB batch that dist.compress_sharded cuts across ranks (here on one GPU)."""
    _float_case(F32, [4 << 20] * 64, 1.0, oracle_elems=[0, 31, 63], expect_kernel="k_compressFused")


def test_config4_fp64_16x1Mi():
    """BASELINE configs[3]: fp64 two-round exponent coding, 16 x 1 Mi words."""
    _float_case(F64, [1 << 20] * 16, 1.0, oracle_elems=range(16))


@pytest.mark.parametrize("ft", [F16, BF16, F32, F64])
def test_reference_large_batch_256x512Ki(ft):
    """dietgpu/float/FloatTest.cu:316-328: 256 x 512 Ki words for every float type."""
    _float_case(ft, [512 * 1024] * 256, 1.0, oracle_elems=[0, 100, 255])


def test_config4_sparse_fp32_16x4Mi():
    """BASELINE configs[3]: sparse fp32, 16 x 4 Mi words, 90 % zeros. Word N-2 is made non-zero so that the
    reference's stream has no uninitialised slot (sparse quirk)."""
    H.ref_temp(3 << 30)
    n = 4 << 20
    ins = []
    for i in range(16):
        g = torch.Generator(device="cuda").manual_seed(300 + i)
        x = torch.randn(n, device="cuda", generator=g)
        x = torch.where(torch.rand(n, device="cuda", generator=g) < 0.9, torch.zeros_like(x), x)
        x[-2], x[-1] = 1.0, 1.5
        ins.append(x.view(torch.int32))
    counts = [n] * 16
    m, sz = H.ours_float_compress(F32, ins, counts, 10, sparse=True)
    rows = [m[i] for i in range(16)]
    for i in (0, 7, 15):
        exp = oracle.sparse_float_compress(F32, ins[i].cpu().numpy().view(np.uint32), 10)
        got = H.row_bytes(m, i, sz[i])
        assert got.size == exp.size and np.array_equal(got, exp), f"sparse archive {i} differs from the oracle"
    if H.ref_lib() is not None:
        rm, rsz = H.ref_float_compress(F32, ins, counts, 10, sparse=True)
        H.assert_rows_match_ref(m, sz, rm, rsz, lambda a: oracle.sparse_float_defined_mask(F32, a), "sparse fp32")
        outs, ok, _ = H.ref_float_decompress(F32, rows, counts, 10, sparse=True)
        assert all(torch.equal(outs[i].view(torch.int32), ins[i]) for i in range(16)), "reference decode of ours"
        outs, ok, _ = H.ours_float_decompress(F32, [rm[i] for i in range(16)], counts, 10, sparse=True)
        assert all(torch.equal(outs[i].view(torch.int32), ins[i]) for i in range(16)), "our decode of the reference's"
    outs, ok, dsz = H.ours_float_decompress(F32, rows, counts, 10, sparse=True)
    assert ok.all() and list(dsz) == counts
    assert all(torch.equal(outs[i].view(torch.int32), ins[i]) for i in range(16))


# ------------------------------------------------------------------------------------ Stride addressing
@pytest.mark.parametrize("num,size", [(13, 8192 + 16), (5, 70000), (300, 4096 + 48)])
def test_ans_batch_stride(num, size):
    """runBatchStride, dietgpu/ans/ANSTest.cu:171-241 (case :277-282: 13 x (8192 + 16), checksum on)."""
    host = datagen.exp_bytes(num * size, 100.0, seed=10)
    in_dev = H.to_dev(host)
    stride_out = codec.get_max_compressed_size(size)
    enc = torch.zeros(num * stride_out, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(num, dtype=torch.int32, device="cuda")
    codec.ans_encode_batch_stride(in_dev, num, size, size, enc, stride_out, sizes, 10, True)
    torch.cuda.synchronize()
    sz = sizes.cpu().numpy()
    assert (sz % 16 == 0).all()
    enc_h = enc.cpu().numpy().reshape(num, stride_out)
    for i in range(num):
        exp = oracle.ans_encode(host[i * size:(i + 1) * size], 10, use_checksum=True)
        assert sz[i] == exp.size and np.array_equal(enc_h[i, :sz[i]], exp), f"stride archive {i} differs from the oracle"
    dec = torch.zeros(num * size, dtype=torch.uint8, device="cuda")
    ok = torch.zeros(num, dtype=torch.uint8, device="cuda")
    dsz = torch.zeros(num, dtype=torch.int32, device="cuda")
    codec.ans_decode_batch_stride(enc, num, stride_out, dec, size, size, ok, dsz, 10, True)
    torch.cuda.synchronize()
    assert ok.cpu().numpy().all() and (dsz.cpu().numpy() == size).all()
    assert torch.equal(dec, in_dev)
    if H.ref_lib() is not None:
        renc = torch.full((num * stride_out,), 0xCD, dtype=torch.uint8, device="cuda")
        rsizes = torch.zeros(num, dtype=torch.int32, device="cuda")
        H.ref_ans_encode_stride(in_dev, num, size, size, renc, stride_out, rsizes, 10, True)
        assert list(rsizes.cpu().numpy()) == list(sz)
        renc_h = renc.cpu().numpy().reshape(num, stride_out)
        for i in range(num):
            mask = oracle.ans_defined_mask(enc_h[i, :sz[i]])
            assert np.array_equal(enc_h[i, :sz[i]][mask], renc_h[i, :sz[i]][mask]), f"stride archive {i} vs reference"
        # the reference decodes our strided archives, we decode its
        rdec = torch.zeros(num * size, dtype=torch.uint8, device="cuda")
        rc = H.ref_ans_decode_stride(enc, num, stride_out, rdec, size, size, ok, dsz, 10, True)
        assert rc == 0 and torch.equal(rdec, in_dev)
        dec.zero_()
        codec.ans_decode_batch_stride(renc, num, stride_out, dec, size, size, ok, dsz, 10, True)
        torch.cuda.synchronize()
        assert ok.cpu().numpy().all() and torch.equal(dec, in_dev)


def test_ans_batch_stride_capacity_too_small():
    num, size = 4, 5000
    in_dev = H.to_dev(datagen.exp_bytes(num * size, 50.0, seed=3))
    stride_out = codec.get_max_compressed_size(size)
    enc = torch.zeros(num * stride_out, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(num, dtype=torch.int32, device="cuda")
    codec.ans_encode_batch_stride(in_dev, num, size, size, enc, stride_out, sizes, 10, False)
    dec = torch.zeros(num * size, dtype=torch.uint8, device="cuda")
    ok = torch.ones(num, dtype=torch.uint8, device="cuda")
    dsz = torch.zeros(num, dtype=torch.int32, device="cuda")
    codec.ans_decode_batch_stride(enc, num, stride_out, dec, size, size - 1, ok, dsz, 10, False)
    torch.cuda.synchronize()
    assert not ok.cpu().numpy().any() and (dsz.cpu().numpy() == size).all()  # required size reported


# ------------------------------------------------------------------------------------ caller-supplied histogram
@pytest.mark.parametrize("sizes", [[70001, 4096, 1 << 20], [3 << 20] * 12])
def test_ans_caller_supplied_histogram(sizes):
    """histogram_dev != nullptr: the tables come from the caller's counts (k_normalize), not from a pass over the
    data (dietgpu/ans/GpuANSEncode.cuh:688-697). Exact counts must reproduce the ordinary archive; the reference
    build fed the same histogram must agree."""
    host = [datagen.exp_bytes(n, 20.0, seed=800 + i) for i, n in enumerate(sizes)]
    dev = [H.to_dev(h) for h in host]
    hist = np.stack([np.bincount(h, minlength=256).astype(np.uint32) for h in host])
    hist_dev = torch.from_numpy(hist.view(np.int32)).cuda()
    m, sz = H.ours_ans_encode(dev, 10, histogram=hist_dev)
    ran = _kernels_run()
    assert "k_normalize" in ran, f"k_normalize did not run: {ran}"
    m0, sz0 = H.ours_ans_encode(dev, 10)
    assert list(sz) == list(sz0)
    for i, h in enumerate(host):
        assert torch.equal(m[i, :sz[i]], m0[i, :sz0[i]]), f"archive {i}: caller histogram != computed histogram"
        exp = oracle.ans_encode(h, 10, hist=hist[i])
        assert np.array_equal(H.row_bytes(m, i, sz[i]), exp), f"archive {i} differs from the oracle"
    if H.ref_lib() is not None:
        rm, rsz = H.ref_ans_encode(dev, 10, histogram=hist_dev)
        H.assert_rows_match_ref(m, sz, rm, rsz, oracle.ans_defined_mask, "caller histogram")
    outs, ok, _ = H.ours_ans_decode([m[i] for i in range(len(sizes))], sizes, 10)
    assert ok.all() and all(torch.equal(outs[i], dev[i]) for i in range(len(sizes)))


# ------------------------------------------------------------------------------------ info calls
def test_ans_compressed_info_sizes_and_checksums():
    sizes = [1, 4096, 10013, 70001, 300000]
    host = [datagen.exp_bytes(n, 10.0, seed=900 + i) for i, n in enumerate(sizes)]
    dev = [H.to_dev(h) for h in host]
    m, sz = H.ours_ans_encode(dev, 10, checksum=True)
    arch = [m[i] for i in range(len(sizes))]
    osz = torch.zeros(len(sizes), dtype=torch.int32, device="cuda")
    ock = torch.zeros(len(sizes), dtype=torch.int32, device="cuda")
    codec.ans_get_compressed_info(arch, osz, ock)
    torch.cuda.synchronize()
    assert list(osz.cpu().numpy()) == sizes
    ck = ock.cpu().numpy().view(np.uint32)
    for i, h in enumerate(host):
        assert ck[i] == oracle.checksum(h), f"stored checksum {i}"
    if H.ref_lib() is not None:
        rsz, rck = H.ref_ans_info(arch)
        assert list(rsz) == sizes and list(rck.view(np.uint32)) == list(ck)


@pytest.mark.parametrize("ft", [F16, BF16, F32, F64])
def test_float_compressed_info_sizes_types_checksums(ft):
    sizes = [13, 4097, 70001]
    host = [datagen.float_words(ft, n, seed=950 + i) for i, n in enumerate(sizes)]
    dev = [H.to_dev(h) for h in host]
    m, sz = H.ours_float_compress(ft, dev, sizes, 10, checksum=True)
    arch = [m[i] for i in range(len(sizes))]
    osz = torch.zeros(len(sizes), dtype=torch.int32, device="cuda")
    oty = torch.zeros(len(sizes), dtype=torch.int32, device="cuda")
    ock = torch.zeros(len(sizes), dtype=torch.int32, device="cuda")
    codec.float_get_compressed_info(arch, osz, oty, ock)
    torch.cuda.synchronize()
    assert list(osz.cpu().numpy()) == sizes and list(oty.cpu().numpy()) == [ft] * len(sizes)
    ck = ock.cpu().numpy().view(np.uint32)
    for i, h in enumerate(host):
        exp = oracle.float_compress(ft, h, 10, use_checksum=True)
        assert ck[i] == exp[12:16].view(np.uint32)[0], f"stored checksum {i}"
    if H.ref_lib() is not None:
        rsz, rty, rck = H.ref_float_info(arch)
        assert list(rsz) == sizes and list(rty) == [ft] * len(sizes) and list(rck.view(np.uint32)) == list(ck)


# ------------------------------------------------------------------------------------ size bounds vs the reference build
@needs_ref
def test_max_compressed_sizes_match_reference_build():
    L = H.ref_lib()
    ns = [0, 1, 15, 16, 17, 4095, 4096, 4097, 32768, 100000, 1 << 20, (1 << 20) + 1, 2 << 20, 64 << 20,
          (256 << 20) + 3, 1 << 30]
    rng = np.random.default_rng(4)
    ns += [int(x) for x in rng.integers(1, 1 << 30, 64)]
    lib = _capi.lib  # the raw C ABI: 0 = "no bound the 32-bit API can express" (codec.py raises on it)
    for n in ns:
        assert lib.dgb_ans_max_compressed_size(n) == L.ref_ans_max_compressed_size(n), f"ANS bound at {n}"
        for ft in (F16, BF16, F32, F64):
            ours = lib.dgb_float_max_compressed_size(ft, n)
            ours_sp = lib.dgb_sparse_float_max_compressed_size(ft, n)
            if ours == 0 or ours_sp == 0:
                with pytest.raises(ValueError):
                    codec.get_max_sparse_float_compressed_size(ft, n)
                # the bound does not fit 32 bits: the reference's uint32 sum wraps silently, we refuse (return 0)
                assert n * codec.WORD_BYTES[ft] > (1 << 31)
                continue
            assert ours == L.ref_float_max_compressed_size(ft, n), f"float bound ft {ft} at {n}"
            assert ours_sp == L.ref_sparse_float_max_compressed_size(ft, n), f"sparse bound ft {ft} at {n}"


# ------------------------------------------------------------------------------------ checksum folded into the data pass
@pytest.mark.parametrize("ft", [F16, BF16, F32, F64])
@pytest.mark.parametrize("big", [False, True], ids=["separate_launches", "fused_kernel"])
def test_float_checksum_rides_the_split_pass(ft, big):
    """useChecksum: the checksum (first N BYTES of the N-word input, dietgpu/float/GpuFloatCompress.cuh:708-710) is
    accumulated by the split pass itself and patched into the header by the CTA that completes the element; no
    k_checksum launch. Sizes put the end of the covered bytes inside a 16-byte vector, at a vector boundary, in the
    first tile and tiles later; a misaligned input takes the scalar path."""
    counts = [0, 1, 13, 16, 4099, 70001, (1 << 20) + 5, 32768 * 3]
    if big:
        counts += [(2 << 20) - 7 * i for i in range(16)]  # >= 1024 tiles and >= 16 elements: the fused kernel (not fp64)
    host = [datagen.float_words(ft, c, seed=600 + i) for i, c in enumerate(counts)]
    wb = host[0].dtype.itemsize
    for mis in (0, wb):
        dev = [H.to_dev(h, mis) for h in host]
        _kernels_run()  # clear
        m, sz = H.ours_float_compress(ft, dev, counts, 10, checksum=True)
        ran = _kernels_run()
        assert "k_checksum" not in ran, f"the separate checksum pass ran: {ran}"
        if big and ft != F64:
            assert "k_compressFused" in ran, ran
        for i, h in enumerate(host):
            exp = oracle.float_compress(ft, h, 10, use_checksum=True)
            got = H.row_bytes(m, i, sz[i])
            assert got.size == exp.size and np.array_equal(got, exp), f"archive {i} (misalign {mis}) differs from the oracle"
        outs, ok, dsz = H.ours_float_decompress(ft, [m[i] for i in range(len(counts))], counts, 10, checksum=True)
        assert ok.all() and list(dsz) == counts
        if H.ref_lib() is not None and mis == 0:
            rm, rsz = H.ref_float_compress(ft, dev, counts, 10, checksum=True)
            H.assert_rows_match_ref(m, sz, rm, rsz, oracle.float_defined_mask, f"checksum ft {ft}")


@pytest.mark.parametrize("big", [False, True], ids=["separate_launches", "fused_kernel"])
def test_ans_checksum_rides_the_histogram_pass(big):
    sizes = [0, 1, 15, 4097, 70001, (1 << 20) + 3]
    if big:
        sizes += [(2 << 20) - 5 * i for i in range(16)]
    host = [datagen.exp_bytes(n, 20.0, seed=700 + i) for i, n in enumerate(sizes)]
    for mis in (0, 3):
        dev = [H.to_dev(h, mis) for h in host]
        _kernels_run()
        m, sz = H.ours_ans_encode(dev, 10, checksum=True)
        ran = _kernels_run()
        assert "k_checksum" not in ran, f"the separate checksum pass ran: {ran}"
        for i, h in enumerate(host):
            exp = oracle.ans_encode(h, 10, use_checksum=True)
            assert np.array_equal(H.row_bytes(m, i, sz[i]), exp), f"ANS archive {i} (misalign {mis}) differs from the oracle"
        outs, ok, _ = H.ours_ans_decode([m[i] for i in range(len(sizes))], sizes, 10, checksum=True)
        assert ok.all()
    # a caller-supplied histogram leaves no pass over the data: the separate kernel computes the same checksum
    dev = [H.to_dev(h) for h in host]
    hist = torch.from_numpy(np.stack([np.bincount(h, minlength=256) for h in host]).astype(np.int32)).cuda()
    _kernels_run()
    m2, sz2 = H.ours_ans_encode(dev, 10, checksum=True, histogram=hist)
    assert "k_checksum" in _kernels_run()
    for i, h in enumerate(host):
        assert np.array_equal(H.row_bytes(m2, i, sz2[i]), oracle.ans_encode(h, 10, use_checksum=True, hist=np.bincount(h, minlength=256).astype(np.uint32)))
