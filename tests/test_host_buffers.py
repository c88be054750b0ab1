"""`-m gpu` tests of the host-buffer entry points (include/dietgpu_b200.h "host buffers", csrc/host_pipe.cu):
CPU tensors in, CPU tensors out, copies and kernels pipelined over element groups inside the library. The archives
must be the oracle's byte for byte (the same as the device-pointer calls'), whatever the grouping, with pinned and
with pageable memory, and the round trip bit exact."""
import numpy as np
import pytest
import torch

import datagen
import oracle
from dietgpu_fork_b200 import _capi, codec

pytestmark = pytest.mark.gpu

F16, BF16, F32, F64 = 1, 2, 3, 4
NP_WORD = {F16: np.uint16, BF16: np.uint16, F32: np.uint32, F64: np.uint64}


def _host(arr, pinned):
    t = torch.from_numpy(np.ascontiguousarray(arr).view(np.uint8).reshape(-1).copy())
    return t.pin_memory() if pinned else t


def _float_case(ft, sizes, pinned, pb=10, checksum=False):
    words = [datagen.float_words(ft, n, seed=300 + i, scale=0.02 if ft != F64 else 1.0) for i, n in enumerate(sizes)]
    ins = [_host(w, pinned) for w in words]
    outs = [_host(np.zeros(codec.get_max_float_compressed_size(ft, n), np.uint8), pinned) for n in sizes]
    csz = codec.float_compress_host(ft, ins, outs, pb, checksum, in_sizes=sizes)
    for i, w in enumerate(words):
        exp = oracle.float_compress(ft, w, pb, checksum)
        assert csz[i] == exp.size, f"archive {i}: size {csz[i]} != {exp.size}"
        assert np.array_equal(outs[i][:csz[i]].numpy(), exp), f"archive {i} differs from the oracle"
    wb = np.dtype(NP_WORD[ft]).itemsize
    dec = [_host(np.full(n * wb, 0xAB, np.uint8), pinned) for n in sizes]
    ok, dsz = codec.float_decompress_host(ft, outs, csz, dec, pb, checksum, out_capacity=sizes)
    assert all(ok) and dsz == list(sizes)
    for i, w in enumerate(words):
        assert np.array_equal(dec[i].numpy().view(NP_WORD[ft]), w), f"round trip {i}"


@pytest.mark.parametrize("pinned", [True, False], ids=["pinned", "pageable"])
@pytest.mark.parametrize("ft", [F16, BF16, F32, F64])
def test_float_host_small_ragged(ft, pinned):
    _float_case(ft, [0, 1, 33, 4096, 4097, 100000, 12345], pinned)


def test_float_host_many_groups_bf16():
    # 40 x 4 MiB: five 32 MiB groups over three lanes, the staging of every lane is reused
    _float_case(BF16, [2 * 1024 * 1024] * 40, True)


def test_float_host_uneven_groups_fp32_checksum():
    # elements larger and smaller than a group, checksum on (decode verifies on the device)
    _float_case(F32, [9 * 1024 * 1024 + 5, 100, 3 * 1024 * 1024, 17, 12 * 1024 * 1024 + 1, 4096 * 3], True,
                pb=11, checksum=True)


@pytest.mark.parametrize("pinned", [True, False], ids=["pinned", "pageable"])
def test_ans_host_matches_oracle(pinned):
    sizes = [0, 1, 4095, 4096, 70001, 1 << 20, (5 << 20) + 3] + [1 << 20] * 70
    data = [datagen.exp_bytes(n, 20.0, seed=500 + i) for i, n in enumerate(sizes)]
    ins = [_host(d, pinned) for d in data]
    outs = [_host(np.zeros(codec.get_max_compressed_size(n), np.uint8), pinned) for n in sizes]
    csz = codec.ans_encode_host(ins, outs, 10, in_sizes=sizes)
    for i, d in enumerate(data):
        exp = oracle.ans_encode(d, 10)
        assert csz[i] == exp.size and np.array_equal(outs[i][:csz[i]].numpy(), exp), f"ans archive {i}"
    dec = [_host(np.full(n, 0xCD, np.uint8), pinned) for n in sizes]
    ok, dsz = codec.ans_decode_host(outs, csz, dec, 10, out_capacity=sizes)
    assert all(ok) and dsz == sizes
    for i, d in enumerate(data):
        assert np.array_equal(dec[i].numpy(), d), f"ans round trip {i}"


def test_host_and_device_calls_give_identical_archives():
    n = 300000
    w = datagen.float_words(BF16, n, seed=9, scale=0.02)
    hin, hout = _host(w, True), _host(np.zeros(codec.get_max_float_compressed_size(BF16, n), np.uint8), True)
    (csz,) = codec.float_compress_host(BF16, [hin], [hout], in_sizes=[n])
    din = hin.cuda().view(torch.bfloat16)
    dout = torch.zeros(hout.numel(), dtype=torch.uint8, device="cuda")
    dsz = torch.zeros(1, dtype=torch.int32, device="cuda")
    codec.float_compress(BF16, [din], [dout], dsz)
    torch.cuda.synchronize()
    assert int(dsz.item()) == csz
    assert torch.equal(dout[:csz].cpu(), hout[:csz])


def test_host_compress_reports_short_output_buffer():
    n = 50000
    w = datagen.float_words(F32, n, seed=3)
    hin = _host(w, True)
    small = torch.zeros(1024, dtype=torch.uint8).pin_memory()
    with pytest.raises(_capi.DietGpuError):
        codec.float_compress_host(F32, [hin], [small], in_sizes=[n])


def test_host_decompress_capacity_too_small_reports_failure():
    n = 50000
    w = datagen.float_words(BF16, n, seed=4, scale=0.02)
    hin, hout = _host(w, True), _host(np.zeros(codec.get_max_float_compressed_size(BF16, n), np.uint8), True)
    (csz,) = codec.float_compress_host(BF16, [hin], [hout], in_sizes=[n])
    dec = torch.zeros(2 * 100, dtype=torch.uint8).pin_memory()
    ok, _ = codec.float_decompress_host(BF16, [hout], [csz], [dec], out_capacity=[100])
    assert ok == [0]
    codec.host_release()


def test_host_decompress_rejects_archives_longer_than_in_bytes():
    """in_bytes is the caller's statement of how much archive there is: headers that claim more (a truncated buffer,
    a corrupted word count) are refused on the host before anything is copied or decoded."""
    from dietgpu_fork_b200 import _capi
    n = 50000
    host = torch.from_numpy(datagen.float_words(2, n, seed=9).view(np.int16).copy()).view(torch.bfloat16)
    comp = torch.zeros(codec.get_max_float_compressed_size(2, n), dtype=torch.uint8)
    size = codec.float_compress_host(2, [host], [comp])[0]
    out = torch.zeros(n, dtype=torch.bfloat16)
    ok, _ = codec.float_decompress_host(2, [comp], [size], [out])  # exact size: fine
    assert ok == [1] and torch.equal(out.view(torch.int16), host.view(torch.int16))
    with pytest.raises(_capi.DietGpuError):
        codec.float_decompress_host(2, [comp], [size - 16], [out])  # truncated buffer
    bad = comp.clone()
    ans_off = 32 + (n + 15) // 16 * 16
    bad[ans_off + 12:ans_off + 16] = torch.tensor([0xFF, 0xFF, 0xFF, 0x0F], dtype=torch.uint8)  # totalCompressedWords
    with pytest.raises(_capi.DietGpuError):
        codec.float_decompress_host(2, [bad], [size], [out])

