"""`-m gpu` out-of-bounds check of our own (compute-sanitizer is closed on this pool): every buffer a call may write -
archives, decoded outputs, the caller's temp region, the size / success arrays - sits between two 4 KiB guard bands
filled with a pattern; after every call under both compress schedules the guards must be intact, bytes beyond the
reported archive size must still hold the fill pattern, and the library must not have needed more temp memory than
dgb_temp_bytes_required promised (i.e. it never fell back to the stream-ordered pool). Inputs get guards too: a kernel
that read past an input would not change them, but one that wrote through a wrong pointer would."""
import numpy as np
import pytest
import torch

import datagen
import oracle
from dietgpu_fork_b200 import _capi, codec

pytestmark = pytest.mark.gpu

GUARD = 4096
PAT = 0x5C
F16, BF16, F32, F64 = 1, 2, 3, 4
WB = {F16: 2, BF16: 2, F32: 4, F64: 8}


class Guarded:
    """A device byte buffer of `nbytes` (16-byte aligned start) between two guard bands."""

    def __init__(self, nbytes, fill=PAT):
        self.n = int(nbytes)
        body = (self.n + 15) // 16 * 16
        self.buf = torch.full((GUARD + body + GUARD + 16,), PAT, dtype=torch.uint8, device="cuda")
        self.off = GUARD + (-(self.buf.data_ptr() + GUARD)) % 16
        self.view = self.buf[self.off:self.off + self.n]
        self.view.fill_(fill)

    def guards_intact(self):
        lo = self.buf[:self.off]
        hi = self.buf[self.off + self.n:]
        return bool((lo == PAT).all()) and bool((hi == PAT).all())


def _put(host):
    b = np.ascontiguousarray(host).view(np.uint8).reshape(-1)
    g = Guarded(max(b.size, 1), fill=0)
    if b.size:
        g.view[:b.size].copy_(torch.from_numpy(b.copy()))
    return g


@pytest.fixture(params=["fused", "separate"])
def schedule(request):
    _capi.lib.dgb_set_fused_min_tiles(0 if request.param == "fused" else 1 << 40)
    yield request.param
    _capi.lib.dgb_set_fused_min_tiles(-1)


def _check(gs, what):
    torch.cuda.synchronize()
    for name, g in gs:
        assert g.guards_intact(), f"{what}: guard band of {name} was overwritten"


@pytest.mark.parametrize("ft", [F16, BF16, F32, F64])
def test_float_codec_stays_inside_its_buffers(ft, schedule):
    counts = [0, 1, 4095, 4097, 32769, 100003, 262144]
    host = [datagen.float_words(ft, c, seed=80 + i) for i, c in enumerate(counts)]
    ins = [_put(h) for h in host]
    mx = max(codec.get_max_float_compressed_size(ft, c) for c in counts)
    outs = [Guarded(mx) for _ in counts]
    sizes = Guarded(4 * len(counts), fill=0)
    temp = Guarded(_capi.lib.dgb_temp_bytes_required(2, ft, len(counts), max(counts)))
    codec.float_compress(ft, [g.view for g in ins], [g.view for g in outs], sizes.view.view(torch.int32), 10, False,
                         temp=temp.view, in_sizes=counts)
    _check([("input", g) for g in ins] + [("archive", g) for g in outs] + [("sizes", sizes), ("temp", temp)],
           f"floatCompress ft {ft}")
    assert codec.last_temp_bytes_used() <= temp.n, "compress needed more temp than dgb_temp_bytes_required promised"
    sz = sizes.view.view(torch.int32).cpu().numpy()
    for i, h in enumerate(host):
        exp = oracle.float_compress(ft, h, 10)
        got = outs[i].view.cpu().numpy()
        assert sz[i] == exp.size and np.array_equal(got[:sz[i]], exp), f"archive {i}"
        assert (got[sz[i]:] == PAT).all(), f"archive {i}: bytes beyond the reported size were written"
    dec = [Guarded(max(c * WB[ft], 1)) for c in counts]
    ok = Guarded(len(counts), fill=0)
    dsz = Guarded(4 * len(counts), fill=0)
    dtemp = Guarded(_capi.lib.dgb_temp_bytes_required(3, ft, len(counts), max(counts)))
    codec.float_decompress(ft, [g.view for g in outs], [g.view for g in dec], ok.view, dsz.view.view(torch.int32), 10,
                           False, temp=dtemp.view, out_capacity=counts)
    _check([("archive", g) for g in outs] + [("output", g) for g in dec] + [("ok", ok), ("sizes", dsz), ("temp", dtemp)],
           f"floatDecompress ft {ft}")
    assert codec.last_temp_bytes_used() <= dtemp.n
    assert bool(ok.view.all())
    for i, h in enumerate(host):
        assert np.array_equal(dec[i].view.cpu().numpy()[:counts[i] * WB[ft]], h.view(np.uint8)), f"decode {i}"


def test_ans_codec_stays_inside_its_buffers(schedule):
    counts = [0, 1, 4096, 4097, 70001, 1 << 20]
    host = [datagen.exp_bytes(c, 20.0, seed=90 + i) for i, c in enumerate(counts)]
    ins = [_put(h) for h in host]
    mx = max(codec.get_max_compressed_size(c) for c in counts)
    outs = [Guarded(mx) for _ in counts]
    sizes = Guarded(4 * len(counts), fill=0)
    temp = Guarded(_capi.lib.dgb_temp_bytes_required(0, 0, len(counts), max(counts)))
    codec.ans_encode_batch_pointer([g.view for g in ins], [g.view for g in outs], sizes.view.view(torch.int32), 10,
                                   True, temp=temp.view, in_sizes=counts)
    _check([("input", g) for g in ins] + [("archive", g) for g in outs] + [("sizes", sizes), ("temp", temp)], "ansEncode")
    assert codec.last_temp_bytes_used() <= temp.n
    sz = sizes.view.view(torch.int32).cpu().numpy()
    for i, h in enumerate(host):
        got = outs[i].view.cpu().numpy()
        assert np.array_equal(got[:sz[i]], oracle.ans_encode(h, 10, use_checksum=True)), f"archive {i}"
        assert (got[sz[i]:] == PAT).all()
    dec = [Guarded(max(c, 1)) for c in counts]
    ok = Guarded(len(counts), fill=0)
    dsz = Guarded(4 * len(counts), fill=0)
    dtemp = Guarded(_capi.lib.dgb_temp_bytes_required(1, 0, len(counts), max(counts)))
    codec.ans_decode_batch_pointer([g.view for g in outs], [g.view for g in dec], ok.view, dsz.view.view(torch.int32),
                                   10, True, temp=dtemp.view, out_capacity=counts)
    _check([("output", g) for g in dec] + [("ok", ok), ("sizes", dsz), ("temp", dtemp)], "ansDecode")
    assert codec.last_temp_bytes_used() <= dtemp.n
    for i, h in enumerate(host):
        assert np.array_equal(dec[i].view.cpu().numpy()[:counts[i]], h)


@pytest.mark.parametrize("ft", [BF16, F32])
def test_sparse_codec_stays_inside_its_buffers(ft):
    counts = [2, 8191, 8193, 300001]
    host = [datagen.float_words(ft, c, seed=70 + i, sparsity=0.9) for i, c in enumerate(counts)]
    ins = [_put(h) for h in host]
    mx = max(codec.get_max_sparse_float_compressed_size(ft, c) for c in counts)
    outs = [Guarded(mx) for _ in counts]
    sizes = Guarded(4 * len(counts), fill=0)
    temp = Guarded(_capi.lib.dgb_temp_bytes_required(4, ft, len(counts), max(counts)))
    codec.float_compress_sparse(ft, [g.view for g in ins], [g.view for g in outs], sizes.view.view(torch.int32), 10,
                                temp=temp.view, in_sizes=counts)
    _check([("input", g) for g in ins] + [("archive", g) for g in outs] + [("sizes", sizes), ("temp", temp)],
           "floatCompressSparse")
    assert codec.last_temp_bytes_used() <= temp.n
    sz = sizes.view.view(torch.int32).cpu().numpy()
    for i, h in enumerate(host):
        got = outs[i].view.cpu().numpy()
        assert np.array_equal(got[:sz[i]], oracle.sparse_float_compress(ft, h, 10)), f"sparse archive {i}"
        assert (got[sz[i]:] == PAT).all()
    dec = [Guarded(c * WB[ft]) for c in counts]
    dtemp = Guarded(_capi.lib.dgb_temp_bytes_required(5, ft, len(counts), max(counts)))
    codec.float_decompress_sparse(ft, [g.view for g in outs], [g.view for g in dec], None, None, 10, temp=dtemp.view,
                                  out_capacity=counts)
    _check([("output", g) for g in dec] + [("temp", dtemp)], "floatDecompressSparse")
    assert codec.last_temp_bytes_used() <= dtemp.n
    for i, h in enumerate(host):
        assert np.array_equal(dec[i].view.cpu().numpy(), h.view(np.uint8))
