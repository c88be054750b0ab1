"""`-m gpu`: archives whose header or block index disagree with themselves must be reported as failed, not decoded
out of bounds (the reference only asserts here, dietgpu/ans/GpuANSDecode.cuh:325-341; a library that replaces
CHECK-aborts with status codes has to bound what it reads from the archive)."""
import numpy as np
import pytest
import torch

import datagen
import gpu_harness as H

pytestmark = pytest.mark.gpu

BF16, F64 = 2, 4


def _ans_archive(n=3 * 4096 + 100):
    host = datagen.exp_bytes(n, 20.0, seed=77)
    m, sz = H.ours_ans_encode([H.to_dev(host)], 10)
    return host, m[0, :sz[0]].cpu().numpy().copy()


def _decode(arch, n):
    outs, ok, dsz = H.ours_ans_decode([H.to_dev(arch)], [n], 10)
    torch.cuda.synchronize()
    return outs[0].cpu().numpy(), ok, dsz


def test_intact_archive_decodes():
    host, arch = _ans_archive()
    out, ok, dsz = _decode(arch, host.size)
    assert ok.all() and np.array_equal(out, host)


@pytest.mark.parametrize("nb", [0, 3, 5, 1 << 20])
def test_wrong_block_count_is_rejected(nb):
    host, arch = _ans_archive()
    arch[4:8].view(np.uint32)[0] = nb  # numBlocks: the header implies 4
    out, ok, dsz = _decode(arch, host.size)
    assert not ok.any()
    assert not out.any(), "nothing may be written for a rejected archive"


def test_misaligned_or_out_of_range_block_start_is_rejected():
    host, arch = _ans_archive()
    nb = 4
    index = arch[544 + 128 * nb:544 + 128 * nb + 8 * nb].view(np.uint32).reshape(nb, 2)
    for bad in (index[1, 1] + 1, 1 << 30):  # not a multiple of 8 words / far past the payload
        a = arch.copy()
        a[544 + 128 * nb:544 + 128 * nb + 8 * nb].view(np.uint32).reshape(nb, 2)[1, 1] = bad
        out, ok, dsz = _decode(a, host.size)
        assert not ok.any()


def test_wrong_block_length_is_rejected():
    host, arch = _ans_archive()
    nb = 4
    a = arch.copy()
    idx = a[544 + 128 * nb:544 + 128 * nb + 8 * nb].view(np.uint32).reshape(nb, 2)
    idx[2, 0] = (0 << 16) | (idx[2, 0] & 0xFFFF)  # block 2 claims 0 symbols but keeps its words
    out, ok, dsz = _decode(a, host.size)
    assert not ok.any()
    a = arch.copy()
    idx = a[544 + 128 * nb:544 + 128 * nb + 8 * nb].view(np.uint32).reshape(nb, 2)
    idx[0, 0] = (4096 << 16) | 0xFFFF  # more words than a block can hold
    out, ok, dsz = _decode(a, host.size)
    assert not ok.any()


def test_fp64_second_plane_header_checked():
    n = 10000
    host = datagen.float_words(F64, n, seed=5)
    m, sz = H.ours_float_compress(F64, [H.to_dev(host)], [n], 10)
    arch = m[0, :sz[0]].cpu().numpy().copy()
    first = int(arch[16:20].view(np.uint32)[0])
    raw = 4 * ((n + 3) // 4 * 4) + 2 * ((n + 7) // 8 * 8)
    second = 32 + raw + first
    assert arch[second:second + 4].view(np.uint32)[0] == 0xD00D0001
    arch[second:second + 4].view(np.uint32)[0] = 0x12345678  # magic of the second exponent plane's archive
    outs, ok, dsz = H.ours_float_decompress(F64, [H.to_dev(arch)], [n], 10)
    assert not ok.any()
