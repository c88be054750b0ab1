"""The reference's own Python tests (/root/reference/dietgpu/ans_test.py:13-139 and
/root/reference/dietgpu/float_test.py:13-178), re-stated against `torch.ops.dietgpu.*` as registered by
dietgpu_fork_b200.torch_ops.install(): same call sequences, same assertions, seeded inputs."""
import random

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ops():
    from dietgpu_fork_b200 import torch_ops
    return torch_ops.install()


@pytest.fixture(scope="module")
def dev():
    return torch.device("cuda:0")


@pytest.fixture(scope="module")
def temp_mem(dev):
    return torch.empty([64 * 1024 * 1024], dtype=torch.uint8, device=dev)


def _round_trip(ops, dev, as_float, ts, checksum, temp_mem):
    comp, sizes, _ = ops.compress_data(as_float, ts, checksum, temp_mem)
    # truncate to exactly the reported sizes: the sizes must be accurate
    truncated = [t.narrow(0, 0, s.item()).clone() for s, t in zip(sizes, [*comp])]
    outs = [torch.empty(t.size(), dtype=t.dtype, device=t.device) for t in ts]
    if temp_mem is not None:
        status = torch.empty([len(ts)], dtype=torch.uint8, device=dev)
        osz = torch.empty([len(ts)], dtype=torch.int32, device=dev)
        ops.decompress_data(as_float, truncated, outs, checksum, temp_mem, status, osz)
        for t, st, sz in zip(ts, status, osz):
            assert st.item()
            assert sz.item() == (t.numel() if as_float else t.numel() * t.element_size())
    else:
        ops.decompress_data(as_float, truncated, outs, checksum)
    for a, b in zip(ts, outs):
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
    return sizes


# ---- ans_test.py -----------------------------------------------------------------------------------
@pytest.mark.parametrize("use_temp", [False, True])
@pytest.mark.parametrize("checksum", [False, True])
def test_ans_codec(ops, dev, temp_mem, use_temp, checksum):
    torch.manual_seed(3)
    ts = [torch.normal(0, 1.0, [n], dtype=torch.float32, device=dev) for n in (10000, 100000, 1000000)]
    _round_trip(ops, dev, False, ts, checksum, temp_mem if use_temp else None)


def test_ans_empty(ops, dev):
    ts = [torch.empty([0], dtype=torch.uint8, device=dev)]
    comp = ops.compress_data_simple(False, ts, True)
    assert comp[0].numel() > 0  # should have a header
    out = ops.decompress_data_simple(False, comp, True)
    assert torch.equal(ts[0], out[0])


def _random_sizes(rng, mult_of_4):
    sizes = []
    for _ in range(rng.randrange(1, 15)):
        s = rng.randrange(1, 10000)
        if mult_of_4:
            s += 4 - (s % 4)
        sizes.append(s)
    return sizes


def test_ans_split_compress(ops, dev, temp_mem):
    rng = random.Random(11)
    torch.manual_seed(11)
    for _ in range(5):
        sizes = _random_sizes(rng, True)
        t = torch.randint(0, 65, [sum(sizes)], dtype=torch.uint8, device=dev)
        comp, _, _ = ops.compress_data_split_size(False, t, torch.IntTensor(sizes), True, temp_mem)
        out = ops.decompress_data_simple(False, comp, True)
        for a, b in zip(torch.split(t, sizes), out):
            assert torch.equal(a, b)


def test_ans_split_decompress(ops, dev, temp_mem):
    rng = random.Random(12)
    torch.manual_seed(12)
    for _ in range(5):
        sizes = _random_sizes(rng, True)
        t = torch.randint(0, 65, [sum(sizes)], dtype=torch.uint8, device=dev)
        comp = ops.compress_data_simple(False, torch.split(t, sizes), True)
        out = torch.empty([sum(sizes)], dtype=torch.uint8, device=dev)
        ops.decompress_data_split_size(False, comp, out, torch.IntTensor(sizes), True, temp_mem)
        assert torch.equal(t, out)


# ---- float_test.py ---------------------------------------------------------------------------------
FLOAT_DTYPES = [torch.bfloat16, torch.float16, torch.float32]


@pytest.mark.parametrize("dt", FLOAT_DTYPES + [torch.float64])
@pytest.mark.parametrize("use_temp", [False, True])
def test_float_codec(ops, dev, temp_mem, dt, use_temp):
    torch.manual_seed(5)
    ts = [torch.normal(0, 1.0, [n], dtype=dt, device=dev) for n in (10000, 100000, 1000000)]
    sizes = _round_trip(ops, dev, True, ts, True, temp_mem if use_temp else None)
    for s, t in zip(sizes.tolist(), ts):
        assert s < t.numel() * t.element_size()


@pytest.mark.parametrize("dt", FLOAT_DTYPES)
@pytest.mark.parametrize("use_temp", [False, True])
def test_float_large(ops, dev, temp_mem, dt, use_temp):
    torch.manual_seed(6)
    ts = [torch.normal(0, 1.0, [123456789], dtype=dt, device=dev)]
    _round_trip(ops, dev, True, ts, True, temp_mem if use_temp else None)


@pytest.mark.parametrize("dt", FLOAT_DTYPES)
def test_float_simple(ops, dev, dt):
    torch.manual_seed(7)
    ts = [torch.normal(0, 1.0, [n], dtype=dt, device=dev) for n in (10000, 100000, 1000000)]
    cts = ops.compress_data_simple(True, ts, True)
    for before, after in zip(ts, cts):
        assert before.numel() * before.element_size() > after.numel() * after.element_size()
    dts = ops.decompress_data_simple(True, cts, True)
    for a, b in zip(ts, dts):
        assert torch.equal(a, b)


@pytest.mark.parametrize("dt", FLOAT_DTYPES)
def test_float_empty(ops, dev, dt):
    ts = [torch.empty([0], dtype=dt, device=dev)]
    comp = ops.compress_data_simple(True, ts, True)
    assert comp[0].numel() > 0
    out = ops.decompress_data_simple(True, comp, True)
    assert torch.equal(ts[0], out[0])


def _float_sizes(rng, align16):
    sizes = []
    for _ in range(rng.randrange(1, 15)):
        s = rng.randrange(1, 10000)
        sizes.append(s * 8 if align16 else s)
    return sizes


@pytest.mark.parametrize("dt", FLOAT_DTYPES)
@pytest.mark.parametrize("align16", [True, False])
def test_float_split_compress(ops, dev, temp_mem, dt, align16):
    rng = random.Random(21)
    torch.manual_seed(21)
    for _ in range(5):
        sizes = _float_sizes(rng, align16)
        t = torch.normal(0, 1.0, [sum(sizes)], dtype=dt, device=dev)
        comp, _, _ = ops.compress_data_split_size(True, t, torch.IntTensor(sizes), True, temp_mem)
        out = ops.decompress_data_simple(True, comp, True)
        for a, b in zip(torch.split(t, sizes), out):
            assert torch.equal(a, b)


@pytest.mark.parametrize("dt", FLOAT_DTYPES)
@pytest.mark.parametrize("align16", [True, False])
def test_float_split_decompress(ops, dev, temp_mem, dt, align16):
    rng = random.Random(22)
    torch.manual_seed(22)
    for _ in range(5):
        sizes = _float_sizes(rng, align16)
        t = torch.normal(0, 1.0, [sum(sizes)], dtype=dt, device=dev)
        comp = ops.compress_data_simple(True, torch.split(t, sizes), True)
        out = torch.empty([sum(sizes)], dtype=dt, device=dev)
        ops.decompress_data_split_size(True, comp, out, torch.IntTensor(sizes), True, temp_mem)
        assert torch.equal(t, out)


# ---- behaviour the reference enforces with TORCH_CHECK (DietGpu.cpp:161-320, :536-700) -------------
def test_checksum_mismatch_raises(ops, dev):
    torch.manual_seed(8)
    ts = [torch.normal(0, 1.0, [50000], dtype=torch.bfloat16, device=dev)]
    comp = ops.compress_data_simple(True, ts, True)
    bad = comp[0].clone()
    bad[12] ^= 0x5A  # GpuFloatHeader checksum word (dietgpu/float/GpuFloatUtils.cuh:19-60)
    with pytest.raises(RuntimeError, match="checksum mismatch"):
        ops.decompress_data(True, [bad], [torch.empty_like(ts[0])], True)


def test_mixed_dtype_rejected(ops, dev):
    ts = [torch.zeros(16, dtype=torch.float16, device=dev), torch.zeros(16, dtype=torch.bfloat16, device=dev)]
    with pytest.raises(RuntimeError):
        ops.compress_data(True, ts, False)


def test_capacity_too_small_reports_failure(ops, dev, temp_mem):
    torch.manual_seed(9)
    ts = [torch.normal(0, 1.0, [4096], dtype=torch.float16, device=dev)]
    comp = ops.compress_data_simple(True, ts, False)
    small = [torch.empty([1000], dtype=torch.float16, device=dev)]
    status = torch.ones([1], dtype=torch.uint8, device=dev)
    osz = torch.zeros([1], dtype=torch.int32, device=dev)
    ops.decompress_data(True, comp, small, False, temp_mem, status, osz)
    assert status.item() == 0 and osz.item() == 4096  # size reported even on failure (GpuFloatDecompress.cuh)


def test_sparse_ops(ops, dev, temp_mem):
    torch.manual_seed(10)
    for dt in FLOAT_DTYPES + [torch.float64]:
        ts = []
        for n in (2, 4097, 300000):
            t = torch.normal(0, 1.0, [n], dtype=dt, device=dev)
            t[torch.rand(n, device=dev) < 0.8] = 0
            ts.append(t)
        comp, sizes, _ = ops.compress_data_sparse(ts, temp_mem)
        outs = [torch.empty_like(t) for t in ts]
        status = torch.zeros([len(ts)], dtype=torch.uint8, device=dev)
        ops.decompress_data_sparse([comp[i, :sizes[i].item()].clone() for i in range(len(ts))], outs, temp_mem, status)
        assert status.cpu().all()
        for a, b in zip(ts, outs):
            assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
