"""Helpers for the `-m gpu` parity tests: run our library and (when oracle/_ref is built) the
reference's own CUDA build on the same device buffers, return host copies of the archives."""
import ctypes as C
import os

import numpy as np
import torch

import oracle
from dietgpu_fork_b200 import codec

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libdietgpu_ref.so")
WORD_DTYPE = oracle.WORD_DTYPE
TORCH_WORD = {1: torch.int16, 2: torch.int16, 3: torch.int32, 4: torch.int64}

_ref = None


def ref_lib():
    global _ref
    if _ref is None:
        if not os.path.exists(REF_PATH):
            return None
        L = C.CDLL(REF_PATH)
        L.ref_ans_max_compressed_size.restype = C.c_uint32
        L.ref_ans_max_compressed_size.argtypes = [C.c_uint32]
        L.ref_float_max_compressed_size.restype = C.c_uint32
        L.ref_float_max_compressed_size.argtypes = [C.c_int, C.c_uint32]
        L.ref_sparse_float_max_compressed_size.restype = C.c_uint32
        L.ref_sparse_float_max_compressed_size.argtypes = [C.c_int, C.c_uint32]
        _ref = L
    return _ref


def to_dev(arr, misalign=0):
    """Device copy of a numpy array whose data pointer is (16-byte aligned + misalign bytes)."""
    b = np.ascontiguousarray(arr).view(np.uint8).reshape(-1)
    buf = torch.zeros(b.size + 64, dtype=torch.uint8, device="cuda")
    base = (-buf.data_ptr()) % 16 + misalign
    view = buf[base:base + b.size]
    if b.size:
        view.copy_(torch.from_numpy(b.copy()))
    return view


def _ptr_array(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


def _u32(vals):
    return (C.c_uint32 * len(vals))(*vals)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


_TEMP = {}


def ref_temp(nbytes=512 << 20):
    if "t" not in _TEMP or _TEMP["t"].numel() < nbytes:
        _TEMP["t"] = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    return _TEMP["t"]


def alloc_outs(n, row_bytes, fill=0):
    rows = (row_bytes + 15) // 16 * 16
    m = torch.full((max(n, 1), rows), fill, dtype=torch.uint8, device="cuda")
    assert m.data_ptr() % 16 == 0
    return m


# ---------------------------------------------------------------- ours
def ours_ans_encode(dev_ins, pb=10, checksum=False, temp=None, fill=0, histogram=None):
    n = len(dev_ins)
    mx = max([codec.get_max_compressed_size(t.numel()) for t in dev_ins])
    m = alloc_outs(n, mx, fill)
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    codec.ans_encode_batch_pointer(dev_ins, [m[i] for i in range(n)], sizes, pb, checksum, histogram=histogram,
                                   temp=temp)
    torch.cuda.synchronize()
    sz = sizes.cpu().numpy().astype(np.int64)
    return m, sz


def ours_ans_decode(archives, caps, pb=10, checksum=False, misalign=0):
    n = len(archives)
    outs = [to_dev(np.zeros(c, np.uint8), misalign) for c in caps]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    codec.ans_decode_batch_pointer(archives, outs, ok, sizes, pb, checksum, out_capacity=list(caps))
    torch.cuda.synchronize()
    return outs, ok.cpu().numpy(), sizes.cpu().numpy()


def ours_float_compress(ft, dev_ins, counts, pb=10, checksum=False, sparse=False, fill=0):
    n = len(dev_ins)
    fmax = codec.get_max_sparse_float_compressed_size if sparse else codec.get_max_float_compressed_size
    mx = max([fmax(ft, c) for c in counts])
    m = alloc_outs(n, mx, fill)
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    outs = [m[i] for i in range(n)]
    if sparse:
        codec.float_compress_sparse(ft, dev_ins, outs, sizes, pb, in_sizes=list(counts))
    else:
        codec.float_compress(ft, dev_ins, outs, sizes, pb, checksum, in_sizes=list(counts))
    torch.cuda.synchronize()
    return m, sizes.cpu().numpy().astype(np.int64)


def ours_float_decompress(ft, archives, caps, pb=10, checksum=False, sparse=False, misalign=0):
    n = len(archives)
    wb = codec.WORD_BYTES[ft]
    outs = [to_dev(np.zeros(c * wb, np.uint8), misalign) for c in caps]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    if sparse:
        codec.float_decompress_sparse(ft, archives, outs, ok, sizes, pb, out_capacity=list(caps))
    else:
        codec.float_decompress(ft, archives, outs, ok, sizes, pb, checksum, out_capacity=list(caps))
    torch.cuda.synchronize()
    return outs, ok.cpu().numpy(), sizes.cpu().numpy()


# ---------------------------------------------------------------- reference CUDA build
def ref_ans_encode(dev_ins, pb=10, checksum=False, fill=0xCD, histogram=None):
    L = ref_lib()
    n = len(dev_ins)
    mx = max([codec.get_max_compressed_size(t.numel()) for t in dev_ins])
    m = alloc_outs(n, mx, fill)
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    t = ref_temp()
    L.ref_ans_encode_batch_pointer(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), pb, int(checksum), n,
                                   _ptr_array(dev_ins), _u32([x.numel() for x in dev_ins]),
                                   None if histogram is None else C.c_void_p(histogram.data_ptr()),
                                   _ptr_array([m[i] for i in range(n)]), C.c_void_p(sizes.data_ptr()),
                                   _stream())
    torch.cuda.synchronize()
    return m, sizes.cpu().numpy().astype(np.int64)


def ref_ans_decode(archives, caps, pb=10, checksum=False):
    L = ref_lib()
    n = len(archives)
    outs = [to_dev(np.zeros(c, np.uint8)) for c in caps]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    t = ref_temp()
    L.ref_ans_decode_batch_pointer(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), pb, int(checksum), n,
                                   _ptr_array(archives), _ptr_array(outs), _u32(list(caps)),
                                   C.c_void_p(ok.data_ptr()), C.c_void_p(sizes.data_ptr()), _stream())
    torch.cuda.synchronize()
    return outs, ok.cpu().numpy(), sizes.cpu().numpy()


def ref_float_compress(ft, dev_ins, counts, pb=10, checksum=False, sparse=False, fill=0xCD):
    L = ref_lib()
    n = len(dev_ins)
    fmax = codec.get_max_sparse_float_compressed_size if sparse else codec.get_max_float_compressed_size
    mx = max([fmax(ft, c) for c in counts])
    m = alloc_outs(n, mx, fill)
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    t = ref_temp()
    fn = L.ref_float_compress_sparse if sparse else L.ref_float_compress
    fn(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), ft, pb, int(checksum), n, _ptr_array(dev_ins),
       _u32(list(counts)), _ptr_array([m[i] for i in range(n)]), C.c_void_p(sizes.data_ptr()), _stream())
    torch.cuda.synchronize()
    return m, sizes.cpu().numpy().astype(np.int64)


def ref_float_decompress(ft, archives, caps, pb=10, checksum=False, sparse=False):
    L = ref_lib()
    n = len(archives)
    wb = codec.WORD_BYTES[ft]
    outs = [to_dev(np.zeros(c * wb, np.uint8)) for c in caps]
    ok = torch.zeros(n, dtype=torch.uint8, device="cuda")
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    t = ref_temp()
    fn = L.ref_float_decompress_sparse if sparse else L.ref_float_decompress
    fn(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), ft, pb, int(checksum), n, _ptr_array(archives),
       _ptr_array(outs), _u32(list(caps)), C.c_void_p(ok.data_ptr()), C.c_void_p(sizes.data_ptr()),
       _stream())
    torch.cuda.synchronize()
    return outs, ok.cpu().numpy(), sizes.cpu().numpy()


def row_bytes(m, i, size):
    return m[i, :size].cpu().numpy()


# ---------------------------------------------------------------- stride addressing / info (both libraries)
def ref_ans_encode_stride(in_dev, num, in_size, in_stride, out_dev, out_stride, sizes, pb=10, checksum=False):
    L = ref_lib()
    t = ref_temp()
    L.ref_ans_encode_batch_stride(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), pb, int(checksum), num,
                                  C.c_void_p(in_dev.data_ptr()), C.c_uint32(in_size), C.c_uint32(in_stride), None,
                                  C.c_void_p(out_dev.data_ptr()), C.c_uint32(out_stride),
                                  C.c_void_p(sizes.data_ptr()), _stream())
    torch.cuda.synchronize()


def ref_ans_decode_stride(in_dev, num, in_stride, out_dev, out_stride, cap, ok, sizes, pb=10, checksum=False):
    L = ref_lib()
    t = ref_temp()
    rc = L.ref_ans_decode_batch_stride(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), pb, int(checksum), num,
                                       C.c_void_p(in_dev.data_ptr()), C.c_uint32(in_stride),
                                       C.c_void_p(out_dev.data_ptr()), C.c_uint32(out_stride), C.c_uint32(cap),
                                       C.c_void_p(ok.data_ptr()), C.c_void_p(sizes.data_ptr()), _stream())
    torch.cuda.synchronize()
    return rc


def ref_ans_info(archives):
    L = ref_lib()
    n = len(archives)
    t = ref_temp()
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    cks = torch.zeros(n, dtype=torch.int32, device="cuda")
    L.ref_ans_get_compressed_info(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), _ptr_array(archives), n,
                                  C.c_void_p(sizes.data_ptr()), C.c_void_p(cks.data_ptr()), _stream())
    torch.cuda.synchronize()
    return sizes.cpu().numpy(), cks.cpu().numpy()


def ref_float_info(archives):
    L = ref_lib()
    n = len(archives)
    t = ref_temp()
    sizes = torch.zeros(n, dtype=torch.int32, device="cuda")
    types = torch.zeros(n, dtype=torch.int32, device="cuda")
    cks = torch.zeros(n, dtype=torch.int32, device="cuda")
    L.ref_float_get_compressed_info(C.c_void_p(t.data_ptr()), C.c_size_t(t.numel()), _ptr_array(archives), n,
                                    C.c_void_p(sizes.data_ptr()), C.c_void_p(types.data_ptr()),
                                    C.c_void_p(cks.data_ptr()), _stream())
    torch.cuda.synchronize()
    return sizes.cpu().numpy(), types.cpu().numpy(), cks.cpu().numpy()


def assert_rows_match_ref(m, sz, rm, rsz, mask_fn, what):
    """Every archive of `m` equals the reference build's under the defined-byte mask (bytes the reference never
    writes are 0 in ours and whatever the fill was in its buffer)."""
    assert list(sz) == list(rsz), f"{what}: compressed sizes differ from the reference build"
    for i in range(len(sz)):
        a, r = row_bytes(m, i, sz[i]), row_bytes(rm, i, sz[i])
        mask = mask_fn(a)
        bad = np.nonzero((a != r) & mask)[0]
        assert bad.size == 0, f"{what}: archive {i} differs from the reference build at {bad[:8]} ({bad.size} bytes)"
        assert not a[~mask].any(), f"{what}: archive {i} has non-zero bytes where the format defines none"
