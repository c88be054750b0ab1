"""Raw-device-pointer codec API over torch tensors (device memory and streams only).

Mirrors the reference's C++ entry points one to one (same names in snake_case, same argument
meaning; /root/reference/dietgpu/ans/GpuANSCodec.h, /root/reference/dietgpu/float/GpuFloatCodec.h):
host-side lists of tensors play the role of the host pointer arrays, `out_sizes` / `out_success`
are device tensors written asynchronously on the current CUDA stream. Everything runs in
libdietgpu_b200.so; there is no fallback.
"""
import ctypes as C

import torch

from . import _capi
from ._capi import BFLOAT16, FLOAT16, FLOAT32, FLOAT64, lib

FLOAT_TYPE_OF_DTYPE = {
    torch.float16: FLOAT16,
    torch.bfloat16: BFLOAT16,
    torch.float32: FLOAT32,
    torch.float64: FLOAT64,
}
WORD_BYTES = {FLOAT16: 2, BFLOAT16: 2, FLOAT32: 4, FLOAT64: 8}


class ChecksumMismatch(RuntimeError):
    def __init__(self, indices):
        super().__init__(f"checksum mismatch in batch elements {indices}")
        self.indices = indices


def _stream(stream):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _temp(temp):
    if temp is None:
        return None, 0
    assert temp.is_cuda and temp.is_contiguous()
    return C.c_void_p(temp.data_ptr()), temp.numel() * temp.element_size()


def _ptr(t):
    return None if t is None else C.c_void_p(t.data_ptr())


def _ptrs(ts):
    return _capi.ptr_array([t.data_ptr() for t in ts])


def _bound(v, what, size):
    # the library returns 0 when the bound does not fit the 32-bit API (the reference CHECK-aborts / wraps there)
    if v == 0 and size > 0:
        raise ValueError(f"{what}: {size} is too large for the 32-bit size API")
    return v


def get_max_compressed_size(uncompressed_bytes):
    if not 0 <= uncompressed_bytes <= 0xFFFFFFFF:
        raise ValueError(f"getMaxCompressedSize: {uncompressed_bytes} is out of range")
    return _bound(lib.dgb_ans_max_compressed_size(uncompressed_bytes), "getMaxCompressedSize", uncompressed_bytes)


def get_max_float_compressed_size(float_type, size):
    if not 0 <= size <= 0xFFFFFFFF:
        raise ValueError(f"getMaxFloatCompressedSize: {size} is out of range")
    return _bound(lib.dgb_float_max_compressed_size(float_type, size), "getMaxFloatCompressedSize", size)


def get_max_sparse_float_compressed_size(float_type, size):
    if not 0 <= size <= 0xFFFFFFFF:
        raise ValueError(f"getMaxSparseFloatCompressedSize: {size} is out of range")
    return _bound(lib.dgb_sparse_float_max_compressed_size(float_type, size), "getMaxSparseFloatCompressedSize", size)


def last_temp_bytes_used():
    return lib.dgb_last_temp_bytes_used()


def _raise_mismatch(mm, nm):
    raise ChecksumMismatch([mm[i] for i in range(min(nm.value, len(mm)))])


def ans_encode_batch_pointer(ins, outs, out_sizes, prob_bits=10, use_checksum=False, histogram=None,
                             temp=None, stream=None, in_sizes=None):
    """ansEncodeBatchPointer. `ins`: cuda tensors viewed as bytes; `in_sizes` defaults to their byte sizes."""
    n = len(ins)
    sizes = in_sizes if in_sizes is not None else [t.numel() * t.element_size() for t in ins]
    tp, tb = _temp(temp)
    rc = lib.dgb_ans_encode_batch_pointer(tp, tb, prob_bits, int(use_checksum), n, _ptrs(ins),
                                          _capi.u32_array(sizes), _ptr(histogram), _ptrs(outs),
                                          _ptr(out_sizes), _stream(stream))
    _capi.check(rc, "ansEncodeBatchPointer")


def ans_decode_batch_pointer(ins, outs, out_success=None, out_sizes=None, prob_bits=10, use_checksum=False,
                             temp=None, stream=None, out_capacity=None):
    n = len(ins)
    caps = out_capacity if out_capacity is not None else [t.numel() * t.element_size() for t in outs]
    tp, tb = _temp(temp)
    mm = (C.c_uint32 * max(n, 1))()
    nm = C.c_uint32(0)
    rc = lib.dgb_ans_decode_batch_pointer(tp, tb, prob_bits, int(use_checksum), n, _ptrs(ins), _ptrs(outs),
                                          _capi.u32_array(caps), _ptr(out_success), _ptr(out_sizes),
                                          _stream(stream), mm, n, C.byref(nm))
    if _capi.check(rc, "ansDecodeBatchPointer"):
        _raise_mismatch(mm, nm)


def ans_encode_batch_stride(in_dev, num, in_size, in_stride, out_dev, out_stride, out_sizes, prob_bits=10,
                            use_checksum=False, histogram=None, temp=None, stream=None):
    """ansEncodeBatchStride (dietgpu/ans/GpuANSCodec.h:68-98): `num` inputs of `in_size` bytes, `in_stride` bytes
    apart in the device tensor `in_dev`; archives `out_stride` bytes apart in `out_dev`."""
    tp, tb = _temp(temp)
    rc = lib.dgb_ans_encode_batch_stride(tp, tb, prob_bits, int(use_checksum), num, _ptr(in_dev), in_size,
                                         in_stride, _ptr(histogram), _ptr(out_dev), out_stride, _ptr(out_sizes),
                                         _stream(stream))
    _capi.check(rc, "ansEncodeBatchStride")


def ans_decode_batch_stride(in_dev, num, in_stride, out_dev, out_stride, out_capacity, out_success=None,
                            out_sizes=None, prob_bits=10, use_checksum=False, temp=None, stream=None):
    """ansDecodeBatchStride (dietgpu/ans/GpuANSCodec.h:222-257)."""
    tp, tb = _temp(temp)
    mm = (C.c_uint32 * max(num, 1))()
    nm = C.c_uint32(0)
    rc = lib.dgb_ans_decode_batch_stride(tp, tb, prob_bits, int(use_checksum), num, _ptr(in_dev), in_stride,
                                         _ptr(out_dev), out_stride, out_capacity, _ptr(out_success),
                                         _ptr(out_sizes), _stream(stream), mm, num, C.byref(nm))
    if _capi.check(rc, "ansDecodeBatchStride"):
        _raise_mismatch(mm, nm)


def float_compress(float_type, ins, outs, out_sizes, prob_bits=10, use_checksum=False, temp=None,
                   stream=None, in_sizes=None):
    """floatCompress. Sizes are in float words."""
    n = len(ins)
    sizes = in_sizes if in_sizes is not None else [t.numel() for t in ins]
    tp, tb = _temp(temp)
    rc = lib.dgb_float_compress(tp, tb, float_type, prob_bits, int(use_checksum), n, _ptrs(ins),
                                _capi.u32_array(sizes), _ptrs(outs), _ptr(out_sizes), _stream(stream))
    _capi.check(rc, "floatCompress")


def float_decompress(float_type, ins, outs, out_success=None, out_sizes=None, prob_bits=10,
                     use_checksum=False, temp=None, stream=None, out_capacity=None):
    n = len(ins)
    caps = out_capacity if out_capacity is not None else [t.numel() for t in outs]
    tp, tb = _temp(temp)
    mm = (C.c_uint32 * max(n, 1))()
    nm = C.c_uint32(0)
    rc = lib.dgb_float_decompress(tp, tb, float_type, prob_bits, int(use_checksum), n, _ptrs(ins),
                                  _ptrs(outs), _capi.u32_array(caps), _ptr(out_success), _ptr(out_sizes),
                                  _stream(stream), mm, n, C.byref(nm))
    if _capi.check(rc, "floatDecompress"):
        _raise_mismatch(mm, nm)


def float_compress_sparse(float_type, ins, outs, out_sizes, prob_bits=10, temp=None, stream=None,
                          in_sizes=None):
    n = len(ins)
    sizes = in_sizes if in_sizes is not None else [t.numel() for t in ins]
    tp, tb = _temp(temp)
    rc = lib.dgb_float_compress_sparse(tp, tb, float_type, prob_bits, 0, n, _ptrs(ins),
                                       _capi.u32_array(sizes), _ptrs(outs), _ptr(out_sizes),
                                       _stream(stream))
    _capi.check(rc, "floatCompressSparse")


def float_decompress_sparse(float_type, ins, outs, out_success=None, out_sizes=None, prob_bits=10,
                            temp=None, stream=None, out_capacity=None):
    n = len(ins)
    caps = out_capacity if out_capacity is not None else [t.numel() for t in outs]
    tp, tb = _temp(temp)
    rc = lib.dgb_float_decompress_sparse(tp, tb, float_type, prob_bits, 0, n, _ptrs(ins), _ptrs(outs),
                                         _capi.u32_array(caps), _ptr(out_success), _ptr(out_sizes),
                                         _stream(stream))
    _capi.check(rc, "floatDecompressSparse")


def ans_get_compressed_info(ins, out_sizes=None, out_checksum=None, stream=None):
    rc = lib.dgb_ans_get_compressed_info(_ptrs(ins), 0, len(ins), _ptr(out_sizes), _ptr(out_checksum),
                                         None, 0, _stream(stream))
    _capi.check(rc, "ansGetCompressedInfo")


def float_get_compressed_info(ins, out_sizes=None, out_types=None, out_checksum=None, stream=None):
    rc = lib.dgb_float_get_compressed_info(_ptrs(ins), 0, len(ins), _ptr(out_sizes), _ptr(out_types),
                                           _ptr(out_checksum), None, 0, _stream(stream))
    _capi.check(rc, "floatGetCompressedInfo")


# ---- host buffers (include/dietgpu_b200.h "host buffers", csrc/host_pipe.cu) ------------------------------
# CPU tensors in, CPU tensors out: the copies, the codec kernels and the size read-back are pipelined over
# groups of batch elements inside the library. Pinned tensors let the copies overlap; pageable ones work.
def _host_ptrs(ts):
    for t in ts:
        assert not t.is_cuda and t.is_contiguous(), "host-buffer calls take contiguous CPU tensors"
    return _capi.ptr_array([t.data_ptr() for t in ts])


def float_compress_host(float_type, ins, outs, prob_bits=10, use_checksum=False, stream=None, in_sizes=None):
    """Compresses the CPU tensors `ins` (float words) into the CPU byte tensors `outs`; returns the archive
    sizes in bytes (list of int). Archives are byte-identical to float_compress'."""
    n = len(ins)
    sizes = in_sizes if in_sizes is not None else [t.numel() for t in ins]
    caps = [t.numel() * t.element_size() for t in outs]
    out_sizes = (C.c_uint32 * max(n, 1))()
    rc = lib.dgb_float_compress_host(float_type, prob_bits, int(use_checksum), n, _host_ptrs(ins),
                                     _capi.u32_array(sizes), _host_ptrs(outs), _capi.u32_array(caps), out_sizes,
                                     _stream(stream))
    _capi.check(rc, "floatCompress (host buffers)")
    return list(out_sizes[:n])


def float_decompress_host(float_type, ins, in_bytes, outs, prob_bits=10, use_checksum=False, stream=None,
                          out_capacity=None):
    """Decompresses the CPU archives `ins` (`in_bytes[i]` bytes each) into the CPU tensors `outs`; returns
    (success flags, decoded sizes in words)."""
    n = len(ins)
    caps = out_capacity if out_capacity is not None else [t.numel() for t in outs]
    ok = (C.c_uint8 * max(n, 1))()
    out_sizes = (C.c_uint32 * max(n, 1))()
    rc = lib.dgb_float_decompress_host(float_type, prob_bits, int(use_checksum), n, _host_ptrs(ins),
                                       _capi.u32_array(in_bytes), _host_ptrs(outs), _capi.u32_array(caps), ok,
                                       out_sizes, _stream(stream))
    if _capi.check(rc, "floatDecompress (host buffers)"):
        raise ChecksumMismatch([])
    return list(ok[:n]), list(out_sizes[:n])


def ans_encode_host(ins, outs, prob_bits=10, use_checksum=False, stream=None, in_sizes=None):
    n = len(ins)
    sizes = in_sizes if in_sizes is not None else [t.numel() * t.element_size() for t in ins]
    caps = [t.numel() * t.element_size() for t in outs]
    out_sizes = (C.c_uint32 * max(n, 1))()
    rc = lib.dgb_ans_encode_host(prob_bits, int(use_checksum), n, _host_ptrs(ins), _capi.u32_array(sizes),
                                 _host_ptrs(outs), _capi.u32_array(caps), out_sizes, _stream(stream))
    _capi.check(rc, "ansEncodeBatchPointer (host buffers)")
    return list(out_sizes[:n])


def ans_decode_host(ins, in_bytes, outs, prob_bits=10, use_checksum=False, stream=None, out_capacity=None):
    n = len(ins)
    caps = out_capacity if out_capacity is not None else [t.numel() * t.element_size() for t in outs]
    ok = (C.c_uint8 * max(n, 1))()
    out_sizes = (C.c_uint32 * max(n, 1))()
    rc = lib.dgb_ans_decode_host(prob_bits, int(use_checksum), n, _host_ptrs(ins), _capi.u32_array(in_bytes),
                                 _host_ptrs(outs), _capi.u32_array(caps), ok, out_sizes, _stream(stream))
    if _capi.check(rc, "ansDecodeBatchPointer (host buffers)"):
        raise ChecksumMismatch([])
    return list(ok[:n]), list(out_sizes[:n])


def host_release():
    lib.dgb_host_release()
