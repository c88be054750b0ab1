"""The reference's `torch.ops.dietgpu.*` operators, same names / arguments / return values
(/root/reference/dietgpu/DietGpu.cpp:921-943 schemas; semantics :117-917), routed through
libdietgpu_b200.so. `install()` additionally exposes them as `torch.ops.dietgpu.<name>` when the
C++ registration (csrc/torch_binding.cpp) has not been loaded, so the reference's Python tests
(ans_test.py, float_test.py) run against this library by changing only the library load line.

Additions the reference's Python surface lacks (its C++ API has them): float64 decompression and
the sparse ops.
"""
import torch

from . import codec
from .codec import FLOAT_TYPE_OF_DTYPE

K_DEFAULT_PRECISION = 10  # DietGpu.cpp:119
_DTYPE_OF_FLOAT_TYPE = {v: k for k, v in FLOAT_TYPE_OF_DTYPE.items()}


def _check(cond, msg="check failed"):
    if not cond:
        raise RuntimeError(msg)


def _float_type(t):
    ft = FLOAT_TYPE_OF_DTYPE.get(t.dtype)
    _check(ft is not None, f"unsupported float dtype {t.dtype}")
    return ft


def _total_and_max(ts):
    total, mx = 0, 0
    for t in ts:
        n = t.numel()
        _check(n * t.element_size() <= 0xFFFFFFFF)
        total += n
        mx = max(mx, n)
    return total, mx


# ---- sizes (DietGpu.cpp:122-153) -----------------------------------------------------------
def max_float_compressed_output_size(ts):
    _, mx = _total_and_max(ts)
    return len(ts), codec.get_max_float_compressed_size(_float_type(ts[0]), mx)


def max_float_compressed_size(dtype_tensor, size):
    return codec.get_max_float_compressed_size(_float_type(dtype_tensor), size)


def max_any_compressed_output_size(ts):
    _, mx = _total_and_max(ts)
    return len(ts), codec.get_max_compressed_size(mx * ts[0].element_size())


def max_any_compressed_size(nbytes):
    return codec.get_max_compressed_size(nbytes)


# ---- compress (DietGpu.cpp:161-320) --------------------------------------------------------
def _validate_out(out_compressed, out_sizes, n, max_bytes, dev):
    if out_compressed is not None:
        _check(out_compressed.dtype == torch.uint8 and out_compressed.is_cuda)
        _check(out_compressed.is_contiguous() and out_compressed.dim() == 2)
        _check(out_compressed.size(0) >= n and out_compressed.size(1) >= max_bytes)
        _check(out_compressed.device == dev)
        comp = out_compressed
    else:
        comp = torch.empty((n, max_bytes), dtype=torch.uint8, device=dev)
    if out_sizes is not None:
        _check(out_sizes.dtype == torch.int32 and out_sizes.is_cuda and out_sizes.dim() == 1)
        _check(out_sizes.is_contiguous() and out_sizes.size(0) >= n and out_sizes.device == dev)
        sizes = out_sizes
    else:
        sizes = torch.empty((n,), dtype=torch.int32, device=dev)
    return comp, sizes


def compress_data(compress_as_float, ts_in, checksum=False, temp_mem=None, out_compressed=None,
                  out_compressed_bytes=None):
    """-> (compressed matrix [B, maxBytes] u8, sizes [B] i32 (device), temp bytes used)"""
    _check(len(ts_in) > 0)
    dev = ts_in[0].device
    for t in ts_in:
        _check(t.is_cuda and t.is_contiguous() and t.device == dev)
        if compress_as_float:
            _check(t.dtype == ts_in[0].dtype)
            _float_type(t)
    _, max_bytes = (max_float_compressed_output_size(ts_in) if compress_as_float
                    else max_any_compressed_output_size(ts_in))
    comp, sizes = _validate_out(out_compressed, out_compressed_bytes, len(ts_in), max_bytes, dev)
    rows = [comp[i] for i in range(len(ts_in))]
    with torch.cuda.device(dev):
        if compress_as_float:
            codec.float_compress(_float_type(ts_in[0]), ts_in, rows, sizes, K_DEFAULT_PRECISION, checksum,
                                 temp=temp_mem)
        else:
            codec.ans_encode_batch_pointer(ts_in, rows, sizes, K_DEFAULT_PRECISION, checksum, temp=temp_mem)
    return comp, sizes, codec.last_temp_bytes_used()


def _matrix_to_tensors(n, matrix, sizes_dev):
    sizes_host = sizes_dev[:n].cpu().tolist()  # DietGpu.cpp:85-111
    flat = matrix.view(-1)
    cols = matrix.size(1)
    return [flat.narrow(0, i * cols, sizes_host[i]) for i in range(n)]


def compress_data_split_size(compress_as_float, t_in, t_split_sizes, checksum=False, temp_mem=None,
                             out_compressed=None, out_compressed_bytes=None):
    """-> (list of views into the compressed matrix, sizes [B] i32 (device), temp bytes used)"""
    dev = t_in.device
    _check(t_in.is_cuda and t_in.is_contiguous())
    _check(t_split_sizes.is_contiguous() and t_split_sizes.device.type == "cpu")
    _check(t_split_sizes.dtype == torch.int32)
    splits = t_split_sizes.tolist()
    n = len(splits)
    if compress_as_float:
        ft = _float_type(t_in)
    else:
        _check(t_in.data_ptr() % 4 == 0,
               "All splits should start on a 16 byte boundary; start pointer is not aligned")
    for i, s in enumerate(splits):
        _check(s > 0)
        if not compress_as_float and i != n - 1:
            _check(s % 4 == 0, "All splits should start on a 16 byte boundary; the size of an interior "
                               "split is not a multiple of 16 bytes")
    mx = max(splits)
    max_bytes = (codec.get_max_float_compressed_size(ft, mx) if compress_as_float
                 else codec.get_max_compressed_size(mx))
    comp, sizes = _validate_out(out_compressed, out_compressed_bytes, n, max_bytes, dev)
    with torch.cuda.device(dev):
        esz = t_in.element_size() if compress_as_float else 1
        base = t_in.view(-1).view(torch.uint8)
        ins, off = [], 0
        for s in splits:
            ins.append(base.narrow(0, off * esz, s * esz))
            off += s
        rows = [comp.view(-1).narrow(0, i * comp.size(1), max_bytes) for i in range(n)]
        if compress_as_float:
            codec.float_compress(ft, ins, rows, sizes, K_DEFAULT_PRECISION, checksum, temp=temp_mem,
                                 in_sizes=splits)
        else:
            codec.ans_encode_batch_pointer(ins, rows, sizes, K_DEFAULT_PRECISION, checksum, temp=temp_mem,
                                           in_sizes=splits)
    return _matrix_to_tensors(n, comp, sizes), sizes, codec.last_temp_bytes_used()


def compress_data_simple(compress_as_float, ts_in, checksum=False, temp_mem=67108864):
    """-> list of exactly-sized compressed tensors (blocks on the sizes read-back, as the reference)."""
    _check(len(ts_in) > 0)
    scratch = None
    if temp_mem and temp_mem > 0:
        scratch = torch.empty((temp_mem,), dtype=torch.uint8, device=ts_in[0].device)
    comp, sizes, _ = compress_data(compress_as_float, ts_in, checksum, scratch)
    sizes_host = sizes.cpu().tolist()
    return [comp[i, :sizes_host[i]].clone() for i in range(len(ts_in))]


# ---- decompress (DietGpu.cpp:536-917) -----------------------------------------------------------
_CHECKSUM_MSG = {True: "floatDecompress: checksum mismatch seen on decoded data; archive cannot be unpacked",
                 False: "ANSDecode: checksum mismatch seen on decoded data; archive cannot be unpacked"}


def _validate_status(out_status, out_sizes, n, dev):
    if out_status is not None:
        _check(out_status.is_contiguous() and out_status.is_cuda and out_status.dtype == torch.uint8)
        _check(out_status.numel() == n and out_status.device == dev)
    if out_sizes is not None:
        _check(out_sizes.is_contiguous() and out_sizes.is_cuda and out_sizes.dtype == torch.int32)
        _check(out_sizes.numel() == n and out_sizes.device == dev)


def decompress_data(compress_as_float, ts_in, ts_out, checksum=False, temp_mem=None, out_status=None,
                    out_decompressed_words=None):
    """-> temp bytes used. Raises RuntimeError on a checksum mismatch (TORCH_CHECK in the reference)."""
    _check(len(ts_in) > 0 and len(ts_in) == len(ts_out))
    dev = ts_in[0].device
    for ti, to in zip(ts_in, ts_out):
        _check(ti.is_cuda and ti.device == dev and ti.is_contiguous() and ti.dtype == torch.uint8)
        _check(to.is_cuda and to.device == dev and to.is_contiguous())
        if compress_as_float:
            _float_type(to)  # float64 accepted here (the reference's binding rejects it, :569-573)
    _validate_status(out_status, out_decompressed_words, len(ts_in), dev)
    with torch.cuda.device(dev):
        try:
            if compress_as_float:
                codec.float_decompress(_float_type(ts_out[0]), ts_in, ts_out, out_status,
                                       out_decompressed_words, K_DEFAULT_PRECISION, checksum, temp=temp_mem)
            else:
                codec.ans_decode_batch_pointer(ts_in, ts_out, out_status, out_decompressed_words,
                                               K_DEFAULT_PRECISION, checksum, temp=temp_mem)
        except codec.ChecksumMismatch:
            raise RuntimeError(_CHECKSUM_MSG[bool(compress_as_float)])
    return codec.last_temp_bytes_used()


def decompress_data_split_size(compress_as_float, ts_in, t_out, t_split_sizes, checksum=False, temp_mem=None,
                               out_status=None, out_decompressed_words=None):
    _check(len(ts_in) > 0)
    dev = ts_in[0].device
    _check(t_split_sizes.is_contiguous() and t_split_sizes.device.type == "cpu")
    _check(t_split_sizes.dtype == torch.int32 and t_split_sizes.numel() == len(ts_in))
    splits = t_split_sizes.tolist()
    for ti, s in zip(ts_in, splits):
        _check(ti.is_cuda and ti.device == dev and ti.is_contiguous() and ti.dtype == torch.uint8)
        _check(s > 0)
    _check(t_out.is_cuda and t_out.device == dev and t_out.is_contiguous())
    if compress_as_float:
        ft = _float_type(t_out)
    _validate_status(out_status, out_decompressed_words, len(ts_in), dev)
    esz = t_out.element_size() if compress_as_float else 1
    base = t_out.view(-1).view(torch.uint8)
    outs, off = [], 0
    for s in splits:
        outs.append(base.narrow(0, off * esz, s * esz))
        off += s
    with torch.cuda.device(dev):
        try:
            if compress_as_float:
                codec.float_decompress(ft, ts_in, outs, out_status, out_decompressed_words,
                                       K_DEFAULT_PRECISION, checksum, temp=temp_mem, out_capacity=splits)
            else:
                codec.ans_decode_batch_pointer(ts_in, outs, out_status, out_decompressed_words,
                                               K_DEFAULT_PRECISION, checksum, temp=temp_mem,
                                               out_capacity=splits)
        except codec.ChecksumMismatch:
            raise RuntimeError(_CHECKSUM_MSG[bool(compress_as_float)])
    return codec.last_temp_bytes_used()


def decompress_data_simple(compress_as_float, ts_in, checksum=False, temp_mem=67108864):
    _check(len(ts_in) > 0)
    dev = ts_in[0].device
    n = len(ts_in)
    scratch = None
    if temp_mem and temp_mem >= 256:
        scratch = torch.empty((temp_mem,), dtype=torch.uint8, device=dev)
    for t in ts_in:
        _check(t.is_cuda and t.device == dev and t.is_contiguous() and t.dtype == torch.uint8)
    with torch.cuda.device(dev):
        sizes = torch.empty((n,), dtype=torch.int32, device=dev)
        types = torch.zeros((n,), dtype=torch.int32, device=dev)
        if compress_as_float:
            codec.float_get_compressed_info(ts_in, sizes, types)
        else:
            codec.ans_get_compressed_info(ts_in, sizes)
        sizes_h, types_h = sizes.cpu().tolist(), types.cpu().tolist()
    outs = []
    for i in range(n):
        if compress_as_float:
            _check(types_h[i] == types_h[0])  # must be consistent dtype
            outs.append(torch.empty((sizes_h[i],), dtype=_DTYPE_OF_FLOAT_TYPE[types_h[i]], device=dev))
        else:
            outs.append(torch.empty((sizes_h[i],), dtype=torch.uint8, device=dev))
    decompress_data(compress_as_float, ts_in, outs, checksum, scratch)
    return outs


# ---- additions: the fork's sparse path, unreachable from the reference's Python ---------------------
def max_sparse_float_compressed_size(dtype_tensor, size):
    return codec.get_max_sparse_float_compressed_size(_float_type(dtype_tensor), size)


def compress_data_sparse(ts_in, temp_mem=None, out_compressed=None, out_compressed_bytes=None):
    """floatCompressSparse (dietgpu/float/GpuFloatCodec.h:166-176) with compress_data's conventions."""
    _check(len(ts_in) > 0)
    dev = ts_in[0].device
    ft = _float_type(ts_in[0])
    for t in ts_in:
        _check(t.is_cuda and t.is_contiguous() and t.device == dev and t.dtype == ts_in[0].dtype)
    _, mx = _total_and_max(ts_in)
    max_bytes = codec.get_max_sparse_float_compressed_size(ft, mx)
    comp, sizes = _validate_out(out_compressed, out_compressed_bytes, len(ts_in), max_bytes, dev)
    with torch.cuda.device(dev):
        codec.float_compress_sparse(ft, ts_in, [comp[i] for i in range(len(ts_in))], sizes,
                                    K_DEFAULT_PRECISION, temp=temp_mem)
    return comp, sizes, codec.last_temp_bytes_used()


def decompress_data_sparse(ts_in, ts_out, temp_mem=None, out_status=None, out_decompressed_words=None):
    _check(len(ts_in) > 0 and len(ts_in) == len(ts_out))
    dev = ts_in[0].device
    _validate_status(out_status, out_decompressed_words, len(ts_in), dev)
    with torch.cuda.device(dev):
        codec.float_decompress_sparse(_float_type(ts_out[0]), ts_in, ts_out, out_status, out_decompressed_words,
                                      K_DEFAULT_PRECISION, temp=temp_mem)
    return codec.last_temp_bytes_used()


OPS = ["max_float_compressed_output_size", "max_float_compressed_size", "max_any_compressed_output_size",
       "max_any_compressed_size", "compress_data", "compress_data_split_size", "compress_data_simple",
       "decompress_data", "decompress_data_split_size", "decompress_data_simple"]

# ---- torch.ops.dietgpu.* registration --------------------------------------------------------------
# Operator schemas, verbatim in meaning from the reference's TORCH_LIBRARY_FRAGMENT
# (/root/reference/dietgpu/DietGpu.cpp:921-943), so that `torch.ops.dietgpu.compress_data(...)` call sites
# (ans_test.py, float_test.py, benchmark.py) resolve to this library.
_SCHEMAS = [
    "max_float_compressed_output_size(Tensor[] ts) -> (int, int)",
    "max_float_compressed_size(Tensor dtype, int size) -> int",
    "max_any_compressed_output_size(Tensor[] ts) -> (int, int)",
    "max_any_compressed_size(int bytes) -> int",
    "compress_data(bool compress_as_float, Tensor[] ts_in, bool checksum=False, Tensor? temp_mem=None, "
    "Tensor? out_compressed=None, Tensor? out_compressed_bytes=None) -> (Tensor, Tensor, int)",
    "compress_data_split_size(bool compress_as_float, Tensor t_in, Tensor t_in_split_sizes, bool checksum=False, "
    "Tensor? temp_mem=None, Tensor? out_compressed=None, Tensor? out_compressed_bytes=None) -> (Tensor[], Tensor, int)",
    "compress_data_simple(bool compress_as_float, Tensor[] ts_in, bool checksum=False, int? temp_mem=67108864) "
    "-> Tensor[]",
    "decompress_data(bool compress_as_float, Tensor[] ts_in, Tensor[] ts_out, bool checksum=False, "
    "Tensor? temp_mem=None, Tensor? out_status=None, Tensor? out_decompressed_words=None) -> (int)",
    "decompress_data_split_size(bool compress_as_float, Tensor[] ts_in, Tensor t_out, Tensor t_out_split_sizes, "
    "bool checksum=False, Tensor? temp_mem=None, Tensor? out_status=None, Tensor? out_decompressed_words=None) "
    "-> (int)",
    "decompress_data_simple(bool compress_as_float, Tensor[] ts_in, bool checksum=False, int? temp_mem=67108864) "
    "-> Tensor[]",
    # additions (the fork's sparse codec; no Python surface in the reference)
    "max_sparse_float_compressed_size(Tensor dtype, int size) -> int",
    "compress_data_sparse(Tensor[] ts_in, Tensor? temp_mem=None, Tensor? out_compressed=None, "
    "Tensor? out_compressed_bytes=None) -> (Tensor, Tensor, int)",
    "decompress_data_sparse(Tensor[] ts_in, Tensor[] ts_out, Tensor? temp_mem=None, Tensor? out_status=None, "
    "Tensor? out_decompressed_words=None) -> (int)",
]
_lib = None


def install():
    """Register the operators under the `dietgpu` namespace (idempotent). Returns torch.ops.dietgpu."""
    global _lib
    if _lib is None:
        lib = torch.library.Library("dietgpu", "DEF")
        g = globals()
        for schema in _SCHEMAS:
            name = schema.split("(", 1)[0]
            lib.define(schema)
            lib.impl(name, g[name], "CompositeImplicitAutograd")
        _lib = lib
    return torch.ops.dietgpu
