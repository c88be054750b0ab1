"""ctypes binding of libdietgpu_b200.so (include/dietgpu_b200.h).

The library is the product: there is no Python or CPU fallback. If the shared object is missing
this module raises at import time so that nothing silently runs without the CUDA kernels.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdietgpu_b200.so")

OK = 0
CHECKSUM_MISMATCH = 1
ERRORS = {-1: "invalid argument", -2: "CUDA error", -3: "misaligned pointer", -4: "internal error",
          -5: "batch too large for the fixed scratch of the no-allocation entry point"}
NOALLOC_MAX_BATCH = 32  # DGB_NOALLOC_MAX_BATCH

FLOAT16, BFLOAT16, FLOAT32, FLOAT64 = 1, 2, 3, 4


class DietGpuError(RuntimeError):
    pass


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `make -C dietgpu_fork_b200/csrc` "
            "(or `python -c 'import __graft_entry__ as g; g.build()'`). There is no fallback path.")
    L = C.CDLL(LIB_PATH)
    u32, vp, i, sz = C.c_uint32, C.c_void_p, C.c_int, C.c_size_t
    L.dgb_version.restype = C.c_char_p
    L.dgb_last_temp_bytes_used.restype = sz
    L.dgb_ans_max_compressed_size.restype = u32
    L.dgb_ans_max_compressed_size.argtypes = [u32]
    L.dgb_float_max_compressed_size.restype = u32
    L.dgb_float_max_compressed_size.argtypes = [i, u32]
    L.dgb_sparse_float_max_compressed_size.restype = u32
    L.dgb_sparse_float_max_compressed_size.argtypes = [i, u32]
    L.dgb_temp_bytes_required.restype = sz
    L.dgb_temp_bytes_required.argtypes = [i, i, u32, u32]
    enc_tail = [vp, sz, i, i, u32]
    L.dgb_ans_encode_batch_pointer.argtypes = enc_tail + [vp, vp, vp, vp, vp, vp]
    L.dgb_ans_encode_batch_stride.argtypes = enc_tail + [vp, u32, u32, vp, vp, u32, vp, vp]
    L.dgb_ans_encode_batch_split_size.argtypes = enc_tail + [vp, vp, vp, vp, u32, vp, vp]
    mm = [vp, u32, vp]
    L.dgb_ans_decode_batch_pointer.argtypes = enc_tail + [vp, vp, vp, vp, vp, vp] + mm
    L.dgb_ans_decode_batch_stride.argtypes = enc_tail + [vp, u32, vp, u32, u32, vp, vp, vp] + mm
    L.dgb_ans_decode_batch_split_size.argtypes = enc_tail + [vp, vp, vp, vp, vp, vp] + mm
    L.dgb_ans_get_compressed_info.argtypes = [vp, i, u32, vp, vp, vp, sz, vp]
    f_tail = [vp, sz, i, i, i, u32]
    L.dgb_float_compress.argtypes = f_tail + [vp, vp, vp, vp, vp]
    L.dgb_float_compress_split_size.argtypes = f_tail + [vp, vp, vp, u32, vp, vp]
    L.dgb_float_compress_sparse.argtypes = f_tail + [vp, vp, vp, vp, vp]
    L.dgb_float_decompress.argtypes = f_tail + [vp, vp, vp, vp, vp, vp] + mm
    L.dgb_float_decompress_split_size.argtypes = f_tail + [vp, vp, vp, vp, vp, vp] + mm
    L.dgb_float_decompress_sparse.argtypes = f_tail + [vp, vp, vp, vp, vp, vp]
    L.dgb_float_get_compressed_info.argtypes = [vp, i, u32, vp, vp, vp, vp, sz, vp]
    L.dgb_float_compress_host.argtypes = [i, i, i, u32, vp, vp, vp, vp, vp, vp]
    L.dgb_float_decompress_host.argtypes = [i, i, i, u32, vp, vp, vp, vp, vp, vp, vp]
    L.dgb_ans_encode_host.argtypes = [i, i, u32, vp, vp, vp, vp, vp, vp]
    L.dgb_ans_decode_host.argtypes = [i, i, u32, vp, vp, vp, vp, vp, vp, vp]
    L.dgb_host_release.restype = None
    L.dgb_decode_batch_noalloc.argtypes = [i, i, u32, vp, vp, vp, vp, vp, vp]
    L.dgb_decode_batch_noalloc.restype = i
    L.dgb_peer_alloc.argtypes = [sz, vp, vp]
    L.dgb_peer_open.argtypes = [vp, vp]
    L.dgb_peer_close.argtypes = [vp]
    L.dgb_peer_free.argtypes = [vp]
    L.dgb_peer_pull_float.argtypes = [u32, vp, vp, vp, u32, vp]
    L.dgb_peer_pull_float.restype = i
    L.dgb_profile_launches.restype = C.c_ulonglong
    L.dgb_profile_enable.argtypes = [i]
    L.dgb_set_fused_min_tiles.argtypes = [C.c_longlong]
    L.dgb_set_fused_min_tiles.restype = None
    L.dgb_profile_report.argtypes = [C.c_char_p, sz]
    L.dgb_profile_report.restype = i
    for name in ("dgb_ans_encode_batch_pointer", "dgb_ans_encode_batch_stride",
                 "dgb_ans_encode_batch_split_size", "dgb_ans_decode_batch_pointer",
                 "dgb_ans_decode_batch_stride", "dgb_ans_decode_batch_split_size",
                 "dgb_ans_get_compressed_info", "dgb_float_compress", "dgb_float_compress_split_size",
                 "dgb_float_compress_sparse", "dgb_float_decompress", "dgb_float_decompress_split_size",
                 "dgb_float_decompress_sparse", "dgb_float_get_compressed_info", "dgb_float_compress_host",
                 "dgb_float_decompress_host", "dgb_ans_encode_host", "dgb_ans_decode_host", "dgb_peer_alloc",
                 "dgb_peer_open", "dgb_peer_close", "dgb_peer_free"):
        getattr(L, name).restype = i
    return L


lib = _load()

# every symbol include/dietgpu_b200.h declares (checked by tests/test_capi_symbols.py)
EXPORTS = [
    "dgb_version", "dgb_last_temp_bytes_used", "dgb_profile_launches", "dgb_profile_enable",
    "dgb_profile_report", "dgb_set_fused_min_tiles", "dgb_ans_max_compressed_size",
    "dgb_float_max_compressed_size", "dgb_sparse_float_max_compressed_size",
    "dgb_temp_bytes_required", "dgb_ans_encode_batch_pointer", "dgb_ans_encode_batch_stride",
    "dgb_ans_encode_batch_split_size", "dgb_ans_decode_batch_pointer", "dgb_ans_decode_batch_stride",
    "dgb_ans_decode_batch_split_size", "dgb_ans_get_compressed_info", "dgb_float_compress",
    "dgb_float_compress_split_size", "dgb_float_compress_sparse", "dgb_float_decompress",
    "dgb_float_decompress_split_size", "dgb_float_decompress_sparse", "dgb_float_get_compressed_info",
    "dgb_float_compress_host", "dgb_float_decompress_host", "dgb_ans_encode_host", "dgb_ans_decode_host",
    "dgb_host_release", "dgb_decode_batch_noalloc", "dgb_peer_alloc", "dgb_peer_open", "dgb_peer_close", "dgb_peer_free", "dgb_peer_pull_float",
]


def check(rc, what):
    """Raise on error codes; returns True when a checksum mismatch was reported."""
    if rc == OK:
        return False
    if rc == CHECKSUM_MISMATCH:
        return True
    raise DietGpuError(f"{what} failed: {ERRORS.get(rc, rc)}")


def ptr_array(ptrs):
    return (C.c_void_p * len(ptrs))(*ptrs)


def u32_array(vals):
    return (C.c_uint32 * len(vals))(*vals)
