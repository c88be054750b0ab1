"""CPU-side checks of the drop-in boundary: libdietgpu_b200.so loads without a GPU, exports every
function include/dietgpu_b200.h declares (and nothing is declared that ctypes does not bind), and the
pure host-side size arithmetic agrees with the oracle."""
import ctypes
import os
import re

import pytest

import oracle

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dietgpu_b200.h")


def _declared():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dgb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_reference_surface():
    names = _declared()
    for need in ["dgb_ans_encode_batch_pointer", "dgb_ans_encode_batch_stride", "dgb_ans_encode_batch_split_size",
                 "dgb_ans_decode_batch_pointer", "dgb_ans_decode_batch_stride", "dgb_ans_decode_batch_split_size",
                 "dgb_ans_get_compressed_info", "dgb_float_compress", "dgb_float_compress_split_size",
                 "dgb_float_compress_sparse", "dgb_float_decompress", "dgb_float_decompress_split_size",
                 "dgb_float_decompress_sparse", "dgb_float_get_compressed_info", "dgb_ans_max_compressed_size",
                 "dgb_float_max_compressed_size", "dgb_sparse_float_max_compressed_size"]:
        assert need in names


def test_library_exports_every_declared_symbol():
    lib = ctypes.CDLL(os.path.join(ROOT, "dietgpu_fork_b200", "libdietgpu_b200.so"))
    for name in _declared():
        assert hasattr(lib, name), f"{name} declared in include/dietgpu_b200.h but not exported"


def test_ctypes_binding_covers_the_header():
    from dietgpu_fork_b200 import _capi
    assert sorted(_capi.EXPORTS) == _declared()


def test_version_and_launch_counter_without_gpu():
    from dietgpu_fork_b200 import _capi
    assert b"sm_100a" in _capi.lib.dgb_version()
    assert _capi.lib.dgb_profile_launches() == 0


@pytest.mark.parametrize("n", [0, 1, 4095, 4096, 4097, 1 << 20, (1 << 22) + 3])
def test_size_bounds_match_oracle(n):
    from dietgpu_fork_b200 import codec
    assert codec.get_max_compressed_size(n) == oracle.ans_max_compressed_size(n)
    for ft in (oracle.FLOAT16, oracle.BFLOAT16, oracle.FLOAT32, oracle.FLOAT64):
        assert codec.get_max_float_compressed_size(ft, n) == oracle.float_max_compressed_size(ft, n)
        assert codec.get_max_sparse_float_compressed_size(ft, n) == oracle.sparse_float_max_compressed_size(ft, n)


def test_invalid_arguments_fail_loudly_without_gpu():
    from dietgpu_fork_b200 import _capi
    # float type 0 (kUndefined) is rejected before any CUDA call
    rc = _capi.lib.dgb_float_compress(None, 0, 0, 10, 0, 1, None, None, None, None, None)
    assert rc < 0
    rc = _capi.lib.dgb_ans_encode_batch_pointer(None, 0, 13, 0, 1, None, None, None, None, None, None)
    assert rc < 0  # probBits outside 9..11 (dietgpu/ans/GpuANSEncode.cuh:700-720)


def test_torch_ops_register_with_reference_schemas():
    import torch
    from dietgpu_fork_b200 import torch_ops
    ops = torch_ops.install()
    assert torch_ops.install() is ops  # idempotent
    for name in torch_ops.OPS:
        assert hasattr(ops, name)
    assert ops.max_any_compressed_size(4096) == oracle.ans_max_compressed_size(4096)
    t16 = torch.empty(0, dtype=torch.float16)
    assert ops.max_float_compressed_size(t16, 1000) == oracle.float_max_compressed_size(oracle.FLOAT16, 1000)
    n, mx = ops.max_float_compressed_output_size([torch.empty(10, dtype=torch.float32),
                                                  torch.empty(99, dtype=torch.float32)])
    assert (n, mx) == (2, oracle.float_max_compressed_size(oracle.FLOAT32, 99))
