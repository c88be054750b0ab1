"""ctypes/numpy front end of the CPU oracle (oracle/dietgpu_oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (dietgpu_fork_b200/) never imports it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libdietgpu_oracle.so")

FLOAT16, BFLOAT16, FLOAT32, FLOAT64 = 1, 2, 3, 4
WORD_DTYPE = {FLOAT16: np.uint16, BFLOAT16: np.uint16, FLOAT32: np.uint32, FLOAT64: np.uint64}


def build(force=False):
    src = os.path.join(_HERE, "dietgpu_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        u32, vp, i = C.c_uint32, C.c_void_p, C.c_int
        L.dgo_ans_max_compressed_size.restype = u32
        L.dgo_ans_max_compressed_size.argtypes = [u32]
        L.dgo_float_max_compressed_size.restype = u32
        L.dgo_float_max_compressed_size.argtypes = [i, u32]
        L.dgo_sparse_float_max_compressed_size.restype = u32
        L.dgo_sparse_float_max_compressed_size.argtypes = [i, u32]
        L.dgo_float_raw_section_bytes.restype = u32
        L.dgo_float_raw_section_bytes.argtypes = [i, u32]
        L.dgo_histogram.argtypes = [vp, u32, vp]
        L.dgo_normalize.restype = i
        L.dgo_normalize.argtypes = [vp, u32, i, vp, vp]
        L.dgo_checksum.restype = u32
        L.dgo_checksum.argtypes = [vp, u32]
        L.dgo_ans_encode.restype = u32
        L.dgo_ans_encode.argtypes = [vp, u32, i, i, vp, vp]
        L.dgo_ans_decode.restype = i
        L.dgo_ans_decode.argtypes = [vp, vp, u32, vp]
        L.dgo_ans_defined_mask.restype = u32
        L.dgo_ans_defined_mask.argtypes = [vp, vp]
        L.dgo_float_compress.restype = u32
        L.dgo_float_compress.argtypes = [i, vp, u32, i, i, vp]
        L.dgo_float_decompress.restype = i
        L.dgo_float_decompress.argtypes = [vp, vp, u32, vp]
        L.dgo_float_defined_mask.restype = u32
        L.dgo_float_defined_mask.argtypes = [vp, vp]
        L.dgo_sparse_float_compress.restype = u32
        L.dgo_sparse_float_compress.argtypes = [i, vp, u32, i, i, vp]
        L.dgo_sparse_float_decompress.restype = i
        L.dgo_sparse_float_decompress.argtypes = [i, vp, vp, u32, vp]
        L.dgo_sparse_float_defined_mask.restype = u32
        L.dgo_sparse_float_defined_mask.argtypes = [i, vp, vp]
        L.dgo_num_threads.restype = i
        L.dgo_set_num_threads.argtypes = [i]
        L.dgo_ans_encode_batch.argtypes = [u32, vp, vp, i, vp, vp]
        L.dgo_ans_decode_batch.argtypes = [u32, vp, vp, vp, vp]
        L.dgo_float_compress_batch.argtypes = [i, u32, vp, vp, i, vp, vp]
        L.dgo_float_decompress_batch.argtypes = [u32, vp, vp, vp, vp]
        _lib = L
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _bytes(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint8).reshape(-1)


# ---- sizes ---------------------------------------------------------------
def ans_max_compressed_size(n):
    return lib().dgo_ans_max_compressed_size(n)


def float_max_compressed_size(ft, n):
    return lib().dgo_float_max_compressed_size(ft, n)


def sparse_float_max_compressed_size(ft, n):
    return lib().dgo_sparse_float_max_compressed_size(ft, n)


def float_raw_section_bytes(ft, n):
    return lib().dgo_float_raw_section_bytes(ft, n)


# ---- statistics ----------------------------------------------------------
def histogram(data):
    d = _bytes(data)
    h = np.zeros(256, np.uint32)
    lib().dgo_histogram(_p(d), d.size, _p(h))
    return h


def normalize(counts, total, prob_bits):
    counts = np.ascontiguousarray(counts, np.uint32)
    pdf = np.zeros(256, np.uint32)
    cdf = np.zeros(256, np.uint32)
    ok = lib().dgo_normalize(_p(counts), int(total), prob_bits, _p(pdf), _p(cdf))
    return (pdf, cdf) if ok else (None, None)


def checksum(data):
    d = _bytes(data)
    return lib().dgo_checksum(_p(d), d.size)


# ---- ANS -----------------------------------------------------------------
def ans_encode(data, prob_bits=10, use_checksum=False, hist=None):
    d = _bytes(data)
    out = np.zeros(ans_max_compressed_size(d.size), np.uint8)
    hp = _p(np.ascontiguousarray(hist, np.uint32)) if hist is not None else None
    n = lib().dgo_ans_encode(_p(d), d.size, prob_bits, int(use_checksum), hp, _p(out))
    return out[:n].copy()


def ans_decode(archive, capacity=None):
    a = _bytes(archive)
    n = int(a[8:12].view(np.uint32)[0])
    cap = n if capacity is None else capacity
    out = np.zeros(max(cap, 1), np.uint8)
    sz = C.c_uint32(0)
    ok = lib().dgo_ans_decode(_p(a), _p(out), cap, C.byref(sz))
    return bool(ok), sz.value, out[:min(cap, sz.value)] if ok else out[:0]


def ans_defined_mask(archive):
    a = _bytes(archive)
    m = np.zeros(a.size, np.uint8)
    n = lib().dgo_ans_defined_mask(_p(a), _p(m))
    return m[:n].astype(bool)


# ---- float ---------------------------------------------------------------
def float_compress(ft, words, prob_bits=10, use_checksum=False):
    w = np.ascontiguousarray(words).view(WORD_DTYPE[ft]).reshape(-1)
    out = np.zeros(float_max_compressed_size(ft, w.size), np.uint8)
    n = lib().dgo_float_compress(ft, _p(w), w.size, prob_bits, int(use_checksum), _p(out))
    return out[:n].copy()


def float_decompress(archive, capacity=None):
    a = _bytes(archive)
    n = int(a[4:8].view(np.uint32)[0])
    ft = int(a[8]) & 0xF
    cap = n if capacity is None else capacity
    out = np.zeros(max(cap, 1), WORD_DTYPE[ft])
    sz = C.c_uint32(0)
    ok = lib().dgo_float_decompress(_p(a), _p(out), cap, C.byref(sz))
    return bool(ok), sz.value, out[:min(cap, sz.value)] if ok else out[:0]


def float_defined_mask(archive):
    a = _bytes(archive)
    m = np.zeros(a.size, np.uint8)
    n = lib().dgo_float_defined_mask(_p(a), _p(m))
    return m[:n].astype(bool)


def sparse_float_compress(ft, words, prob_bits=10, use_checksum=False):
    w = np.ascontiguousarray(words).view(WORD_DTYPE[ft]).reshape(-1)
    out = np.zeros(sparse_float_max_compressed_size(ft, w.size), np.uint8)
    n = lib().dgo_sparse_float_compress(ft, _p(w), w.size, prob_bits, int(use_checksum), _p(out))
    return out[:n].copy()


def sparse_float_decompress(ft, archive, capacity=None):
    a = _bytes(archive)
    n = int(a[0:4].view(np.uint32)[0])
    cap = n if capacity is None else capacity
    out = np.zeros(max(cap, 1), WORD_DTYPE[ft])
    sz = C.c_uint32(0)
    ok = lib().dgo_sparse_float_decompress(ft, _p(a), _p(out), cap, C.byref(sz))
    return bool(ok), sz.value, out[:min(cap, sz.value)] if ok else out[:0]


def sparse_float_defined_mask(ft, archive):
    a = _bytes(archive)
    m = np.zeros(a.size, np.uint8)
    n = lib().dgo_sparse_float_defined_mask(ft, _p(a), _p(m))
    return m[:n].astype(bool)


# ---- batched multithreaded drivers (the timed CPU baseline) ----------------
def num_threads():
    return lib().dgo_num_threads()


def _ptr_array(arrs):
    return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])


def ans_encode_batch(ins, outs, prob_bits=10):
    n = np.array([a.size for a in ins], np.uint32)
    ob = np.zeros(len(ins), np.uint32)
    lib().dgo_ans_encode_batch(len(ins), _ptr_array(ins), _p(n), prob_bits, _ptr_array(outs), _p(ob))
    return ob


def ans_decode_batch(archives, outs):
    cap = np.array([a.size for a in outs], np.uint32)
    sz = np.zeros(len(outs), np.uint32)
    lib().dgo_ans_decode_batch(len(outs), _ptr_array(archives), _ptr_array(outs), _p(cap), _p(sz))
    return sz


def float_compress_batch(ft, ins, outs, prob_bits=10):
    n = np.array([a.size for a in ins], np.uint32)
    ob = np.zeros(len(ins), np.uint32)
    lib().dgo_float_compress_batch(ft, len(ins), _ptr_array(ins), _p(n), prob_bits,
                                   _ptr_array(outs), _p(ob))
    return ob


def float_decompress_batch(archives, outs):
    cap = np.array([a.size for a in outs], np.uint32)
    sz = np.zeros(len(outs), np.uint32)
    lib().dgo_float_decompress_batch(len(outs), _ptr_array(archives), _ptr_array(outs), _p(cap), _p(sz))
    return sz
