"""Seeded synthetic inputs shared by the CPU and GPU tests and bench.py.

The generator families follow the reference's own tests: exponential bytes
(dietgpu/ans/ANSTest.cu:18-31) and N(0,1) floats (dietgpu/float/FloatTest.cu:118-129), with
numpy RNGs so that the same data is produced on every machine.
"""
import numpy as np

FLOAT16, BFLOAT16, FLOAT32, FLOAT64 = 1, 2, 3, 4


def exp_bytes(n, lam=20.0, seed=10):
    """b = trunc(min(Exp(lam), 1) * 256) as u8 (255.99.. truncates to 255; 256 wraps to 0 as in C)."""
    rng = np.random.default_rng(seed)
    s = np.minimum(rng.exponential(1.0 / lam, size=n).astype(np.float32), np.float32(1.0))
    return (s.astype(np.float64) * 256.0).astype(np.int64).astype(np.uint8)


def float_words(ft, n, seed=10, scale=1.0, sparsity=0.0):
    """Bit patterns (unsigned words) of N(0, scale) samples in the given float type."""
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(n) * scale)
    if sparsity > 0:
        x = np.where(rng.random(n) < sparsity, 0.0, x)
    if ft == FLOAT64:
        return x.astype(np.float64).view(np.uint64)
    if ft == FLOAT32:
        return x.astype(np.float32).view(np.uint32)
    if ft == FLOAT16:
        return x.astype(np.float16).view(np.uint16)
    if ft == BFLOAT16:  # truncation, as the reference's test helper (FloatTest.cu:21-29)
        return (x.astype(np.float32).view(np.uint32) >> 16).astype(np.uint16)
    raise ValueError(ft)
