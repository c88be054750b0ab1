"""dietgpu_fork_b200 — B200-native (sm_100a) DietGPU codec: batched byte-oriented rANS and the
float codecs layered on it, behind the reference's own interfaces.

Layers:
  csrc/ + libdietgpu_b200.so   hand-written CUDA kernels and the C ABI (include/dietgpu_b200.h)
  codec                        the reference's C++ API over torch tensors (raw device pointers)
  torch_ops                    the reference's `torch.ops.dietgpu.*` operators (same names/semantics)
"""
from . import _capi  # noqa: F401  (fails loudly when the CUDA library is missing)
from . import codec  # noqa: F401
from ._capi import BFLOAT16, FLOAT16, FLOAT32, FLOAT64, DietGpuError  # noqa: F401

__all__ = ["codec", "FLOAT16", "BFLOAT16", "FLOAT32", "FLOAT64", "DietGpuError"]
