// Launch accounting for bench.py: every kernel launch of the library goes through a ProfScope,
// which counts it and, when profiling is enabled, brackets it with CUDA events on the launching
// stream so that per-kernel device time can be read back without a profiler.
#pragma once
#include <cuda_runtime.h>

namespace dgb {
struct ProfScope {
  ProfScope(const char* name, cudaStream_t s);
  ~ProfScope();
  const char* name_;
  cudaStream_t s_;
  cudaEvent_t start_ = nullptr;
};
} // namespace dgb
