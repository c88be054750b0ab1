#!/usr/bin/env bash
# Builds the UNMODIFIED reference (NSagan271/dietgpu_fork) from its sources where they lie under
# /root/reference into oracle/_ref/libdietgpu_ref.so, for sm_100a. TEST INFRASTRUCTURE ONLY: it is
# the bit-exact oracle on the GPU box and the "reference CUDA build" timed beside ours.
#
# The reference's own CMake build does not configure on this toolchain: we
# compile its .cu/.cpp files directly. glog is replaced by oracle/ref_shim/glog/logging.h (the
# reference only uses CHECK*), and `__thrust_exec_check_disable__` (gone from CUDA 12.9's CCCL)
# is defined away. -DNDEBUG mirrors the reference's default RelWithDebInfo build.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${DIETGPU_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
OBJ="$OUT/obj"
if [ ! -d "$REF/dietgpu" ]; then
  echo "build_ref.sh: $REF not present; keeping any prebuilt $OUT/libdietgpu_ref.so" >&2
  exit 0
fi
mkdir -p "$OBJ"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-std=c++17 -O2 -DNDEBUG "-D__thrust_exec_check_disable__=" -gencode arch=compute_100a,code=sm_100a
       -I"$HERE/ref_shim" -I"$REF" -Xcompiler -fPIC -w)
SRCS=(dietgpu/ans/GpuANSEncode.cu dietgpu/ans/GpuANSDecode.cu dietgpu/ans/GpuANSInfo.cu
      dietgpu/float/GpuFloatCompress.cu dietgpu/float/GpuFloatDecompress.cu dietgpu/float/GpuFloatInfo.cu
      dietgpu/float/GpuSparseFloatCompress.cu dietgpu/float/GpuSparseFloatDecompress.cu
      dietgpu/utils/DeviceUtils.cpp dietgpu/utils/StackDeviceMemory.cpp)
pids=()
objs=()
for s in "${SRCS[@]}"; do
  o="$OBJ/$(echo "$s" | tr '/' '_').o"
  objs+=("$o")
  if [ ! -f "$o" ] || [ "$REF/$s" -nt "$o" ]; then
    "$NVCC" "${FLAGS[@]}" -x cu -c "$REF/$s" -o "$o" &
    pids+=($!)
  fi
done
o="$OBJ/ref_capi.o"
objs+=("$o")
"$NVCC" "${FLAGS[@]}" -c "$HERE/ref_shim/ref_capi.cu" -o "$o" &
pids+=($!)
for p in "${pids[@]}"; do wait "$p"; done
# The reference defines a few non-template __global__ helpers in headers that two of its own
# shared libraries include (its CMake links them separately); the definitions are identical.
"$NVCC" -shared -Xlinker --allow-multiple-definition -o "$OUT/libdietgpu_ref.so" "${objs[@]}"
echo "built $OUT/libdietgpu_ref.so"
