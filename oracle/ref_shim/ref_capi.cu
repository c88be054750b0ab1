/*
 * extern "C" face of the UNMODIFIED reference library (compiled from /root/reference in place by
 * oracle/build_ref.sh). TEST INFRASTRUCTURE ONLY: tests/ and bench.py load
 * oracle/_ref/libdietgpu_ref.so through ctypes to (a) pin byte-stream parity of our kernels and
 * (b) time the reference's CUDA build next to ours. Nothing in dietgpu_fork_b200/ touches it.
 *
 * Each wrapper forwards to the reference entry point named in its comment with a
 * StackDeviceMemory built over the caller's temp buffer.
 */
#include <cuda_runtime.h>
#include <cassert>
#include <cstdint>
#include <string>
#include <vector>
#include "dietgpu/ans/GpuANSCodec.h"
#include "dietgpu/float/GpuFloatCodec.h"
#include "dietgpu/utils/StackDeviceMemory.h"

using namespace dietgpu;

namespace {
int curDev() {
  int d = 0;
  cudaGetDevice(&d);
  return d;
}
} // namespace

extern "C" {

/* dietgpu/ans/GpuANSCodec.h:22 */
uint32_t ref_ans_max_compressed_size(uint32_t n) { return getMaxCompressedSize(n); }
/* dietgpu/float/GpuFloatCodec.h:27-30 */
uint32_t ref_float_max_compressed_size(int ft, uint32_t n) {
  return getMaxFloatCompressedSize((FloatType)ft, n);
}
uint32_t ref_sparse_float_max_compressed_size(int ft, uint32_t n) {
  return getMaxSparseFloatCompressedSize((FloatType)ft, n);
}

/* ansEncodeBatchPointer, dietgpu/ans/GpuANSCodec.h:157-176 */
void ref_ans_encode_batch_pointer(void* temp, size_t tempBytes, int probBits, int useChecksum,
                                  uint32_t numInBatch, const void** in, const uint32_t* inSize,
                                  const uint32_t* histogram_dev, void** out, uint32_t* outSize_dev,
                                  cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  ansEncodeBatchPointer(res, ANSCodecConfig(probBits, useChecksum != 0), numInBatch, in, inSize,
                        histogram_dev, out, outSize_dev, stream);
}

/* ansDecodeBatchPointer, dietgpu/ans/GpuANSCodec.h:259-279; returns the error enum */
int ref_ans_decode_batch_pointer(void* temp, size_t tempBytes, int probBits, int useChecksum,
                                 uint32_t numInBatch, const void** in, void** out,
                                 const uint32_t* outCapacity, uint8_t* outSuccess_dev,
                                 uint32_t* outSize_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  auto st = ansDecodeBatchPointer(res, ANSCodecConfig(probBits, useChecksum != 0), numInBatch, in,
                                  out, outCapacity, outSuccess_dev, outSize_dev, stream);
  return (int)st.error;
}

/* floatCompress, dietgpu/float/GpuFloatCodec.h:105-140 */
void ref_float_compress(void* temp, size_t tempBytes, int floatType, int probBits, int useChecksum,
                        uint32_t numInBatch, const void** in, const uint32_t* inSize, void** out,
                        uint32_t* outSize_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  FloatCompressConfig cfg((FloatType)floatType, ANSCodecConfig(probBits, false), false,
                          useChecksum != 0);
  floatCompress(res, cfg, numInBatch, in, inSize, out, outSize_dev, stream);
}

/* floatDecompress, dietgpu/float/GpuFloatCodec.h:190-224 */
int ref_float_decompress(void* temp, size_t tempBytes, int floatType, int probBits,
                         int useChecksum, uint32_t numInBatch, const void** in, void** out,
                         const uint32_t* outCapacity, uint8_t* outSuccess_dev,
                         uint32_t* outSize_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  FloatDecompressConfig cfg((FloatType)floatType, ANSCodecConfig(probBits, false), false,
                            useChecksum != 0);
  auto st = floatDecompress(res, cfg, numInBatch, in, out, outCapacity, outSuccess_dev,
                            outSize_dev, stream);
  return (int)st.error;
}

/* floatCompressSparse / floatDecompressSparse, dietgpu/float/GpuFloatCodec.h:166-176, :262-273 */
void ref_float_compress_sparse(void* temp, size_t tempBytes, int floatType, int probBits,
                               int useChecksum, uint32_t numInBatch, const void** in,
                               const uint32_t* inSize, void** out, uint32_t* outSize_dev,
                               cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  FloatCompressConfig cfg((FloatType)floatType, ANSCodecConfig(probBits, false), false,
                          useChecksum != 0);
  floatCompressSparse(res, cfg, numInBatch, in, inSize, out, outSize_dev, stream);
}

int ref_float_decompress_sparse(void* temp, size_t tempBytes, int floatType, int probBits,
                                int useChecksum, uint32_t numInBatch, const void** in, void** out,
                                const uint32_t* outCapacity, uint8_t* outSuccess_dev,
                                uint32_t* outSize_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  FloatDecompressConfig cfg((FloatType)floatType, ANSCodecConfig(probBits, false), false,
                            useChecksum != 0);
  auto st = floatDecompressSparse(res, cfg, numInBatch, in, out, outCapacity, outSuccess_dev,
                                  outSize_dev, stream);
  return (int)st.error;
}

/* ansEncodeBatchStride / ansDecodeBatchStride, dietgpu/ans/GpuANSCodec.h:68-98, :222-257 */
void ref_ans_encode_batch_stride(void* temp, size_t tempBytes, int probBits, int useChecksum,
                                 uint32_t numInBatch, const void* in_dev, uint32_t inPerBatchSize,
                                 uint32_t inPerBatchStride, const uint32_t* histogram_dev,
                                 void* out_dev, uint32_t outPerBatchStride, uint32_t* outSize_dev,
                                 cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  ansEncodeBatchStride(res, ANSCodecConfig(probBits, useChecksum != 0), numInBatch, in_dev,
                       inPerBatchSize, inPerBatchStride, histogram_dev, out_dev, outPerBatchStride,
                       outSize_dev, stream);
}

int ref_ans_decode_batch_stride(void* temp, size_t tempBytes, int probBits, int useChecksum,
                                uint32_t numInBatch, const void* in_dev, uint32_t inPerBatchStride,
                                void* out_dev, uint32_t outPerBatchStride,
                                uint32_t outPerBatchCapacity, uint8_t* outSuccess_dev,
                                uint32_t* outSize_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  auto st = ansDecodeBatchStride(res, ANSCodecConfig(probBits, useChecksum != 0), numInBatch,
                                 in_dev, inPerBatchStride, out_dev, outPerBatchStride,
                                 outPerBatchCapacity, outSuccess_dev, outSize_dev, stream);
  return (int)st.error;
}

/* ansGetCompressedInfo, dietgpu/ans/GpuANSCodec.h:315-328 */
void ref_ans_get_compressed_info(void* temp, size_t tempBytes, const void** in,
                                 uint32_t numInBatch, uint32_t* outSizes_dev,
                                 uint32_t* outChecksum_dev, cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  ansGetCompressedInfo(res, in, numInBatch, outSizes_dev, outChecksum_dev, stream);
}

/* floatGetCompressedInfo, dietgpu/float/GpuFloatCodec.h:281-300 */
void ref_float_get_compressed_info(void* temp, size_t tempBytes, const void** in,
                                   uint32_t numInBatch, uint32_t* outSizes_dev,
                                   uint32_t* outTypes_dev, uint32_t* outChecksum_dev,
                                   cudaStream_t stream) {
  StackDeviceMemory res(curDev(), temp, tempBytes);
  floatGetCompressedInfo(res, in, numInBatch, outSizes_dev, outTypes_dev, outChecksum_dev, stream);
}

} // extern "C"
