// dietgpu_b200.hpp — header-only C++ adaptor that re-exports the reference's C++ API
// (namespace dietgpu; /root/reference/dietgpu/ans/GpuANSCodec.h:22-341,
// /root/reference/dietgpu/float/GpuFloatCodec.h:32-322,
// /root/reference/dietgpu/utils/StackDeviceMemory.h:33-298) on top of the C ABI in
// dietgpu_b200.h. A caller of the reference replaces
//     #include "dietgpu/float/GpuFloatCodec.h"   +   -lgpu_float_compress -lgpu_ans -ldietgpu_utils
// by
//     #include "dietgpu_b200.hpp"                +   -ldietgpu_b200
// and keeps its code: same function names, argument order, units, host/device placement of every
// array, and the same status structs. Differences, all deliberate:
//   * host-side precondition failures throw std::runtime_error instead of glog CHECK-aborting;
//   * StackDeviceMemory only carries the caller's scratch region (the library bump-allocates
//     inside it and takes any shortfall from the stream-ordered pool instead of cudaMalloc);
//     alloc<T>() / copyAlloc() / copyToHost() are provided for callers that used them.
#pragma once
#include <cuda_runtime.h>

#include <cassert>
#include <cstdint>
#include <stdexcept>
#include <string>
#include <utility>
#include <vector>

#include "dietgpu_b200.h"

namespace dietgpu {

// ---- utils/DeviceUtils.h (the subset callers of the codec use: :17-23, :41-49, :61-67, :100, :111-119) ------
#ifndef CUDA_VERIFY
#define CUDA_VERIFY(X)                                                                              \
  do {                                                                                              \
    auto err__ = (X);                                                                               \
    if (err__ != cudaSuccess)                                                                       \
      throw std::runtime_error(std::string("CUDA error ") + cudaGetErrorName(err__) + " at " + __FILE__ + ":" + \
                               std::to_string(__LINE__));                                           \
  } while (0)
#endif
#ifndef CUDA_TEST_ERROR
#define CUDA_TEST_ERROR() CUDA_VERIFY(cudaGetLastError())
#endif

inline int getCurrentDevice() {
  int dev = -1;
  CUDA_VERIFY(cudaGetDevice(&dev));
  return dev;
}
inline void setCurrentDevice(int device) { CUDA_VERIFY(cudaSetDevice(device)); }
inline int getNumDevices() {
  int n = 0;
  cudaError_t err = cudaGetDeviceCount(&n);
  if (err == cudaErrorNoDevice) return 0;
  CUDA_VERIFY(err);
  return n;
}
inline int getDeviceForAddress(const void* p) {
  if (!p) return -1;
  cudaPointerAttributes att;
  cudaError_t err = cudaPointerGetAttributes(&att, p);
  if (err != cudaSuccess) {
    cudaGetLastError();
    return -1;
  }
  return att.type == cudaMemoryTypeDevice || att.type == cudaMemoryTypeManaged ? att.device : -1;
}
class DeviceScope {
 public:
  explicit DeviceScope(int device) : prev_(getCurrentDevice()) {
    if (prev_ != device) setCurrentDevice(device);
    else prev_ = -1;
  }
  ~DeviceScope() {
    if (prev_ != -1) cudaSetDevice(prev_);
  }
 private:
  int prev_;
};

// ---- utils/StackDeviceMemory.h --------------------------------------------------------------
template <typename T>
struct GpuMemoryReservation {
  GpuMemoryReservation() = default;
  GpuMemoryReservation(T* p, size_t n, cudaStream_t s, bool owned) : ptr_(p), num_(n), stream_(s), owned_(owned) {}
  GpuMemoryReservation(GpuMemoryReservation&& o) noexcept { *this = std::move(o); }
  GpuMemoryReservation& operator=(GpuMemoryReservation&& o) noexcept {
    release();
    ptr_ = o.ptr_; num_ = o.num_; stream_ = o.stream_; owned_ = o.owned_;
    o.ptr_ = nullptr; o.owned_ = false;
    return *this;
  }
  GpuMemoryReservation(const GpuMemoryReservation&) = delete;
  GpuMemoryReservation& operator=(const GpuMemoryReservation&) = delete;
  ~GpuMemoryReservation() { release(); }
  T* data() { return ptr_; }
  const T* data() const { return ptr_; }
  size_t size() const { return num_; }
  std::vector<T> copyToHost(cudaStream_t stream) const {
    std::vector<T> out(num_);
    cudaMemcpyAsync(out.data(), ptr_, num_ * sizeof(T), cudaMemcpyDeviceToHost, stream);
    cudaStreamSynchronize(stream);
    return out;
  }
 private:
  void release() {
    if (owned_ && ptr_) cudaFreeAsync(ptr_, stream_);
    ptr_ = nullptr;
  }
  T* ptr_ = nullptr;
  size_t num_ = 0;
  cudaStream_t stream_ = nullptr;
  bool owned_ = false;
};

// All allocations are aligned to this boundary (utils/StackDeviceMemory.h:25-28)
constexpr size_t kSDMAlignment = 256;

class StackDeviceMemory {
 public:
  // owning: allocates `allocPerDevice` bytes on `device`
  StackDeviceMemory(int device, size_t allocPerDevice) : device_(device), bytes_(allocPerDevice), owned_(true) {
    int prev = 0;
    cudaGetDevice(&prev);
    cudaSetDevice(device);
    if (bytes_ && cudaMalloc(&base_, bytes_) != cudaSuccess) throw std::runtime_error("StackDeviceMemory: cudaMalloc failed");
    cudaSetDevice(prev);
  }
  // non-owning: wraps caller memory
  StackDeviceMemory(int device, void* p, size_t size) : device_(device), base_(p), bytes_(p ? size : 0), owned_(false) {}
  StackDeviceMemory(const StackDeviceMemory&) = delete;
  StackDeviceMemory& operator=(const StackDeviceMemory&) = delete;
  ~StackDeviceMemory() {
    if (owned_ && base_) cudaFree(base_);
  }
  int getDevice() const { return device_; }
  void* data() const { return base_; }
  size_t size() const { return bytes_; }
  // Temporary device arrays for the CALLER's own use (stream ordered, freed with the reservation).
  template <typename T>
  GpuMemoryReservation<T> alloc(cudaStream_t stream, size_t num) {
    T* p = nullptr;
    if (cudaMallocAsync((void**)&p, (num ? num : 1) * sizeof(T), stream) != cudaSuccess)
      throw std::runtime_error("StackDeviceMemory::alloc failed");
    return GpuMemoryReservation<T>(p, num, stream, true);
  }
  template <typename T>
  GpuMemoryReservation<T> copyAlloc(cudaStream_t stream, const T* hostPtr, size_t num) {
    auto r = alloc<T>(stream, num);
    cudaMemcpyAsync(r.data(), hostPtr, num * sizeof(T), cudaMemcpyHostToDevice, stream);
    return r;
  }
  template <typename T>
  GpuMemoryReservation<T> copyAlloc(cudaStream_t stream, const std::vector<T>& v) {
    return copyAlloc(stream, v.data(), v.size());
  }
  // high-water mark of the library's scratch use in the last call on this thread
  size_t getMaxMemoryUsage() const { return dgb_last_temp_bytes_used(); }
 private:
  int device_;
  void* base_ = nullptr;
  size_t bytes_;
  bool owned_;
};

inline StackDeviceMemory makeStackMemory(size_t bytes = 256 * 1024 * 1024) {
  int dev = 0;
  cudaGetDevice(&dev);
  return StackDeviceMemory(dev, bytes);
}

namespace detail {
inline void check(int rc, const char* what) {
  if (rc < 0) throw std::runtime_error(std::string("dietgpu_b200: ") + what + " failed with code " + std::to_string(rc));
}
} // namespace detail

// ---- ans/GpuANSCodec.h ----------------------------------------------------------------------
constexpr int kANSRequiredAlignment = 4;
constexpr int kANSDefaultProbBits = 10;

inline uint32_t getMaxCompressedSize(uint32_t uncompressedBytes) { return dgb_ans_max_compressed_size(uncompressedBytes); }

struct ANSCodecConfig {
  inline ANSCodecConfig() : probBits(kANSDefaultProbBits), useChecksum(false) {}
  explicit inline ANSCodecConfig(int pb, bool checksum = false) : probBits(pb), useChecksum(checksum) {}
  int probBits;
  bool useChecksum;
};

enum class ANSDecodeError : uint32_t { None = 0, ChecksumMismatch = 1 };

struct ANSDecodeStatus {
  inline ANSDecodeStatus() : error(ANSDecodeError::None) {}
  ANSDecodeError error;
  std::vector<std::pair<int, std::string>> errorInfo;
};

namespace detail {
template <typename Status, typename Err>
inline Status makeStatus(int rc, const std::vector<uint32_t>& idx, uint32_t n, const char* what) {
  check(rc, what);
  Status st;
  if (rc == DGB_CHECKSUM_MISMATCH) {
    st.error = Err::ChecksumMismatch;
    for (uint32_t i = 0; i < n && i < idx.size(); ++i)
      st.errorInfo.emplace_back((int)idx[i], "Checksum mismatch in batch member " + std::to_string(idx[i]));
  }
  return st;
}
} // namespace detail

inline void ansEncodeBatchStride(StackDeviceMemory& res, const ANSCodecConfig& config, uint32_t numInBatch,
                                 const void* in_dev, uint32_t inPerBatchSize, uint32_t inPerBatchStride,
                                 const uint32_t* histogram_dev, void* out_dev, uint32_t outPerBatchStride,
                                 uint32_t* outBatchSize_dev, cudaStream_t stream) {
  detail::check(dgb_ans_encode_batch_stride(res.data(), res.size(), config.probBits, config.useChecksum, numInBatch,
                                            in_dev, inPerBatchSize, inPerBatchStride, histogram_dev, out_dev,
                                            outPerBatchStride, outBatchSize_dev, stream),
                "ansEncodeBatchStride");
}

inline void ansEncodeBatchPointer(StackDeviceMemory& res, const ANSCodecConfig& config, uint32_t numInBatch,
                                  const void** in, const uint32_t* inSize, const uint32_t* histogram_dev, void** out,
                                  uint32_t* outSize_dev, cudaStream_t stream) {
  detail::check(dgb_ans_encode_batch_pointer(res.data(), res.size(), config.probBits, config.useChecksum, numInBatch,
                                             in, inSize, histogram_dev, out, outSize_dev, stream),
                "ansEncodeBatchPointer");
}

inline void ansEncodeBatchSplitSize(StackDeviceMemory& res, const ANSCodecConfig& config, uint32_t numInBatch,
                                    const void* in_dev, const uint32_t* inSplitSizes, const uint32_t* histogram_dev,
                                    void* out_dev, uint32_t outStride, uint32_t* outSize_dev, cudaStream_t stream) {
  detail::check(dgb_ans_encode_batch_split_size(res.data(), res.size(), config.probBits, config.useChecksum,
                                                numInBatch, in_dev, inSplitSizes, histogram_dev, out_dev, outStride,
                                                outSize_dev, stream),
                "ansEncodeBatchSplitSize");
}

inline ANSDecodeStatus ansDecodeBatchStride(StackDeviceMemory& res, const ANSCodecConfig& config, uint32_t numInBatch,
                                            const void* in_dev, uint32_t inPerBatchStride, void* out_dev,
                                            uint32_t outPerBatchStride, uint32_t outPerBatchCapacity,
                                            uint8_t* outSuccess_dev, uint32_t* outSize_dev, cudaStream_t stream) {
  std::vector<uint32_t> idx(numInBatch);
  uint32_t n = 0;
  int rc = dgb_ans_decode_batch_stride(res.data(), res.size(), config.probBits, config.useChecksum, numInBatch, in_dev,
                                       inPerBatchStride, out_dev, outPerBatchStride, outPerBatchCapacity,
                                       outSuccess_dev, outSize_dev, stream, idx.data(), numInBatch, &n);
  return detail::makeStatus<ANSDecodeStatus, ANSDecodeError>(rc, idx, n, "ansDecodeBatchStride");
}

inline ANSDecodeStatus ansDecodeBatchPointer(StackDeviceMemory& res, const ANSCodecConfig& config, uint32_t numInBatch,
                                             const void** in, void** out, const uint32_t* outCapacity,
                                             uint8_t* outSuccess_dev, uint32_t* outSize_dev, cudaStream_t stream) {
  std::vector<uint32_t> idx(numInBatch);
  uint32_t n = 0;
  int rc = dgb_ans_decode_batch_pointer(res.data(), res.size(), config.probBits, config.useChecksum, numInBatch, in,
                                        out, outCapacity, outSuccess_dev, outSize_dev, stream, idx.data(), numInBatch,
                                        &n);
  return detail::makeStatus<ANSDecodeStatus, ANSDecodeError>(rc, idx, n, "ansDecodeBatchPointer");
}

inline ANSDecodeStatus ansDecodeBatchSplitSize(StackDeviceMemory& res, const ANSCodecConfig& config,
                                               uint32_t numInBatch, const void** in, void* out_dev,
                                               const uint32_t* outSplitSizes, uint8_t* outSuccess_dev,
                                               uint32_t* outSize_dev, cudaStream_t stream) {
  std::vector<uint32_t> idx(numInBatch);
  uint32_t n = 0;
  int rc = dgb_ans_decode_batch_split_size(res.data(), res.size(), config.probBits, config.useChecksum, numInBatch, in,
                                           out_dev, outSplitSizes, outSuccess_dev, outSize_dev, stream, idx.data(),
                                           numInBatch, &n);
  return detail::makeStatus<ANSDecodeStatus, ANSDecodeError>(rc, idx, n, "ansDecodeBatchSplitSize");
}

inline void ansGetCompressedInfo(StackDeviceMemory& res, const void** in, uint32_t numInBatch, uint32_t* outSizes_dev,
                                 uint32_t* outChecksum_dev, cudaStream_t stream) {
  detail::check(dgb_ans_get_compressed_info(in, 0, numInBatch, outSizes_dev, outChecksum_dev, res.data(), res.size(),
                                            stream),
                "ansGetCompressedInfo");
}

inline void ansGetCompressedInfoDevice(StackDeviceMemory& res, const void** in_dev, uint32_t numInBatch,
                                       uint32_t* outSizes_dev, uint32_t* outChecksum_dev, cudaStream_t stream) {
  detail::check(dgb_ans_get_compressed_info(in_dev, 1, numInBatch, outSizes_dev, outChecksum_dev, res.data(),
                                            res.size(), stream),
                "ansGetCompressedInfoDevice");
}

// ---- float/GpuFloatCodec.h -----------------------------------------------------------------
enum class FloatType : uint32_t { kUndefined = 0, kFloat16 = 1, kBFloat16 = 2, kFloat32 = 3, kFloat64 = 4 };

inline uint32_t getMaxFloatCompressedSize(FloatType floatType, uint32_t size) {
  return dgb_float_max_compressed_size((int)floatType, size);
}
inline uint32_t getMaxSparseFloatCompressedSize(FloatType floatType, uint32_t size) {
  return dgb_sparse_float_max_compressed_size((int)floatType, size);
}

struct FloatCodecConfig {
  inline FloatCodecConfig() : floatType(FloatType::kFloat16), useChecksum(false), is16ByteAligned(false) {}
  inline FloatCodecConfig(FloatType ft, const ANSCodecConfig& ansConf, bool align, bool checksum = false)
      : floatType(ft), useChecksum(checksum), ansConfig(ansConf), is16ByteAligned(align) {
    assert(!ansConf.useChecksum);
  }
  FloatType floatType;
  bool useChecksum;
  ANSCodecConfig ansConfig;
  bool is16ByteAligned; // recomputed internally, as in the reference (GpuFloatDecompress.cu:36-47)
};
using FloatCompressConfig = FloatCodecConfig;
using FloatDecompressConfig = FloatCodecConfig;

enum class FloatDecompressError : uint32_t { None = 0, ChecksumMismatch = 1 };

struct FloatDecompressStatus {
  inline FloatDecompressStatus() : error(FloatDecompressError::None) {}
  FloatDecompressError error;
  std::vector<std::pair<int, std::string>> errorInfo;
};

inline void floatCompress(StackDeviceMemory& res, const FloatCompressConfig& config, uint32_t numInBatch,
                          const void** in, const uint32_t* inSize, void** out, uint32_t* outSize_dev,
                          cudaStream_t stream) {
  detail::check(dgb_float_compress(res.data(), res.size(), (int)config.floatType, config.ansConfig.probBits,
                                   config.useChecksum, numInBatch, in, inSize, out, outSize_dev, stream),
                "floatCompress");
}

inline void floatCompressSplitSize(StackDeviceMemory& res, const FloatCompressConfig& config, uint32_t numInBatch,
                                   const void* in_dev, const uint32_t* inSplitSizes, void* out_dev, uint32_t outStride,
                                   uint32_t* outSize_dev, cudaStream_t stream) {
  detail::check(dgb_float_compress_split_size(res.data(), res.size(), (int)config.floatType,
                                              config.ansConfig.probBits, config.useChecksum, numInBatch, in_dev,
                                              inSplitSizes, out_dev, outStride, outSize_dev, stream),
                "floatCompressSplitSize");
}

inline void floatCompressSparse(StackDeviceMemory& res, const FloatCompressConfig& config, uint32_t numInBatch,
                                const void** in, const uint32_t* inSize, void** out, uint32_t* outSize_dev,
                                cudaStream_t stream) {
  detail::check(dgb_float_compress_sparse(res.data(), res.size(), (int)config.floatType, config.ansConfig.probBits,
                                          config.useChecksum, numInBatch, in, inSize, out, outSize_dev, stream),
                "floatCompressSparse");
}

inline FloatDecompressStatus floatDecompress(StackDeviceMemory& res, const FloatDecompressConfig& config,
                                             uint32_t numInBatch, const void** in, void** out,
                                             const uint32_t* outCapacity, uint8_t* outSuccess_dev,
                                             uint32_t* outSize_dev, cudaStream_t stream) {
  std::vector<uint32_t> idx(numInBatch);
  uint32_t n = 0;
  int rc = dgb_float_decompress(res.data(), res.size(), (int)config.floatType, config.ansConfig.probBits,
                                config.useChecksum, numInBatch, in, out, outCapacity, outSuccess_dev, outSize_dev,
                                stream, idx.data(), numInBatch, &n);
  return detail::makeStatus<FloatDecompressStatus, FloatDecompressError>(rc, idx, n, "floatDecompress");
}

inline FloatDecompressStatus floatDecompressSplitSize(StackDeviceMemory& res, const FloatDecompressConfig& config,
                                                      uint32_t numInBatch, const void** in, void* out_dev,
                                                      const uint32_t* outSplitSizes, uint8_t* outSuccess_dev,
                                                      uint32_t* outSize_dev, cudaStream_t stream) {
  std::vector<uint32_t> idx(numInBatch);
  uint32_t n = 0;
  int rc = dgb_float_decompress_split_size(res.data(), res.size(), (int)config.floatType, config.ansConfig.probBits,
                                           config.useChecksum, numInBatch, in, out_dev, outSplitSizes, outSuccess_dev,
                                           outSize_dev, stream, idx.data(), numInBatch, &n);
  return detail::makeStatus<FloatDecompressStatus, FloatDecompressError>(rc, idx, n, "floatDecompressSplitSize");
}

inline FloatDecompressStatus floatDecompressSparse(StackDeviceMemory& res, const FloatDecompressConfig& config,
                                                   uint32_t numInBatch, const void** in, void** out,
                                                   const uint32_t* outCapacity, uint8_t* outSuccess_dev,
                                                   uint32_t* outSize_dev, cudaStream_t stream) {
  detail::check(dgb_float_decompress_sparse(res.data(), res.size(), (int)config.floatType, config.ansConfig.probBits,
                                            config.useChecksum, numInBatch, in, out, outCapacity, outSuccess_dev,
                                            outSize_dev, stream),
                "floatDecompressSparse");
  return FloatDecompressStatus();
}

inline void floatGetCompressedInfo(StackDeviceMemory& res, const void** in, uint32_t numInBatch,
                                   uint32_t* outSizes_dev, uint32_t* outTypes_dev, uint32_t* outChecksum_dev,
                                   cudaStream_t stream) {
  detail::check(dgb_float_get_compressed_info(in, 0, numInBatch, outSizes_dev, outTypes_dev, outChecksum_dev,
                                              res.data(), res.size(), stream),
                "floatGetCompressedInfo");
}

inline void floatGetCompressedInfoDevice(StackDeviceMemory& res, const void** in_dev, uint32_t numInBatch,
                                         uint32_t* outSizes_dev, uint32_t* outTypes_dev, uint32_t* outChecksum_dev,
                                         cudaStream_t stream) {
  detail::check(dgb_float_get_compressed_info(in_dev, 1, numInBatch, outSizes_dev, outTypes_dev, outChecksum_dev,
                                              res.data(), res.size(), stream),
                "floatGetCompressedInfoDevice");
}

} // namespace dietgpu
