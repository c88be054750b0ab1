// Reference-style C++ caller (namespace dietgpu) over include/dietgpu_b200.hpp: the batch round trips of
// /root/reference/dietgpu/ans/ANSTest.cu:98-200 and /root/reference/dietgpu/float/FloatTest.cu:60-190,
// written against the adaptor so that a passing run shows the reference's C++ call sites keep compiling
// and working. Exit code 0 = all round trips bit-exact.
#include <cmath>
#include <cstdio>
#include <cstring>
#include <random>

#include "dietgpu_b200.hpp"

using namespace dietgpu;

#define CK(x)                                                                  \
  do {                                                                         \
    cudaError_t e__ = (x);                                                     \
    if (e__ != cudaSuccess) {                                                  \
      std::fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e__)); \
      return 2;                                                                \
    }                                                                          \
  } while (0)

static int ansRound(StackDeviceMemory& res, cudaStream_t stream, int probBits, const std::vector<uint32_t>& sizes) {
  std::mt19937 gen(10 + probBits);
  std::exponential_distribution<double> dist(20.0 / 256.0);
  uint32_t n = sizes.size();
  std::vector<void*> in(n), enc(n), dec(n);
  std::vector<std::vector<uint8_t>> host(n);
  std::vector<uint32_t> cap(n);
  for (uint32_t i = 0; i < n; ++i) {
    host[i].resize(sizes[i]);
    for (auto& b : host[i]) b = (uint8_t)std::min(255.0, dist(gen));
    CK(cudaMalloc(&in[i], sizes[i] + 16));
    CK(cudaMalloc(&enc[i], getMaxCompressedSize(sizes[i])));
    CK(cudaMalloc(&dec[i], sizes[i] + 16));
    CK(cudaMemcpy(in[i], host[i].data(), sizes[i], cudaMemcpyHostToDevice));
    cap[i] = sizes[i];
  }
  auto encSize = res.alloc<uint32_t>(stream, n);
  auto outSize = res.alloc<uint32_t>(stream, n);
  auto outOk = res.alloc<uint8_t>(stream, n);
  ANSCodecConfig cfg(probBits, true);
  ansEncodeBatchPointer(res, cfg, n, (const void**)in.data(), sizes.data(), nullptr, enc.data(), encSize.data(), stream);
  auto st = ansDecodeBatchPointer(res, cfg, n, (const void**)enc.data(), dec.data(), cap.data(), outOk.data(),
                                  outSize.data(), stream);
  if (st.error != ANSDecodeError::None) return 3;
  auto ok = outOk.copyToHost(stream);
  auto sz = outSize.copyToHost(stream);
  for (uint32_t i = 0; i < n; ++i) {
    if (!ok[i] || sz[i] != sizes[i]) return 4;
    std::vector<uint8_t> back(sizes[i]);
    CK(cudaMemcpy(back.data(), dec[i], sizes[i], cudaMemcpyDeviceToHost));
    if (std::memcmp(back.data(), host[i].data(), sizes[i]) != 0) return 5;
    cudaFree(in[i]); cudaFree(enc[i]); cudaFree(dec[i]);
  }
  return 0;
}

template <typename Word>
static int floatRound(StackDeviceMemory& res, cudaStream_t stream, FloatType ft, const std::vector<uint32_t>& sizes,
                      bool sparse) {
  std::mt19937 gen(77 + (int)ft);
  std::normal_distribution<float> dist(0.f, 0.02f);
  uint32_t n = sizes.size();
  std::vector<void*> in(n), enc(n), dec(n);
  std::vector<std::vector<Word>> host(n);
  for (uint32_t i = 0; i < n; ++i) {
    host[i].resize(sizes[i]);
    for (auto& w : host[i]) {
      float f = dist(gen);
      if (sparse && (gen() & 3)) f = 0.f;
      if (sizeof(Word) == 8) { double d = f; std::memcpy(&w, &d, 8); }
      else if (sizeof(Word) == 4) std::memcpy(&w, &f, 4);
      else { uint32_t u; std::memcpy(&u, &f, 4); w = (Word)(u >> 16); } // bf16 bit pattern, also a valid fp16 word
    }
    uint32_t maxC = sparse ? getMaxSparseFloatCompressedSize(ft, sizes[i]) : getMaxFloatCompressedSize(ft, sizes[i]);
    CK(cudaMalloc(&in[i], sizes[i] * sizeof(Word) + 16));
    CK(cudaMalloc(&enc[i], maxC));
    CK(cudaMalloc(&dec[i], sizes[i] * sizeof(Word) + 16));
    CK(cudaMemcpy(in[i], host[i].data(), sizes[i] * sizeof(Word), cudaMemcpyHostToDevice));
  }
  auto encSize = res.alloc<uint32_t>(stream, n);
  auto outSize = res.alloc<uint32_t>(stream, n);
  auto outOk = res.alloc<uint8_t>(stream, n);
  FloatCodecConfig cfg(ft, ANSCodecConfig(10, false), false, !sparse);
  FloatDecompressStatus st;
  if (sparse) {
    floatCompressSparse(res, cfg, n, (const void**)in.data(), sizes.data(), enc.data(), encSize.data(), stream);
    st = floatDecompressSparse(res, cfg, n, (const void**)enc.data(), dec.data(), sizes.data(), outOk.data(),
                               outSize.data(), stream);
  } else {
    floatCompress(res, cfg, n, (const void**)in.data(), sizes.data(), enc.data(), encSize.data(), stream);
    st = floatDecompress(res, cfg, n, (const void**)enc.data(), dec.data(), sizes.data(), outOk.data(),
                         outSize.data(), stream);
  }
  if (st.error != FloatDecompressError::None) return 6;
  auto ok = outOk.copyToHost(stream);
  auto sz = outSize.copyToHost(stream);
  auto esz = encSize.copyToHost(stream);
  for (uint32_t i = 0; i < n; ++i) {
    if (!ok[i] || sz[i] != sizes[i]) return 7;
    uint32_t maxC = sparse ? getMaxSparseFloatCompressedSize(ft, sizes[i]) : getMaxFloatCompressedSize(ft, sizes[i]);
    if (esz[i] > maxC) return 8;
    std::vector<Word> back(sizes[i]);
    CK(cudaMemcpy(back.data(), dec[i], sizes[i] * sizeof(Word), cudaMemcpyDeviceToHost));
    if (std::memcmp(back.data(), host[i].data(), sizes[i] * sizeof(Word)) != 0) return 9;
    cudaFree(in[i]); cudaFree(enc[i]); cudaFree(dec[i]);
  }
  return 0;
}

int main() {
  cudaStream_t stream;
  CK(cudaStreamCreate(&stream));
  auto res = makeStackMemory(64 << 20);
  std::vector<uint32_t> sizes = {1, 13, 4096, 4097, 100000, 1 << 20, 0, 333333};
  for (int pb : {9, 10, 11}) {
    int rc = ansRound(res, stream, pb, sizes);
    if (rc) { std::printf("ans probBits=%d FAILED rc=%d\n", pb, rc); return 1; }
  }
  int rc;
  if ((rc = floatRound<uint16_t>(res, stream, FloatType::kFloat16, sizes, false))) { std::printf("f16 rc=%d\n", rc); return 1; }
  if ((rc = floatRound<uint16_t>(res, stream, FloatType::kBFloat16, sizes, false))) { std::printf("bf16 rc=%d\n", rc); return 1; }
  if ((rc = floatRound<uint32_t>(res, stream, FloatType::kFloat32, sizes, false))) { std::printf("f32 rc=%d\n", rc); return 1; }
  if ((rc = floatRound<uint64_t>(res, stream, FloatType::kFloat64, sizes, false))) { std::printf("f64 rc=%d\n", rc); return 1; }
  std::vector<uint32_t> ssizes = {2, 13, 4096, 100000, 1 << 20};
  if ((rc = floatRound<uint16_t>(res, stream, FloatType::kBFloat16, ssizes, true))) { std::printf("sparse bf16 rc=%d\n", rc); return 1; }
  if ((rc = floatRound<uint32_t>(res, stream, FloatType::kFloat32, ssizes, true))) { std::printf("sparse f32 rc=%d\n", rc); return 1; }
  std::printf("adaptor round trips OK (scratch high water %zu bytes)\n", res.getMaxMemoryUsage());
  return 0;
}
