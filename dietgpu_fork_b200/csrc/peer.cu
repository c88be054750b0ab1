// Peer-visible device buffers for the compressed collectives (include/dietgpu_b200.h, "peer buffers"): one
// cudaMalloc'd archive buffer per rank, exported with CUDA IPC and mapped by every other rank of the node, so that
// the decoder reads a peer's archives straight over NVLink / NVSwitch (its compressed-word ring and raw-section
// staging are asynchronous copies from global memory, local or peer alike). The exchange of compressed bytes is
// thereby fused into the decompress kernel: no staging copy, no size exchange (archives are self-describing,
// dietgpu/ans/GpuANSUtils.cuh header), nothing on the wire but the bytes the decoder consumes.
// The reference has no multi-GPU code; its README names compressed collectives as the purpose (README.md:92-96).
#include <string.h>

#include "../../include/dietgpu_b200.h"
#include "internal.h"
#include "prof.h"

using namespace dgb;

namespace dgb {

// ---------------------------------------------------------------------------------------------
// Pull: copies float archives out of peer memory into local memory at streaming speed, taking each archive's
// length from its own headers on the device - still no size exchange and no host synchronisation. Measured
// (round 1, scripts/peer_breakdown.py): the decoder reading a peer's archive in place gets 325 GB/s, because every
// 4096-symbol block starts with dependent reads (states, index entry, first ring piece) that each cost an NVLink
// round trip; a plain copy with thousands of independent 16-byte loads in flight runs at link speed, and the local
// decode that follows runs at its usual 1.9 TB/s. dist.py pipelines the two source by source (pull of peer k + 1 on a
// copy stream while peer k's archive is being decoded).
// Archive length (dietgpu/float/GpuFloatUtils.cuh:19-96, dietgpu/ans/GpuANSUtils.cuh:68-86):
//   32 + rawBytes(type, n) + [fp64: firstCompSegmentBytes +] ansOverhead(nb) + 2 * totalCompressedWords of the last ANS
// A source whose headers are not sane, or whose length exceeds the destination's capacity, is copied as its 32-byte
// float header with the magic cleared: the decode that follows then reports that element as failed.
// ---------------------------------------------------------------------------------------------
struct PullArgs {
  const void* const* src; // device array [num]: archive in peer (or local) memory, 16-byte aligned
  void* const* dst;       // device array [num]: local destination, 16-byte aligned
  const uint32_t* cap;    // device array [num]: destination capacity in bytes
  uint32_t num;
};

__global__ void __launch_bounds__(256) k_peerPull(const PullArgs a) {
  __shared__ uint32_t sBytes;
  const uint32_t b = blockIdx.y, tid = threadIdx.x;
  const uint8_t* src = (const uint8_t*)a.src[b];
  uint8_t* dst = (uint8_t*)a.dst[b];
  if (tid == 0) {
    const uint4 h0 = *(const uint4*)src;        // magic | n | type | checksum
    const uint4 h1 = *(const uint4*)(src + 16); // firstCompSegmentBytes
    uint32_t bytes = 0;
    const int ft = (int)(h0.z & 0xfu);
    if (h0.x == kFloatMagic && ft >= kFtF16 && ft <= kFtF64) {
      uint64_t off = (uint64_t)kFloatHeaderBytes + floatRawBytes(ft, h0.y);
      if (ft == kFtF64) off += h1.x;
      if (off + kAnsFixedBytes <= a.cap[b]) {
        const uint4 ah = *(const uint4*)(src + off); // magic | numBlocks | n | totalCompressedWords
        const uint64_t total = off + ansOverhead(ah.y) + 2ull * ah.w;
        if (ah.x == kAnsMagic && ah.y == divUp(h0.y, kBlockSyms) && total <= a.cap[b]) bytes = (uint32_t)total;
      }
    }
    sBytes = bytes;
  }
  __syncthreads();
  const uint32_t bytes = sBytes;
  if (bytes == 0) { // not copyable: leave a header the decoder rejects
    if (blockIdx.x == 0 && tid == 0 && a.cap[b] >= 32) {
      ((uint4*)dst)[0] = make_uint4(0u, 0u, 0u, 0u);
      ((uint4*)dst)[1] = make_uint4(0u, 0u, 0u, 0u);
    }
    return;
  }
  const uint4* s4 = (const uint4*)src;
  uint4* d4 = (uint4*)dst;
  const uint32_t nv = bytes >> 4; // archive lengths are multiples of 16
  const uint32_t stride = gridDim.x * 256u;
  uint32_t i = blockIdx.x * 256u + tid;
  for (; i + 7 * stride < nv; i += 8 * stride) { // 8 independent 16-byte loads per thread in flight
    uint4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = s4[i + k * stride];
#pragma unroll
    for (int k = 0; k < 8; ++k) d4[i + k * stride] = v[k];
  }
  for (; i < nv; i += stride) d4[i] = s4[i];
}

} // namespace dgb

extern "C" {

int dgb_peer_pull_float(uint32_t num, const void* const* src_dev, void* const* dst_dev, const uint32_t* dst_capacity_dev,
                        uint32_t max_bytes, void* stream) {
  if (num == 0) return kOk;
  if (!src_dev || !dst_dev || !dst_capacity_dev) return kErrInvalidArg;
  cudaStream_t s = (cudaStream_t)stream;
  PullArgs a{src_dev, dst_dev, dst_capacity_dev, num};
  // enough CTAs to keep the link busy, shared among the archives of the call; sized from the caller's bound
  uint32_t per = (uint32_t)smCount() * 4u / num;
  const uint32_t useful = max_bytes / (16u * 256u * 8u) + 1u; // one 8-deep iteration per thread at least
  if (per > useful) per = useful;
  if (per < 1) per = 1;
  {
    ProfScope ps__("k_peerPull", s);
    k_peerPull<<<dim3(per, num), 256, 0, s>>>(a);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

int dgb_peer_alloc(size_t bytes, void** devPtr, void* handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == DGB_PEER_HANDLE_BYTES, "IPC handle size");
  if (!devPtr || !handle64 || bytes == 0) return kErrInvalidArg;
  void* p = nullptr;
  DGB_CUDA(cudaMalloc(&p, bytes));
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, p);
  if (e != cudaSuccess) {
    fprintf(stderr, "dietgpu_b200: cudaIpcGetMemHandle: %s\n", cudaGetErrorString(e));
    cudaFree(p);
    return kErrCuda;
  }
  memcpy(handle64, &h, sizeof(h));
  *devPtr = p;
  return kOk;
}

int dgb_peer_open(const void* handle64, void** peerPtr) {
  if (!handle64 || !peerPtr) return kErrInvalidArg;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, sizeof(h));
  DGB_CUDA(cudaIpcOpenMemHandle(peerPtr, h, cudaIpcMemLazyEnablePeerAccess));
  return kOk;
}

int dgb_peer_close(void* peerPtr) {
  if (!peerPtr) return kOk;
  DGB_CUDA(cudaIpcCloseMemHandle(peerPtr));
  return kOk;
}

int dgb_peer_free(void* devPtr) {
  if (!devPtr) return kOk;
  DGB_CUDA(cudaFree(devPtr));
  return kOk;
}

} // extern "C"
