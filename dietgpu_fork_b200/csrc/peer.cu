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

using namespace dgb;

extern "C" {

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
