// Host-buffer entry points of the C ABI (include/dietgpu_b200.h, "host buffers"): what a caller with HOST tensors
// has to do around the reference's device-pointer API (cudaMemcpy in, floatCompress / ansEncodeBatchPointer,
// cudaMemcpy the sizes, cudaMemcpy out; dietgpu/float/GpuFloatCodec.h:105-140, :190-224,
// dietgpu/ans/GpuANSCodec.h:157-176, :259-279), folded into one call and pipelined over groups of batch elements.
//
// The batch is cut into groups of consecutive elements (about 16 MiB of input each: measured on B200 with bf16
// 128 x 4 MiB, 4 / 8 / 16 / 32 / 64 MiB groups give 26.7 / 25.2 / 24.2 / 26.7 / 27.2 ms per round trip; small groups
// keep the fill and drain of the pipeline short, the kernels are never the bottleneck). Group g runs on lane g % 3; a lane is a stream with its own device staging
// (inputs, archive rows, scratch). On its lane a group is strictly ordered  H2D -> codec -> D2H, across lanes nothing
// is: while group g's result leaves through the device-to-host copy engine, group g + 1's input arrives through the
// host-to-device engine and the SMs work on whichever group has its data. PCIe is full duplex, so a step costs about
// max(bytes in, bytes out) / link rate instead of their sum.
//
// Compression has one host dependency: the exact archive sizes exist only on the device, so each group's sizes come
// back first (one small D2H into pinned memory + an event); the host waits for that event - the two younger groups
// are already enqueued on the other lanes - and then enqueues the exact-size archive copies.
// The codec work itself goes through the public device-pointer entry points of this library, unchanged.
#include <stdlib.h>
#include <string.h>

#include <mutex>
#include <vector>

#include "../../include/dietgpu_b200.h"
#include "internal.h"
#include "knobs.h"

namespace dgb {
namespace {

constexpr int kPipeLanes = 4; // lanes that exist; lanesInUse() of them take groups
int lanesInUse() { // DGB_HOST_LANES is an experiment knob
  int v = DGB_KNOB("DGB_HOST_LANES", 3);
  if (v < 2) v = 2;
  if (v > kPipeLanes) v = kPipeLanes;
  return v;
}
// input bytes per group (at least one element); DGB_HOST_GROUP_MB is an experiment knob
size_t groupBytes() {
  const int mb = DGB_KNOB("DGB_HOST_GROUP_MB", 16);
  return (size_t)(mb > 0 ? mb : 16) << 20;
}

struct Lane {
  cudaStream_t s = nullptr;
  cudaEvent_t sizesReady = nullptr;
  uint8_t* dev = nullptr;
  size_t devCap = 0;
  uint8_t* pinned = nullptr; // sizes / success flags
  size_t pinnedCap = 0;
};

struct Pipe {
  Lane lane[kPipeLanes];
  bool ready = false;
};

// one pipeline and one lock per device: calls on different GPUs (one thread each) do not serialise each other; two
// calls on the same GPU share its lanes and take turns
std::mutex g_mutex[kMaxDevices];
Pipe g_pipe[kMaxDevices];

int lanePrepare(Lane& l, size_t devBytes, size_t pinnedBytes) {
  if (!l.s) {
    DGB_CUDA(cudaStreamCreateWithFlags(&l.s, cudaStreamNonBlocking));
    DGB_CUDA(cudaEventCreateWithFlags(&l.sizesReady, cudaEventDisableTiming));
  }
  if (devBytes > l.devCap) {
    if (l.dev) {
      DGB_CUDA(cudaStreamSynchronize(l.s));
      DGB_CUDA(cudaFree(l.dev));
      l.dev = nullptr;
      l.devCap = 0;
    }
    DGB_CUDA(cudaMalloc(&l.dev, devBytes));
    l.devCap = devBytes;
  }
  if (pinnedBytes > l.pinnedCap) {
    if (l.pinned) {
      DGB_CUDA(cudaStreamSynchronize(l.s));
      DGB_CUDA(cudaFreeHost(l.pinned));
      l.pinned = nullptr;
      l.pinnedCap = 0;
    }
    DGB_CUDA(cudaMallocHost(&l.pinned, pinnedBytes));
    l.pinnedCap = pinnedBytes;
  }
  return kOk;
}

struct Group {
  uint32_t e0, e1;
};

// consecutive elements until the group holds kGroupBytes of input
std::vector<Group> makeGroups(uint32_t num, const uint32_t* units, size_t unitBytes) {
  std::vector<Group> g;
  uint32_t e0 = 0;
  size_t acc = 0;
  for (uint32_t i = 0; i < num; ++i) {
    acc += (size_t)units[i] * unitBytes;
    if (acc >= groupBytes() || i + 1 == num || i + 1 - e0 >= 4096) {
      g.push_back(Group{e0, i + 1});
      e0 = i + 1;
      acc = 0;
    }
  }
  return g;
}

size_t unitBytesOf(int ft) { return ft == kFtNone ? 1 : floatWordBytes(ft); }
uint32_t maxCompressed(int ft, uint32_t n) {
  return ft == kFtNone ? dgb_ans_max_compressed_size(n) : dgb_float_max_compressed_size(ft, n);
}

// Device staging of one group: [ins | outs | sizes u32 | success u8 | temp]
struct Layout {
  std::vector<size_t> inOff, outOff;
  size_t sizesOff = 0, successOff = 0, tempOff = 0, tempBytes = 0, total = 0;
};

Layout layoutGroup(const Group& g, const std::vector<size_t>& inBytes, const std::vector<size_t>& outBytes, int op,
                   int ft, uint32_t maxUnits) {
  Layout L;
  size_t off = 0;
  for (uint32_t i = g.e0; i < g.e1; ++i) {
    L.inOff.push_back(off);
    off += roundUpSz(inBytes[i], 16);
  }
  off = roundUpSz(off, 256);
  for (uint32_t i = g.e0; i < g.e1; ++i) {
    L.outOff.push_back(off);
    off += roundUpSz(outBytes[i], 16);
  }
  off = roundUpSz(off, 256);
  const uint32_t cnt = g.e1 - g.e0;
  L.sizesOff = off;
  off += roundUpSz((size_t)cnt * 4, 256);
  L.successOff = off;
  off += roundUpSz(cnt, 256);
  L.tempOff = off;
  L.tempBytes = dgb_temp_bytes_required(op, ft, cnt, maxUnits) + 4096;
  L.total = off + L.tempBytes;
  return L;
}

// Copies of one group in one direction; neighbours that are contiguous on both sides leave as one transfer
// (a batch that is one host tensor, as in torch's compress_data, then moves in group-sized pieces). Two host
// buffers can be adjacent without belonging to one allocation; CUDA rejects a copy spanning both, and the
// pieces then go one by one.
struct CopyList {
  struct Item {
    uint8_t* dst;
    const uint8_t* src;
    size_t bytes;
  };
  std::vector<Item> pieces;
  void add(void* dst, const void* src, size_t bytes) {
    if (bytes) pieces.push_back(Item{(uint8_t*)dst, (const uint8_t*)src, bytes});
  }
  int issue(cudaMemcpyKind kind, cudaStream_t s) {
    size_t i = 0;
    while (i < pieces.size()) {
      size_t j = i + 1;
      size_t bytes = pieces[i].bytes;
      while (j < pieces.size() && pieces[j].dst == pieces[i].dst + bytes && pieces[j].src == pieces[i].src + bytes)
        bytes += pieces[j++].bytes;
      if (j > i + 1 && cudaMemcpyAsync(pieces[i].dst, pieces[i].src, bytes, kind, s) != cudaSuccess) {
        (void)cudaGetLastError();
        j = i + 1; // not one allocation: this piece alone, the rest is tried again from the next one
      }
      if (j == i + 1) DGB_CUDA(cudaMemcpyAsync(pieces[i].dst, pieces[i].src, pieces[i].bytes, kind, s));
      i = j;
    }
    pieces.clear();
    return kOk;
  }
};

struct EntryGuard { // the caller's stream orders the call: lanes start after it, the call returns when they are done
  int dev = 0;
  Pipe* pipe = nullptr;
  std::unique_lock<std::mutex> lock;
  int begin(cudaStream_t user) {
    DGB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return kErrInvalidArg;
    lock = std::unique_lock<std::mutex>(g_mutex[dev]);
    pipe = &g_pipe[dev];
    for (int k = 0; k < kPipeLanes; ++k) {
      int rc = lanePrepare(pipe->lane[k], 0, 0);
      if (rc) return rc;
    }
    cudaEvent_t ev;
    DGB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    DGB_CUDA(cudaEventRecord(ev, user));
    for (int k = 0; k < kPipeLanes; ++k) DGB_CUDA(cudaStreamWaitEvent(pipe->lane[k].s, ev, 0));
    DGB_CUDA(cudaEventDestroy(ev));
    return kOk;
  }
  int finish() {
    for (int k = 0; k < kPipeLanes; ++k) DGB_CUDA(cudaStreamSynchronize(pipe->lane[k].s));
    return kOk;
  }
};

// ft == kFtNone: raw ANS (sizes in bytes); else the float codec (sizes in words)
int compressHost(int ft, int pb, int useChecksum, uint32_t num, const void* const* in, const uint32_t* inSize,
                 void* const* out, const uint32_t* outCapacity, uint32_t* outSize, cudaStream_t user) {
  if (num == 0) return kOk;
  if (!in || !inSize || !out || !outSize) return kErrInvalidArg;
  if (ft != kFtNone && (ft < kFtF16 || ft > kFtF64)) return kErrInvalidArg;
  if (pb < 9 || pb > 11) return kErrInvalidArg;
  const size_t ub = unitBytesOf(ft);
  std::vector<size_t> inBytes(num), outBytes(num);
  for (uint32_t i = 0; i < num; ++i) {
    inBytes[i] = (size_t)inSize[i] * ub;
    outBytes[i] = maxCompressed(ft, inSize[i]);
  }
  EntryGuard guard;
  int rc = guard.begin(user);
  if (rc) return rc;
  const std::vector<Group> groups = makeGroups(num, inSize, ub);
  const int G = (int)groups.size();
  const int lanes = lanesInUse();
  std::vector<Layout> layouts(G);
  int status = kOk;

  auto phase1 = [&](int g) -> int { // inputs in, codec, sizes out
    const Group& gr = groups[g];
    const uint32_t cnt = gr.e1 - gr.e0;
    uint32_t maxUnits = 0;
    for (uint32_t i = gr.e0; i < gr.e1; ++i) maxUnits = inSize[i] > maxUnits ? inSize[i] : maxUnits;
    Layout& L = layouts[g];
    L = layoutGroup(gr, inBytes, outBytes, ft == kFtNone ? 0 : 2, ft, maxUnits);
    Lane& l = guard.pipe->lane[g % lanes];
    int r = lanePrepare(l, L.total, (size_t)cnt * 8);
    if (r) return r;
    std::vector<const void*> inDev(cnt);
    std::vector<void*> outDev(cnt);
    CopyList cl;
    for (uint32_t j = 0; j < cnt; ++j) {
      inDev[j] = l.dev + L.inOff[j];
      outDev[j] = l.dev + L.outOff[j];
      cl.add(l.dev + L.inOff[j], in[gr.e0 + j], inBytes[gr.e0 + j]);
    }
    if ((r = cl.issue(cudaMemcpyHostToDevice, l.s))) return r;
    uint32_t* sizesDev = (uint32_t*)(l.dev + L.sizesOff);
    if (ft == kFtNone)
      r = dgb_ans_encode_batch_pointer(l.dev + L.tempOff, L.tempBytes, pb, useChecksum, cnt, inDev.data(),
                                       inSize + gr.e0, nullptr, outDev.data(), sizesDev, l.s);
    else
      r = dgb_float_compress(l.dev + L.tempOff, L.tempBytes, ft, pb, useChecksum, cnt, inDev.data(), inSize + gr.e0,
                             outDev.data(), sizesDev, l.s);
    if (r) return r;
    DGB_CUDA(cudaMemcpyAsync(l.pinned, sizesDev, (size_t)cnt * 4, cudaMemcpyDeviceToHost, l.s));
    DGB_CUDA(cudaEventRecord(l.sizesReady, l.s));
    return kOk;
  };
  auto phase2 = [&](int g) -> int { // exact-size archive copies once the sizes are on the host
    const Group& gr = groups[g];
    const uint32_t cnt = gr.e1 - gr.e0;
    Lane& l = guard.pipe->lane[g % lanes];
    const Layout& L = layouts[g];
    DGB_CUDA(cudaEventSynchronize(l.sizesReady));
    const uint32_t* sz = (const uint32_t*)l.pinned;
    for (uint32_t j = 0; j < cnt; ++j) {
      const uint32_t bytes = sz[j];
      outSize[gr.e0 + j] = bytes;
      if (outCapacity && bytes > outCapacity[gr.e0 + j]) {
        status = kErrInvalidArg; // archive does not fit the caller's buffer: size reported, nothing copied
        continue;
      }
      DGB_CUDA(cudaMemcpyAsync(out[gr.e0 + j], l.dev + L.outOff[j], bytes, cudaMemcpyDeviceToHost, l.s));
    }
    return kOk;
  };
  // group g's phase 1 reuses the lane of group g - lanes, whose phase 2 was enqueued one iteration earlier
  for (int g = 0; g < G + lanes - 1; ++g) {
    if (g < G && (rc = phase1(g))) break;
    const int done = g - (lanes - 1);
    if (done >= 0 && done < G && (rc = phase2(done))) break;
  }
  const int rcSync = guard.finish();
  return rc ? rc : rcSync ? rcSync : status;
}

// Length an archive in HOST memory claims for itself (headers only; dietgpu/ans/GpuANSUtils.cuh:68-86,
// dietgpu/float/GpuFloatUtils.cuh:19-96), or 0 when a header it needs lies beyond `avail` bytes or is not sane. The
// device decoder trusts the header's word count for the extent of what it reads, so a caller-visible size is checked
// against it here, before anything is copied (a wrong magic is left to the decoder, which reports success = 0).
static uint64_t claimedArchiveBytes(int ft, const void* archive, uint64_t avail) {
  const uint8_t* a = (const uint8_t*)archive;
  uint64_t off = 0;
  if (ft != kFtNone) {
    if (avail < kFloatHeaderBytes) return 0;
    uint32_t h[5];
    memcpy(h, a, sizeof(h));
    if (h[0] != kFloatMagic || (int)(h[2] & 0xfu) != ft) return avail; // not ours to size: the decoder rejects it
    off = (uint64_t)kFloatHeaderBytes + floatRawBytes(ft, h[1]);
    if (ft == kFtF64) {
      if (off + kAnsHeaderBytes > avail) return 0;
      uint32_t ah[4];
      memcpy(ah, a + off, sizeof(ah));
      if (ah[0] != kAnsMagic) return avail;
      if (h[4] != (uint64_t)ansOverhead(ah[1]) + 2ull * ah[3]) return 0; // firstCompSegmentBytes must be that archive's length
      off += h[4];
    }
  }
  if (off + kAnsHeaderBytes > avail) return 0;
  uint32_t ah[4];
  memcpy(ah, a + off, sizeof(ah));
  if (ah[0] != kAnsMagic) return avail;
  return off + ansOverhead(ah[1]) + 2ull * ah[3];
}

int decompressHost(int ft, int pb, int useChecksum, uint32_t num, const void* const* in, const uint32_t* inBytesArg,
                   void* const* out, const uint32_t* outCapacity, uint8_t* outSuccess, uint32_t* outSize,
                   cudaStream_t user) {
  if (num == 0) return kOk;
  if (!in || !inBytesArg || !out || !outCapacity) return kErrInvalidArg;
  if (ft != kFtNone && (ft < kFtF16 || ft > kFtF64)) return kErrInvalidArg;
  if (pb < 9 || pb > 11) return kErrInvalidArg;
  const size_t ub = unitBytesOf(ft);
  std::vector<size_t> inBytes(num), outBytes(num);
  for (uint32_t i = 0; i < num; ++i) {
    if (inBytesArg[i] < (ft == kFtNone ? kAnsHeaderBytes : kFloatHeaderBytes)) return kErrInvalidArg;
    const uint64_t claimed = claimedArchiveBytes(ft, in[i], inBytesArg[i]);
    if (claimed == 0 || claimed > inBytesArg[i]) {
      fprintf(stderr, "dietgpu_b200: host-buffer decode: archive %u claims %llu bytes but in_bytes is %u\n", i,
              (unsigned long long)claimed, inBytesArg[i]);
      return kErrInvalidArg;
    }
    inBytes[i] = inBytesArg[i];
    outBytes[i] = (size_t)outCapacity[i] * ub;
  }
  EntryGuard guard;
  int rc = guard.begin(user);
  if (rc) return rc;
  const std::vector<Group> groups = makeGroups(num, outCapacity, ub);
  const int G = (int)groups.size();
  const int lanes = lanesInUse();
  int status = kOk;
  struct Pending {
    uint32_t e0, cnt;
    int laneIdx;
  };
  auto collect = [&](const Pending& p) -> int { // flags and sizes of a finished group -> caller's arrays
    Lane& l = guard.pipe->lane[p.laneIdx];
    DGB_CUDA(cudaEventSynchronize(l.sizesReady));
    if (outSize) memcpy(outSize + p.e0, l.pinned, (size_t)p.cnt * 4);
    if (outSuccess) memcpy(outSuccess + p.e0, l.pinned + (size_t)p.cnt * 4, p.cnt);
    return kOk;
  };
  std::vector<Pending> pending;
  for (int g = 0; g < G; ++g) {
    const Group& gr = groups[g];
    const uint32_t cnt = gr.e1 - gr.e0;
    const int laneIdx = g % lanes;
    if (g >= lanes && (rc = collect(pending[g - lanes]))) break; // the lane's pinned words are about to be reused
    uint32_t maxUnits = 0;
    for (uint32_t i = gr.e0; i < gr.e1; ++i) maxUnits = outCapacity[i] > maxUnits ? outCapacity[i] : maxUnits;
    const Layout L = layoutGroup(gr, inBytes, outBytes, ft == kFtNone ? 1 : 3, ft, maxUnits);
    Lane& l = guard.pipe->lane[laneIdx];
    if ((rc = lanePrepare(l, L.total, (size_t)cnt * 8))) break;
    std::vector<const void*> inDev(cnt);
    std::vector<void*> outDev(cnt);
    CopyList cl;
    for (uint32_t j = 0; j < cnt; ++j) {
      inDev[j] = l.dev + L.inOff[j];
      outDev[j] = l.dev + L.outOff[j];
      cl.add(l.dev + L.inOff[j], in[gr.e0 + j], inBytes[gr.e0 + j]);
    }
    if ((rc = cl.issue(cudaMemcpyHostToDevice, l.s))) break;
    uint32_t* sizesDev = (uint32_t*)(l.dev + L.sizesOff);
    uint8_t* successDev = l.dev + L.successOff;
    int r;
    if (ft == kFtNone)
      r = dgb_ans_decode_batch_pointer(l.dev + L.tempOff, L.tempBytes, pb, useChecksum, cnt, inDev.data(),
                                       outDev.data(), outCapacity + gr.e0, successDev, sizesDev, l.s, nullptr, 0, nullptr);
    else
      r = dgb_float_decompress(l.dev + L.tempOff, L.tempBytes, ft, pb, useChecksum, cnt, inDev.data(), outDev.data(),
                               outCapacity + gr.e0, successDev, sizesDev, l.s, nullptr, 0, nullptr);
    if (r == kChecksumMismatch) status = kChecksumMismatch;
    else if (r) { rc = r; break; }
    // decoded length from the archive header, which is host memory here (float: words at [1], ANS: bytes at [2]);
    // an element the device rejects (bad magic, capacity) reports success 0 and its bytes are not meaningful
    for (uint32_t j = 0; j < cnt; ++j) {
      const uint32_t* h = (const uint32_t*)in[gr.e0 + j];
      uint32_t n = ft == kFtNone ? h[2] : h[1];
      if (n > outCapacity[gr.e0 + j]) n = 0;
      cl.add(out[gr.e0 + j], l.dev + L.outOff[j], (size_t)n * ub);
    }
    if ((rc = cl.issue(cudaMemcpyDeviceToHost, l.s))) break;
    DGB_CUDA(cudaMemcpyAsync(l.pinned, sizesDev, (size_t)cnt * 4, cudaMemcpyDeviceToHost, l.s));
    DGB_CUDA(cudaMemcpyAsync(l.pinned + (size_t)cnt * 4, successDev, cnt, cudaMemcpyDeviceToHost, l.s));
    DGB_CUDA(cudaEventRecord(l.sizesReady, l.s));
    pending.push_back(Pending{gr.e0, cnt, laneIdx});
  }
  if (!rc)
    for (int g = G > lanes ? G - lanes : 0; g < (int)pending.size(); ++g)
      if ((rc = collect(pending[g]))) break;
  const int rcSync = guard.finish();
  return rc ? rc : rcSync ? rcSync : status;
}

} // namespace
} // namespace dgb

using namespace dgb;

extern "C" {

int dgb_float_compress_host(int ft, int pb, int cs, uint32_t num, const void* const* in, const uint32_t* inSize,
                            void* const* out, const uint32_t* outCapacity, uint32_t* outSize, void* stream) {
  if (ft < kFtF16 || ft > kFtF64) return kErrInvalidArg;
  return compressHost(ft, pb, cs, num, in, inSize, out, outCapacity, outSize, (cudaStream_t)stream);
}

int dgb_float_decompress_host(int ft, int pb, int cs, uint32_t num, const void* const* in, const uint32_t* inBytes,
                              void* const* out, const uint32_t* outCapacity, uint8_t* outSuccess, uint32_t* outSize,
                              void* stream) {
  if (ft < kFtF16 || ft > kFtF64) return kErrInvalidArg;
  return decompressHost(ft, pb, cs, num, in, inBytes, out, outCapacity, outSuccess, outSize, (cudaStream_t)stream);
}

int dgb_ans_encode_host(int pb, int cs, uint32_t num, const void* const* in, const uint32_t* inSize, void* const* out,
                        const uint32_t* outCapacity, uint32_t* outSize, void* stream) {
  return compressHost(kFtNone, pb, cs, num, in, inSize, out, outCapacity, outSize, (cudaStream_t)stream);
}

int dgb_ans_decode_host(int pb, int cs, uint32_t num, const void* const* in, const uint32_t* inBytes, void* const* out,
                        const uint32_t* outCapacity, uint8_t* outSuccess, uint32_t* outSize, void* stream) {
  return decompressHost(kFtNone, pb, cs, num, in, inBytes, out, outCapacity, outSuccess, outSize, (cudaStream_t)stream);
}

void dgb_host_release(void) {
  int cur = 0;
  cudaGetDevice(&cur);
  for (int d = 0; d < kMaxDevices; ++d) {
    std::lock_guard<std::mutex> lock(g_mutex[d]);
    Pipe& p = g_pipe[d];
    bool any = false;
    for (int k = 0; k < kPipeLanes; ++k) any = any || p.lane[k].s;
    if (!any) continue;
    cudaSetDevice(d);
    for (int k = 0; k < kPipeLanes; ++k) {
      Lane& l = p.lane[k];
      if (l.s) cudaStreamSynchronize(l.s);
      if (l.dev) cudaFree(l.dev);
      if (l.pinned) cudaFreeHost(l.pinned);
      if (l.sizesReady) cudaEventDestroy(l.sizesReady);
      if (l.s) cudaStreamDestroy(l.s);
      l = Lane{};
    }
  }
  cudaSetDevice(cur);
}

} // extern "C"
