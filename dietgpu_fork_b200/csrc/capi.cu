// C ABI of libdietgpu_b200.so (include/dietgpu_b200.h): host-side marshalling only. Builds the
// per-element descriptor table from the caller's host arrays, ships it with one H2D copy and
// enqueues the kernels of ans_encode.cu / ans_decode.cu / float_split.cu / sparse.cu on the
// caller's stream. Mirrors the host wrappers of the reference (dietgpu/ans/GpuANSEncode.cu:27-179,
// dietgpu/ans/GpuANSDecode.cu:19-193, dietgpu/float/GpuFloatCompress.cu:50-162,
// dietgpu/float/GpuFloatDecompress.cu:22-179, dietgpu/float/GpuSparseFloat{Compress,Decompress}.cu).
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <vector>

#include "../../include/dietgpu_b200.h"
#include "internal.h"
#include "knobs.h"
#include "prof.h"

namespace dgb {

int launchSparseCompact(Temp& tmp, int ft, const Elem* elems, uint32_t numElems, uint32_t totalTiles,
                        uint8_t* dense, uint32_t* nnz, cudaStream_t s);
int launchSparseScatter(Temp& tmp, int ft, const Elem* elems, uint32_t numElems, uint32_t totalTiles,
                        const uint8_t* dense, uint8_t* outSuccess, uint32_t* outSize, cudaStream_t s);

namespace {

bool validFloatType(int ft) { return ft >= kFtF16 && ft <= kFtF64; }

uint64_t ansMaxBytes64(uint64_t n);
uint64_t floatMaxBytes64(int ft, uint64_t n);
// every element must have a size bound the 32-bit API can express (see dgb_*_max_compressed_size)
bool ansSizesOk(uint32_t num, const uint32_t* n) {
  for (uint32_t i = 0; i < num; ++i)
    if (ansMaxBytes64(n[i]) > 0x7fffffffull) return false;
  return true;
}
bool floatSizesOk(int ft, uint32_t num, const uint32_t* n, uint64_t extra = 0) {
  for (uint32_t i = 0; i < num; ++i) {
    const uint64_t f = floatMaxBytes64(ft, n[i]);
    if (f > 0xffffffffull || f + extra + n[i] / 8 > 0xffffffffull) return false;
  }
  return true;
}

// Host-side work list: tiles of `tileUnits` symbols/words per element, at least one per element.
struct WorkList {
  std::vector<Elem> elems; // numElems + 1 (sentinel)
  uint32_t totalTiles = 0;
  uint32_t totalBlocks = 0; // 4096-unit blocks, at least one per element
  uint32_t totalGroups = 0; // 256-block size groups, at least one per element
  uint64_t planeUnits = 0; // sum of roundUp(n, 16)
  uint32_t maxN = 0;
};

int buildWorkList(WorkList& wl, uint32_t num, const void* const* in, void* const* out, const uint32_t* n,
                  uint32_t tileUnits) {
  wl.elems.resize((size_t)num + 1);
  uint64_t tiles = 0, plane = 0, blocks = 0, groups = 0;
  for (uint32_t i = 0; i < num; ++i) {
    Elem& e = wl.elems[i];
    e.in = in[i];
    e.out = out[i];
    e.n = n[i];
    e.tileBase = (uint32_t)tiles;
    e.planeOff = plane;
    e.sizeBias = 0;
    e.blockBase = (uint32_t)blocks;
    e.groupBase = (uint32_t)groups;
    e.pad = 0;
    uint32_t t = divUp(n[i], tileUnits);
    tiles += t ? t : 1;
    uint32_t b = divUp(n[i], kBlockSyms);
    blocks += b ? b : 1;
    groups += divUp(b ? b : 1, 256);
    plane += roundUpSz(n[i], 16);
    if (n[i] > wl.maxN) wl.maxN = n[i];
    if (tiles > 0x7fffffffull || blocks > 0x7fffffffull) return kErrInvalidArg;
  }
  Elem& s = wl.elems[num];
  memset(&s, 0, sizeof(Elem));
  s.tileBase = (uint32_t)tiles;
  s.blockBase = (uint32_t)blocks;
  wl.totalTiles = (uint32_t)tiles;
  wl.totalBlocks = (uint32_t)blocks;
  wl.totalGroups = (uint32_t)groups;
  s.groupBase = (uint32_t)groups;
  wl.planeUnits = plane;
  return kOk;
}

WorkRange wholeRange(const WorkList& wl, uint32_t num) {
  return WorkRange{0, num, 0, wl.totalTiles, 0, wl.totalBlocks};
}

// The descriptor table is shipped with one small pageable-source cudaMemcpyAsync. The CUDA runtime stages a pageable
// source before the call returns (documented: "the function will return once the pageable buffer has been copied to
// the staging memory"), so the host vector may die after the call, and for a few KB the driver sends the data inline
// with the command stream: the first kernel starts 4-10 us sooner than behind a copy-engine DMA from pinned memory
// (round 2 tried a pinned staging ring here: bf16 32 x 2 MiB compress 0.111 -> 0.121 ms, ANS 16 x 1 MiB 62 -> 64 us;
// reverted). Consequence: these calls read the caller's host arrays at call time and are NOT capturable into CUDA
// graphs (neither are the reference's); dgb_decode_batch_noalloc ships its table as a kernel argument instead.
int upload(Temp& tmp, const WorkList& wl, Elem** dev, cudaStream_t s) {
  *dev = tmp.alloc<Elem>(wl.elems.size());
  if (tmp.failed()) return kErrCuda;
  DGB_CUDA(cudaMemcpyAsync(*dev, wl.elems.data(), wl.elems.size() * sizeof(Elem), cudaMemcpyHostToDevice, s));
  return kOk;
}

// ---- checksum verification after decode -----------------------------------------------------
__global__ void k_verifyChecksum(const Elem* __restrict__ elems, uint32_t num, int floatMode,
                                 const uint32_t* __restrict__ computed, const uint8_t* __restrict__ success,
                                 uint8_t* __restrict__ mismatch) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= num) return;
  const uint32_t* h = (const uint32_t*)elems[i].in;
  // ANS header: options at [4], checksum at [5]; float header: options at [2], checksum at [3]
  uint32_t opt = floatMode ? h[2] : h[4];
  uint32_t cs = floatMode ? h[3] : h[5];
  bool bad = false;
  if ((opt & 16u) && (!success || success[i])) bad = (cs & 0xffu) != (computed[i] & 0xffu);
  mismatch[i] = bad ? 1 : 0;
}

int verifyChecksums(Temp& tmp, const Elem* elemsDev, uint32_t num, uint32_t maxUnits, int floatMode,
                    const uint32_t* sizeDev, const uint8_t* successDev, cudaStream_t s, uint32_t* mismatchIdx,
                    uint32_t mismatchCap, uint32_t* numMismatch) {
  uint32_t* computed = tmp.alloc<uint32_t>(num);
  uint8_t* flags = tmp.alloc<uint8_t>(num);
  if (tmp.failed()) return kErrCuda;
  // The reference checksums bytes [0, count) of the output where count is in the provider's units
  // (bytes for ANS, WORDS for floats: dietgpu/float/GpuFloatDecompress.cuh:1078-1112 with
  // dietgpu/ans/GpuChecksum.cuh:95-102), mirroring what compression checksummed.
  int rc = launchChecksum(elemsDev, num, maxUnits, 1, sizeDev, 1, computed, s);
  if (rc) return rc;
  {
    ProfScope ps__("k_verifyChecksum", s);
    k_verifyChecksum<<<divUp(num, 128), 128, 0, s>>>(elemsDev, num, floatMode, computed, successDev, flags);
  }
  DGB_CUDA(cudaGetLastError());
  std::vector<uint8_t> host(num);
  DGB_CUDA(cudaMemcpyAsync(host.data(), flags, num, cudaMemcpyDeviceToHost, s));
  DGB_CUDA(cudaStreamSynchronize(s));
  uint32_t bad = 0;
  for (uint32_t i = 0; i < num; ++i) {
    if (host[i]) {
      if (mismatchIdx && bad < mismatchCap) mismatchIdx[bad] = i;
      ++bad;
    }
  }
  if (numMismatch) *numMismatch = bad;
  return bad ? kChecksumMismatch : kOk;
}

// ---- ANS ----------------------------------------------------------------------------------
bool fusedCompressEnabled();
uint32_t fusedMinTiles();
bool fusedWanted(uint32_t totalTiles, uint32_t numElems);

int ansEncode(void* temp, size_t tempBytes, int pb, int useChecksum, uint32_t num, const void* const* in,
              const uint32_t* inSize, const uint32_t* histDev, void* const* out, uint32_t* outSizeDev,
              cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in || !inSize || !out || pb < 9 || pb > 11) return kErrInvalidArg;
  if (!ansSizesOk(num, inSize)) return kErrInvalidArg;
  for (uint32_t i = 0; i < num; ++i)
    if (((uintptr_t)out[i]) & 15u) return kErrAlignment;
  Temp tmp(temp, tempBytes, s);
  WorkList wl;
  int rc = buildWorkList(wl, num, in, out, inSize, kTileSyms);
  if (rc) return rc;
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  uint32_t* checksum = nullptr;
  const uint32_t* hist = histDev;
  const bool fused = !histDev && fusedWanted(wl.totalTiles, num) && fusedCompressEnabled(); // histogram, tables, encoding and compaction in one launch
  uint32_t* ownHist = nullptr;
  if (!hist) {
    ownHist = tmp.alloc<uint32_t>((size_t)num * 256);
    if (tmp.failed()) return kErrCuda;
    DGB_CUDA(cudaMemsetAsync(ownHist, 0, (size_t)num * 1024, s));
    hist = ownHist;
  }
  // The checksum rides the histogram pass (one XOR per loaded vector, one atomicXor per warp and element) whenever
  // that pass runs; only a caller-supplied histogram leaves no pass over the data and keeps the separate k_checksum.
  const bool foldCk = useChecksum && ownHist && DGB_KNOB("DGB_FOLD_CHECKSUM", 1) != 0;
  if (useChecksum) {
    checksum = tmp.alloc<uint32_t>(num);
    if (tmp.failed()) return kErrCuda;
    if (foldCk) DGB_CUDA(cudaMemsetAsync(checksum, 0, sizeof(uint32_t) * num, s));
    else if ((rc = launchChecksum(elemsDev, num, wl.maxN, 1, nullptr, 0, checksum, s))) return rc;
  }
  EncodeArgs a;
  a.elems = elemsDev;
  a.numElems = num;
  a.totalTiles = wl.totalBlocks;
  a.nDev = nullptr;
  a.planeBase = nullptr;
  a.hist = hist;
  a.checksum = checksum;
  a.outSize = outSizeDev;
  a.floatType = kFtNone;
  a.round = 0;
  a.probBits = pb;
  a.useChecksum = useChecksum ? 1 : 0;
  EncodeScratch es;
  if ((rc = allocEncodeScratch(tmp, a, wl.totalGroups, es, s))) return rc;
  if (fused) return launchFusedCompress(kFtNone, a, es, wl.totalTiles, foldCk ? checksum : nullptr, 0, s, foldCk);
  // separate launches: the histogram kernel also builds the tables (unless the caller supplied histograms)
  if (ownHist && (rc = launchHistogram(elemsDev, num, wl.totalTiles, ownHist, s, es.fusedCtrl + kFusedCtrlFixed,
                                       es.table, pb, foldCk ? checksum : nullptr)))
    return rc;
  return launchAnsEncode(a, es, wholeRange(wl, num), s, ownHist != nullptr);
}

int ansDecode(void* temp, size_t tempBytes, int pb, int useChecksum, uint32_t num, const void* const* in,
              void* const* out, const uint32_t* outCapacity, uint8_t* successDev, uint32_t* sizeDev,
              cudaStream_t s, uint32_t* mismatchIdx, uint32_t mismatchCap, uint32_t* numMismatch) {
  if (numMismatch) *numMismatch = 0;
  if (num == 0) return kOk;
  if (!in || !out || !outCapacity || pb < 9 || pb > 11) return kErrInvalidArg;
  for (uint32_t i = 0; i < num; ++i)
    if (((uintptr_t)in[i]) & 15u) return kErrAlignment;
  Temp tmp(temp, tempBytes, s);
  WorkList wl;
  int rc = buildWorkList(wl, num, in, out, outCapacity, kTileSyms);
  if (rc) return rc;
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  if (useChecksum && !sizeDev) sizeDev = tmp.alloc<uint32_t>(num);
  if (useChecksum && !successDev) successDev = tmp.alloc<uint8_t>(num);
  if (tmp.failed()) return kErrCuda;
  DecodeArgs a;
  a.elems = elemsDev;
  a.numElems = num;
  a.totalTiles = wl.totalTiles;
  a.floatType = kFtNone;
  a.sparseNested = 0;
  a.probBits = pb;
  a.outSuccess = successDev;
  a.outSize = sizeDev;
  a.denseBase = nullptr;
  if ((rc = launchDecode(tmp, a, s))) return rc;
  if (useChecksum)
    return verifyChecksums(tmp, elemsDev, num, wl.maxN, 0, sizeDev, successDev, s, mismatchIdx, mismatchCap,
                           numMismatch);
  return kOk;
}

// ---- float ---------------------------------------------------------------------------------
// Shared by floatCompress and the sparse path (which feeds the dense non-zeros with device-side
// counts). wl.elems: in/out/n/tileBase/planeOff/sizeBias filled; elemsDev already uploaded.
// Second stream (highest priority) on which the split kernels of the later element groups run
// while the caller's stream encodes the earlier ones.
cudaStream_t sideStream() {
  static thread_local cudaStream_t side[kMaxDevices] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDevices) return nullptr;
  if (!side[dev]) {
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    if (cudaStreamCreateWithPriority(&side[dev], cudaStreamNonBlocking, hi) != cudaSuccess) side[dev] = nullptr;
  }
  return side[dev];
}

// DGB_FUSED=0 selects the separate split / normalise / encode launches (kept for the sparse and fp64 paths).
// The fused persistent kernel pays a start-up bubble (encode warps wait for the first element's table) and a
// compaction tail; measured on B200 it loses below about 1024 8-block tiles (32 MiB of symbols; bf16 16 x 1 MiB:
// 0.068 ms fused vs 0.048 ms as separate launches) and, above that, wins when there are elements to overlap
// (fusedWanted below).
constexpr uint32_t kFusedMinElems = 1;
static long long g_fusedMinTiles = -1; // < 0: default (or DGB_FUSED_MIN_TILES); set by dgb_set_fused_min_tiles (atomic accesses)
static long long fusedMinTilesOverride() { return __atomic_load_n(&g_fusedMinTiles, __ATOMIC_RELAXED); }
uint32_t fusedMinTiles() {
  const long long o = fusedMinTilesOverride();
  if (o >= 0) return o > 0xffffffffll ? 0xffffffffu : (uint32_t)o;
  const int v = DGB_KNOB("DGB_FUSED_MIN_TILES", -1); // experiment knob
  return v >= 0 ? (uint32_t)v : 1024u;
}

// The fused kernel wins by overlapping the split of later elements with the encoding of earlier ones; with few
// elements there is little to overlap and its phases (split, encode, compaction of one element) run one after the
// other inside one grid that is tuned for the mix. Measured on B200, bf16, fused / separate launches in ms:
//   1 x 64 MiB 0.128 / 0.101   4 x 16 MiB 0.114 / 0.104   3 x 32 MiB 0.160 / 0.141   6 x 16 MiB 0.150 / 0.143
//   1 x 256 MiB 0.325 / 0.289  2 x 64 MiB 0.179 / 0.191   8 x 16 MiB 0.180 / 0.193   16 x 8 MiB 0.165 / 0.192
//   32 x 2 MiB 0.107 / 0.112   2 x 256 MiB 0.483 / 0.530  1 x 1 GiB 0.911 / 0.995    fp32 1 x 1 GiB 0.669 / 0.769
// An explicit threshold (dgb_set_fused_min_tiles / DGB_FUSED_MIN_TILES) is taken literally.
bool fusedWanted(uint32_t totalTiles, uint32_t numElems) {
  if (fusedMinTilesOverride() >= 0 || DGB_KNOB("DGB_FUSED_MIN_TILES", -1) >= 0) return totalTiles >= fusedMinTiles();
  if (totalTiles < 1024) return false;
  if (numElems >= 16) return true;
  if (numElems >= 2 && totalTiles >= 2048) return true;
  return totalTiles >= 8192;
}

bool fusedCompressEnabled() {
  return DGB_KNOB("DGB_FUSED", 1) != 0;
}

int overlapGroups() {
  int g = DGB_KNOB("DGB_OVERLAP_GROUPS", 1); // measured on B200: 2/4/8 groups are 1-19 % slower than 1 (kernels of two streams do not co-run well)
  if (g < 1) g = 1;
  if (g > 16) g = 16;
  return g;
}

int floatCompressCore(Temp& tmp, int ft, int pb, int useChecksum, const WorkList& wl, uint32_t num,
                      const Elem* elemsDev, const uint32_t* nDev, const uint8_t* denseIn,
                      uint32_t* outSizeDev, cudaStream_t s) {
  const int planes = ft == kFtF64 ? 2 : 1;
  const size_t planeBytes = roundUpSz(wl.planeUnits + 16, 256);
  uint8_t* plane = (uint8_t*)tmp.alloc(planeBytes * planes);
  uint32_t* hist = tmp.alloc<uint32_t>((size_t)planes * num * 256);
  uint32_t* checksum = useChecksum ? tmp.alloc<uint32_t>(num) : nullptr;
  if (tmp.failed()) return kErrCuda;
  DGB_CUDA(cudaMemsetAsync(hist, 0, (size_t)planes * num * 1024, s));
  int rc;
  // Split is HBM-bound, the encoder issue-bound: cut the batch into element groups and run
  // split(g + 1) on a side stream while the caller's stream encodes group g.
  int groups = overlapGroups();
  cudaStream_t side = nullptr;
  if (num < 8 || wl.totalTiles < 1024 || planes > 1) groups = 1;
  if (groups > 1) side = sideStream();
  if (!side) groups = 1;
  // The checksum rides the split pass (the CTA completing an element patches its header); the grouped schedule has
  // no completion protocol and keeps the separate k_checksum pass.
  const bool foldCk = useChecksum && groups == 1 && DGB_KNOB("DGB_FOLD_CHECKSUM", 1) != 0;
  if (useChecksum) {
    // the float-level checksum covers the first N BYTES of the N-word input
    // (dietgpu/float/GpuFloatCompress.cuh:708-710 with dietgpu/ans/GpuChecksum.cuh:95-102)
    if (denseIn) return kErrInvalidArg; // sparse + checksum: not supported (reference: untested)
    if (foldCk) DGB_CUDA(cudaMemsetAsync(checksum, 0, sizeof(uint32_t) * num, s));
    else if ((rc = launchChecksum(elemsDev, num, wl.maxN, 1, nDev, 0, checksum, s))) return rc;
  }
  EncodeArgs args[2];
  EncodeScratch es[2];
  for (int r = 0; r < planes; ++r) {
    EncodeArgs& a = args[r];
    a.elems = elemsDev;
    a.numElems = num;
    a.totalTiles = wl.totalBlocks;
    a.nDev = nDev;
    a.planeBase = plane + planeBytes * r;
    a.hist = hist + (size_t)r * num * 256;
    a.checksum = nullptr;
    a.outSize = outSizeDev;
    a.floatType = ft;
    a.round = r;
    a.probBits = pb;
    a.useChecksum = 0; // not allowed in float mode (dietgpu/float/GpuFloatCodec.h:52)
    if ((rc = allocEncodeScratch(tmp, a, wl.totalGroups, es[r], s))) return rc;
  }

  std::vector<WorkRange> ranges;
  {
    uint32_t e0 = 0;
    for (int g = 0; g < groups; ++g) {
      const uint32_t target = (uint32_t)(((uint64_t)wl.totalTiles * (g + 1)) / groups);
      uint32_t e1 = e0;
      while (e1 < num && (g == groups - 1 || wl.elems[e1 + 1].tileBase <= target)) ++e1;
      if (g == groups - 1) e1 = num;
      WorkRange r{e0, e1, wl.elems[e0].tileBase, wl.elems[e1].tileBase, wl.elems[e0].blockBase,
                  wl.elems[e1].blockBase};
      if (e1 > e0) ranges.push_back(r);
      e0 = e1;
    }
  }
  uint32_t minElemTiles = 0xffffffffu;
  for (uint32_t i = 0; i < num; ++i)
    minElemTiles = std::min(minElemTiles, wl.elems[i + 1].tileBase - wl.elems[i].tileBase);
  auto split = [&](const WorkRange& r, cudaStream_t st, bool withTables) {
    return launchFloatSplit(ft, elemsDev, num, r.tile0, r.tile1, nDev, denseIn, plane, plane + planeBytes, hist,
                            hist + (size_t)num * 256, checksum, useChecksum, st,
                            withTables ? es[0].fusedCtrl + kFusedCtrlFixed : nullptr, es[0].table,
                            planes == 2 ? es[1].table : nullptr, pb, minElemTiles, foldCk && withTables);
  };
  if (ranges.size() <= 1) {
    const WorkRange all = wholeRange(wl, num);
    if (planes == 1 && !nDev && !denseIn && num >= kFusedMinElems && fusedWanted(wl.totalTiles, num) &&
        fusedCompressEnabled())
      return launchFusedCompress(ft, args[0], es[0], wl.totalTiles, checksum, useChecksum, s, foldCk);
    if ((rc = split(all, s, true))) return rc; // the split kernel also builds the encode tables
    for (int r = 0; r < planes; ++r)
      if ((rc = launchAnsEncode(args[r], es[r], all, s, true))) return rc;
    return kOk;
  }
  cudaEvent_t ev;
  DGB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
  DGB_CUDA(cudaEventRecord(ev, s));
  DGB_CUDA(cudaStreamWaitEvent(side, ev, 0));
  DGB_CUDA(cudaEventDestroy(ev));
  for (size_t g = 0; g < ranges.size(); ++g) {
    if ((rc = split(ranges[g], side, false))) return rc;
    DGB_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    DGB_CUDA(cudaEventRecord(ev, side));
    DGB_CUDA(cudaStreamWaitEvent(s, ev, 0));
    DGB_CUDA(cudaEventDestroy(ev));
    if ((rc = launchAnsEncode(args[0], es[0], ranges[g], s))) return rc;
  }
  return kOk;
}

int floatCompress(void* temp, size_t tempBytes, int ft, int pb, int useChecksum, uint32_t num,
                  const void* const* in, const uint32_t* inSize, void* const* out, uint32_t* outSizeDev,
                  cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in || !inSize || !out || !validFloatType(ft) || pb < 9 || pb > 11) return kErrInvalidArg;
  if (!floatSizesOk(ft, num, inSize)) return kErrInvalidArg;
  const uint32_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    if (((uintptr_t)out[i]) & 15u) return kErrAlignment;
    if (((uintptr_t)in[i]) & (wb - 1)) return kErrAlignment;
  }
  Temp tmp(temp, tempBytes, s);
  WorkList wl;
  int rc = buildWorkList(wl, num, in, out, inSize, kTileSyms);
  if (rc) return rc;
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  return floatCompressCore(tmp, ft, pb, useChecksum, wl, num, elemsDev, nullptr, nullptr, outSizeDev, s);
}

int floatDecompress(void* temp, size_t tempBytes, int ft, int pb, int useChecksum, uint32_t num,
                    const void* const* in, void* const* out, const uint32_t* outCapacity,
                    uint8_t* successDev, uint32_t* sizeDev, cudaStream_t s, uint32_t* mismatchIdx,
                    uint32_t mismatchCap, uint32_t* numMismatch) {
  if (numMismatch) *numMismatch = 0;
  if (num == 0) return kOk;
  if (!in || !out || !outCapacity || !validFloatType(ft) || pb < 9 || pb > 11) return kErrInvalidArg;
  const uint32_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    if (((uintptr_t)in[i]) & 15u) return kErrAlignment;
    if (((uintptr_t)out[i]) & (wb - 1)) return kErrAlignment;
  }
  Temp tmp(temp, tempBytes, s);
  WorkList wl;
  int rc = buildWorkList(wl, num, in, out, outCapacity, kTileSyms);
  if (rc) return rc;
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  if (useChecksum && !sizeDev) sizeDev = tmp.alloc<uint32_t>(num);
  if (useChecksum && !successDev) successDev = tmp.alloc<uint8_t>(num);
  if (tmp.failed()) return kErrCuda;
  DecodeArgs a;
  a.elems = elemsDev;
  a.numElems = num;
  a.totalTiles = wl.totalTiles;
  a.floatType = ft;
  a.sparseNested = 0;
  a.probBits = pb;
  a.outSuccess = successDev;
  a.outSize = sizeDev;
  a.denseBase = nullptr;
  if ((rc = launchDecode(tmp, a, s))) return rc;
  if (useChecksum)
    return verifyChecksums(tmp, elemsDev, num, wl.maxN, 1, sizeDev, successDev, s, mismatchIdx, mismatchCap,
                           numMismatch);
  return kOk;
}

// ---- sparse float ----------------------------------------------------------------------------
constexpr uint32_t kSparseTileWords = 8192; // must match kSpTileWords in sparse.cu

int floatCompressSparse(void* temp, size_t tempBytes, int ft, int pb, int useChecksum, uint32_t num,
                        const void* const* in, const uint32_t* inSize, void* const* out,
                        uint32_t* outSizeDev, cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in || !inSize || !out || !validFloatType(ft) || pb < 9 || pb > 11) return kErrInvalidArg;
  if (useChecksum) return kErrInvalidArg;
  if (!floatSizesOk(ft, num, inSize, 32)) return kErrInvalidArg;
  const uint32_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    if (((uintptr_t)out[i]) & 15u) return kErrAlignment;
    if (((uintptr_t)in[i]) & (wb - 1)) return kErrAlignment;
  }
  Temp tmp(temp, tempBytes, s);
  // pass 1: bitmap + compaction, tiles of 8192 words
  WorkList spl;
  int rc = buildWorkList(spl, num, in, out, inSize, kSparseTileWords);
  if (rc) return rc;
  Elem* splDev;
  if ((rc = upload(tmp, spl, &splDev, s))) return rc;
  uint8_t* dense = (uint8_t*)tmp.alloc((spl.planeUnits + 16) * wb);
  uint32_t* nnz = tmp.alloc<uint32_t>(num);
  if (tmp.failed()) return kErrCuda;
  if ((rc = launchSparseCompact(tmp, ft, splDev, num, spl.totalTiles, dense, nnz, s))) return rc;
  // pass 2: the dense float codec over the non-zeros, sizes known only on the device
  WorkList wl;
  if ((rc = buildWorkList(wl, num, in, out, inSize, kTileSyms))) return rc;
  for (uint32_t i = 0; i < num; ++i) wl.elems[i].sizeBias = 16 + sparseBitmapBytes(inSize[i]);
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  return floatCompressCore(tmp, ft, pb, 0, wl, num, elemsDev, nnz, dense, outSizeDev, s);
}

int floatDecompressSparse(void* temp, size_t tempBytes, int ft, int pb, int useChecksum, uint32_t num,
                          const void* const* in, void* const* out, const uint32_t* outCapacity,
                          uint8_t* successDev, uint32_t* sizeDev, cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in || !out || !outCapacity || !validFloatType(ft) || pb < 9 || pb > 11) return kErrInvalidArg;
  if (useChecksum) return kErrInvalidArg;
  const uint32_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    if (((uintptr_t)in[i]) & 15u) return kErrAlignment;
    if (((uintptr_t)out[i]) & (wb - 1)) return kErrAlignment;
  }
  Temp tmp(temp, tempBytes, s);
  WorkList wl;
  int rc = buildWorkList(wl, num, in, out, outCapacity, kTileSyms);
  if (rc) return rc;
  Elem* elemsDev;
  if ((rc = upload(tmp, wl, &elemsDev, s))) return rc;
  uint8_t* dense = (uint8_t*)tmp.alloc((wl.planeUnits + 16) * wb);
  if (tmp.failed()) return kErrCuda;
  DecodeArgs a;
  a.elems = elemsDev;
  a.numElems = num;
  a.totalTiles = wl.totalTiles;
  a.floatType = ft;
  a.sparseNested = 1;
  a.probBits = pb;
  a.outSuccess = nullptr;
  a.outSize = nullptr;
  a.denseBase = dense;
  if ((rc = launchDecode(tmp, a, s))) return rc;
  WorkList spl;
  if ((rc = buildWorkList(spl, num, in, out, outCapacity, kSparseTileWords))) return rc;
  for (uint32_t i = 0; i < num; ++i) spl.elems[i].planeOff = wl.elems[i].planeOff;
  Elem* splDev;
  if ((rc = upload(tmp, spl, &splDev, s))) return rc;
  return launchSparseScatter(tmp, ft, splDev, num, spl.totalTiles, dense, successDev, sizeDev, s);
}

// ---- info ------------------------------------------------------------------------------------
// dietgpu/ans/GpuANSInfo.cuh:16-37, dietgpu/float/GpuFloatInfo.cuh:18-41
__global__ void k_info(const void* const* __restrict__ in, uint32_t num, int floatMode,
                       uint32_t* __restrict__ sizes, uint32_t* __restrict__ types, uint32_t* __restrict__ checksums) {
  uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= num) return;
  const uint32_t* h = (const uint32_t*)in[i];
  if (floatMode) {
    if (sizes) sizes[i] = h[1];
    if (types) types[i] = h[2] & 0xfu;
    if (checksums) checksums[i] = h[3];
  } else {
    if (sizes) sizes[i] = h[2];
    if (checksums) checksums[i] = h[5];
  }
}

int getInfo(const void* const* in, int inIsDevice, uint32_t num, int floatMode, uint32_t* sizes, uint32_t* types,
            uint32_t* checksums, void* temp, size_t tempBytes, cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in) return kErrInvalidArg;
  Temp tmp(temp, tempBytes, s);
  const void* const* inDev = in;
  if (!inIsDevice) {
    void** d = tmp.alloc<void*>(num);
    if (tmp.failed()) return kErrCuda;
    DGB_CUDA(cudaMemcpyAsync(d, in, sizeof(void*) * num, cudaMemcpyHostToDevice, s));
    inDev = d;
  }
  {
    ProfScope ps__("k_info", s);
    k_info<<<divUp(num, 128), 128, 0, s>>>(inDev, num, floatMode, sizes, types, checksums);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

// ---- size bounds in 64 bits (dietgpu/ans/GpuANSEncode.cu:13-25, dietgpu/float/GpuFloatCompress.cu:23-47) --------
uint64_t ansMaxBytes64(uint64_t n) {
  const uint64_t blocks = (n + kBlockSyms - 1) / kBlockSyms;
  uint64_t raw = ansOverhead(kBlockSyms);
  raw += (uint64_t)roundUp(kBlockSyms + kBlockSyms / 4, 16) * blocks;
  return roundUpSz(raw, 16);
}
uint64_t floatRawBytes64(int ft, uint64_t n) {
  switch (ft) {
    case kFtF16:
    case kFtBF16: return roundUpSz(n, 16);
    case kFtF32: return 2 * roundUpSz(n, 8) + roundUpSz(n, 16);
    case kFtF64: return 4 * roundUpSz(n, 4) + 2 * roundUpSz(n, 8);
    default: return 0;
  }
}
// > UINT32_MAX also when the ANS part alone would have made the reference abort
uint64_t floatMaxBytes64(int ft, uint64_t n) {
  const uint64_t a = ansMaxBytes64(n);
  if (a > 0x7fffffffull) return ~0ull;
  return kFloatHeaderBytes + a + floatRawBytes64(ft, n) + (ft == kFtF64 ? a : 0);
}

// ---- stride / split-size addressing -> pointer arrays -------------------------------------------
struct PtrArrays {
  std::vector<const void*> in;
  std::vector<void*> out;
  std::vector<uint32_t> n;
};

} // namespace
} // namespace dgb

using namespace dgb;

extern "C" {

const char* dgb_version(void) { return "dietgpu_b200 r2 sm_100a"; }

size_t dgb_last_temp_bytes_used(void) { return Temp::lastHighWater(); }

void dgb_set_fused_min_tiles(long long tiles) { __atomic_store_n(&g_fusedMinTiles, tiles, __ATOMIC_RELAXED); }

// dietgpu/ans/GpuANSEncode.cu:13-25: the reference passes the block SIZE (4096) where
// getCompressedOverhead expects a block COUNT; the value is part of the API (Python sizes and
// validates buffers with it), so it is reproduced.
// All three bounds are computed in 64 bits and return 0 ("no valid bound") when the result leaves the range the
// 32-bit API can express: the reference CHECK-aborts when the ANS bound exceeds INT32_MAX (same file, :22) and lets
// the uint32 float sums wrap silently; callers here must treat 0 for n > 0 as an error (codec.py / torch_ops.py raise).
uint32_t dgb_ans_max_compressed_size(uint32_t n) {
  const uint64_t v = ansMaxBytes64(n);
  return v > 0x7fffffffull ? 0u : (uint32_t)v;
}

uint32_t dgb_float_max_compressed_size(int ft, uint32_t n) {
  const uint64_t v = floatMaxBytes64(ft, n);
  return v > 0xffffffffull ? 0u : (uint32_t)v;
}

uint32_t dgb_sparse_float_max_compressed_size(int ft, uint32_t n) {
  const uint64_t f = floatMaxBytes64(ft, n);
  const uint64_t v = f + 16u + roundUpSz(((uint64_t)n + 7) / 8, 16);
  return (f > 0xffffffffull || v > 0xffffffffull) ? 0u : (uint32_t)v;
}

size_t dgb_temp_bytes_required(int op, int ft, uint32_t num, uint32_t maxSize) {
  const size_t B = num ? num : 1;
  const size_t tiles = B * (divUp(maxSize, kTileSyms) + 1);
  const size_t spTiles = B * (divUp(maxSize, 8192) + 1);
  const size_t desc = (B + 1) * sizeof(Elem) + 512;
  const size_t blocks = B * (divUp(maxSize, kBlockSyms) + 1);
  const size_t enc = B * 256 * 8 + tiles * 8 + B * 32 + blocks * (4 + 5632) + 6144; // tables, control words, scratch slots
  const size_t dec = B * (sizeof(DecInfo) + (4u << 11)) * 2 + 1024;
  const size_t row = roundUpSz(maxSize, 16);
  const size_t wb = floatWordBytes(ft);
  const int planes = ft == kFtF64 ? 2 : 1;
  switch (op) {
    case 0: return desc + B * 1024 + B * 4 + enc + 2048;
    case 1: return desc + dec + B * 16 + 2048;
    case 2: return desc + planes * (B * row + 512) + planes * B * 1024 + B * 4 + planes * enc + 4096;
    case 3: return desc + dec + B * 16 + 2048;
    case 4:
      return 2 * desc + (B * row + 16) * wb + B * 4 + spTiles * 8 + planes * (B * row + 512) +
             planes * B * 1024 + planes * enc + 8192;
    case 5: return 2 * desc + (B * row + 16) * wb + dec + spTiles * 8 + 4096;
    default: return 0;
  }
}

int dgb_ans_encode_batch_pointer(void* temp, size_t tempBytes, int pb, int cs, uint32_t num,
                                 const void* const* in, const uint32_t* inSize, const uint32_t* hist,
                                 void* const* out, uint32_t* outSize, void* stream) {
  return ansEncode(temp, tempBytes, pb, cs, num, in, inSize, hist, out, outSize, (cudaStream_t)stream);
}

int dgb_ans_encode_batch_stride(void* temp, size_t tempBytes, int pb, int cs, uint32_t num, const void* in,
                                uint32_t inSize, uint32_t inStride, const uint32_t* hist, void* out,
                                uint32_t outStride, uint32_t* outSize, void* stream) {
  PtrArrays p;
  for (uint32_t i = 0; i < num; ++i) {
    p.in.push_back((const uint8_t*)in + (size_t)i * inStride);
    p.out.push_back((uint8_t*)out + (size_t)i * outStride);
    p.n.push_back(inSize);
  }
  return ansEncode(temp, tempBytes, pb, cs, num, p.in.data(), p.n.data(), hist, p.out.data(), outSize,
                   (cudaStream_t)stream);
}

int dgb_ans_encode_batch_split_size(void* temp, size_t tempBytes, int pb, int cs, uint32_t num, const void* in,
                                    const uint32_t* splits, const uint32_t* hist, void* out,
                                    uint32_t outStride, uint32_t* outSize, void* stream) {
  if (num && !splits) return kErrInvalidArg;
  PtrArrays p;
  size_t off = 0;
  for (uint32_t i = 0; i < num; ++i) {
    p.in.push_back((const uint8_t*)in + off);
    p.out.push_back((uint8_t*)out + (size_t)i * outStride);
    p.n.push_back(splits[i]);
    off += splits[i];
  }
  return ansEncode(temp, tempBytes, pb, cs, num, p.in.data(), p.n.data(), hist, p.out.data(), outSize,
                   (cudaStream_t)stream);
}

int dgb_ans_decode_batch_pointer(void* temp, size_t tempBytes, int pb, int cs, uint32_t num,
                                 const void* const* in, void* const* out, const uint32_t* cap,
                                 uint8_t* success, uint32_t* size, void* stream, uint32_t* mi, uint32_t mc,
                                 uint32_t* nm) {
  return ansDecode(temp, tempBytes, pb, cs, num, in, out, cap, success, size, (cudaStream_t)stream, mi, mc, nm);
}

int dgb_ans_decode_batch_stride(void* temp, size_t tempBytes, int pb, int cs, uint32_t num, const void* in,
                                uint32_t inStride, void* out, uint32_t outStride, uint32_t outCap,
                                uint8_t* success, uint32_t* size, void* stream, uint32_t* mi, uint32_t mc,
                                uint32_t* nm) {
  PtrArrays p;
  for (uint32_t i = 0; i < num; ++i) {
    p.in.push_back((const uint8_t*)in + (size_t)i * inStride);
    p.out.push_back((uint8_t*)out + (size_t)i * outStride);
    p.n.push_back(outCap);
  }
  return ansDecode(temp, tempBytes, pb, cs, num, p.in.data(), p.out.data(), p.n.data(), success, size,
                   (cudaStream_t)stream, mi, mc, nm);
}

int dgb_ans_decode_batch_split_size(void* temp, size_t tempBytes, int pb, int cs, uint32_t num,
                                    const void* const* in, void* out, const uint32_t* splits,
                                    uint8_t* success, uint32_t* size, void* stream, uint32_t* mi,
                                    uint32_t mc, uint32_t* nm) {
  if (num && !splits) return kErrInvalidArg;
  PtrArrays p;
  size_t off = 0;
  for (uint32_t i = 0; i < num; ++i) {
    p.out.push_back((uint8_t*)out + off);
    off += splits[i];
  }
  return ansDecode(temp, tempBytes, pb, cs, num, in, p.out.data(), splits, success, size, (cudaStream_t)stream,
                   mi, mc, nm);
}

int dgb_ans_get_compressed_info(const void* const* in, int inIsDevice, uint32_t num, uint32_t* sizes,
                                uint32_t* checksums, void* temp, size_t tempBytes, void* stream) {
  return getInfo(in, inIsDevice, num, 0, sizes, nullptr, checksums, temp, tempBytes, (cudaStream_t)stream);
}

int dgb_float_compress(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                       const void* const* in, const uint32_t* inSize, void* const* out, uint32_t* outSize,
                       void* stream) {
  return floatCompress(temp, tempBytes, ft, pb, cs, num, in, inSize, out, outSize, (cudaStream_t)stream);
}

int dgb_float_compress_split_size(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                                  const void* in, const uint32_t* splits, void* out, uint32_t outStride,
                                  uint32_t* outSize, void* stream) {
  if (num && !splits) return kErrInvalidArg;
  PtrArrays p;
  size_t off = 0;
  const size_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    p.in.push_back((const uint8_t*)in + off * wb);
    p.out.push_back((uint8_t*)out + (size_t)i * outStride);
    off += splits[i];
  }
  return floatCompress(temp, tempBytes, ft, pb, cs, num, p.in.data(), splits, p.out.data(), outSize,
                       (cudaStream_t)stream);
}

int dgb_float_compress_sparse(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                              const void* const* in, const uint32_t* inSize, void* const* out,
                              uint32_t* outSize, void* stream) {
  return floatCompressSparse(temp, tempBytes, ft, pb, cs, num, in, inSize, out, outSize, (cudaStream_t)stream);
}

int dgb_float_decompress(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                         const void* const* in, void* const* out, const uint32_t* cap, uint8_t* success,
                         uint32_t* size, void* stream, uint32_t* mi, uint32_t mc, uint32_t* nm) {
  return floatDecompress(temp, tempBytes, ft, pb, cs, num, in, out, cap, success, size, (cudaStream_t)stream, mi,
                         mc, nm);
}

int dgb_float_decompress_split_size(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                                    const void* const* in, void* out, const uint32_t* splits,
                                    uint8_t* success, uint32_t* size, void* stream, uint32_t* mi,
                                    uint32_t mc, uint32_t* nm) {
  if (num && !splits) return kErrInvalidArg;
  PtrArrays p;
  size_t off = 0;
  const size_t wb = floatWordBytes(ft);
  for (uint32_t i = 0; i < num; ++i) {
    p.out.push_back((uint8_t*)out + off * wb);
    off += splits[i];
  }
  return floatDecompress(temp, tempBytes, ft, pb, cs, num, in, p.out.data(), splits, success, size,
                         (cudaStream_t)stream, mi, mc, nm);
}

int dgb_float_decompress_sparse(void* temp, size_t tempBytes, int ft, int pb, int cs, uint32_t num,
                                const void* const* in, void* const* out, const uint32_t* cap,
                                uint8_t* success, uint32_t* size, void* stream) {
  return floatDecompressSparse(temp, tempBytes, ft, pb, cs, num, in, out, cap, success, size,
                               (cudaStream_t)stream);
}

int dgb_float_get_compressed_info(const void* const* in, int inIsDevice, uint32_t num, uint32_t* sizes,
                                  uint32_t* types, uint32_t* checksums, void* temp, size_t tempBytes,
                                  void* stream) {
  return getInfo(in, inIsDevice, num, 1, sizes, types, checksums, temp, tempBytes, (cudaStream_t)stream);
}

} // extern "C"
