// ANS encode side for sm_100a: byte histogram, checksum, probability normalisation, block encoder and
// output compaction kernels plus their launchers. The device-side building blocks live in ans_encode.cuh.
//
// What is computed follows the reference (citations relative to /root/reference/):
//   histogram            dietgpu/ans/GpuANSStatistics.cuh:21-143
//   normalisation        dietgpu/ans/GpuANSStatistics.cuh:178-367
//   per-warp rANS coder  dietgpu/ans/GpuANSEncode.cuh:49-211
//   archive assembly     dietgpu/ans/GpuANSEncode.cuh:511-624
// How it is computed does not: the reference encodes into an uncoalesced temp buffer, runs a 1-3 kernel
// prefix sum and copies every compressed byte a second time with one CTA per block; here warps encode
// runs of blocks as one stream into L2-resident scratch slots and a single compaction kernel that already
// knows every size writes payload, index, pdf and header.
#include <cstdlib>

#include "ans_encode.cuh"
#include "float_split.cuh"
#include "internal.h"
#include "knobs.h"
#include "prof.h"

namespace dgb {

int currentDevice() {
  int dev = 0;
  cudaGetDevice(&dev);
  return dev < 0 ? 0 : (dev >= kMaxDevices ? kMaxDevices - 1 : dev);
}

int smCount() {
  static int cached[kMaxDevices] = {}; // written once per device with the same value: a race is benign
  const int dev = currentDevice();
  int v = __atomic_load_n(&cached[dev], __ATOMIC_RELAXED);
  if (!v) {
    cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    if (v <= 0) v = 148;
    __atomic_store_n(&cached[dev], v, __ATOMIC_RELAXED);
  }
  return v;
}

// ---------------------------------------------------------------------------------------------
// Byte histogram. Each CTA owns a contiguous range of tiles, accumulates into per-warp,
// shared-memory bins (kSplitRep replicas, float_split.cuh) and flushes to global
// memory whenever the element changes.
// ---------------------------------------------------------------------------------------------
constexpr int kHistRep = kSplitRep; // same layout and increment as the float split's bins (splitHistAddAt)


// sums the 8 warps x kHistRep replicas of bin `tid`, adds to global, clears
__device__ __forceinline__ void histFlush(uint32_t* sh, uint32_t* gHist, uint32_t tid) {
  __syncthreads();
  uint32_t sum = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
#pragma unroll
    for (int r = 0; r < kHistRep; ++r) {
      uint32_t idx = w * 256 * kHistRep + tid * kHistRep + r;
      sum += sh[idx];
      sh[idx] = 0;
    }
  }
  if (sum) atomicAdd(&gHist[tid], sum);
  __syncthreads();
}

constexpr uint32_t kHistBytesSmem = kSplitHistAlign + 8 * 256 * kHistRep * 4; // < 48 KB: no opt-in needed
__global__ void __launch_bounds__(256) k_histBytes(const Elem* __restrict__ elems, uint32_t numElems,
                                                   uint32_t totalTiles, uint32_t tilesPerCta,
                                                   uint32_t* __restrict__ hist, uint32_t* __restrict__ normTilesDone,
                                                   uint2* __restrict__ normTable, int normProbBits,
                                                   uint32_t* __restrict__ checksumAcc) {
  extern __shared__ __align__(16) uint8_t smemRaw[]; // kHistBytesSmem: the bins, aligned inside the block
  uint32_t* const sh = (uint32_t*)splitHistBase(smemRaw);
  __shared__ uint32_t sLast;
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < 8 * 256 * kHistRep; i += 256) sh[i] = 0;
  __syncthreads();
  const uint32_t wh = smemAddr(sh) + (tid >> 5) * kSplitHistAlign + (tid & (kHistRep - 1)) * 4u; // see splitHistAddAt

  uint32_t t0 = blockIdx.x * tilesPerCta;
  uint32_t t1 = min(t0 + tilesPerCta, totalTiles);
  uint32_t cur = 0xffffffffu;
  uint32_t ck = 0; // checksum folded into this pass (every byte of a raw ANS input counts); flushed per element
  // counts of element `cur` -> global bins; the CTA whose tiles complete the element also builds its encode table
  // (saves the k_normalize launch: small batches are launch-latency bound)
  auto finishElement = [&]() {
    histFlush(sh, hist + cur * 256, tid);
    if (checksumAcc) ckFlushWarp(ck, checksumAcc + cur);
    if (normTilesDone) {
      const uint32_t eb = elems[cur].tileBase, ee = elems[cur + 1].tileBase;
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        const uint32_t count = min(t1, ee) - max(t0, eb);
        sLast = atomicAdd(&normTilesDone[cur], count) + count == ee - eb ? 1u : 0u;
      }
      __syncthreads();
      if (sLast) {
        __threadfence();
        normalizeElement<256>(hist + (size_t)cur * 256, elems[cur].n, normProbBits, normTable + (size_t)cur * 256, sh);
        __syncthreads();
        for (uint32_t i = tid; i < 256 + 16; i += 256) sh[i] = 0; // bins used as scratch
      }
      __syncthreads();
    }
  };
  for (uint32_t tile = t0; tile < t1; ++tile) {
    uint32_t e = findElement(&elems[0].tileBase, sizeof(Elem) / 4, numElems, tile, cur);
    if (e != cur) {
      if (cur != 0xffffffffu) finishElement();
      cur = e;
    }
    const uint8_t* base = (const uint8_t*)elems[e].in;
    uint32_t n = elems[e].n;
    uint32_t off = (tile - elems[e].tileBase) * kTileSyms;
    if (off >= n) continue;
    uint32_t len = min(kTileSyms, n - off);
    const uint8_t* p = base + off;
    const uint8_t* end = p + len;
    const uint8_t* a0 = (const uint8_t*)(((uintptr_t)p + 15) & ~(uintptr_t)15);
    if (a0 > end) a0 = end;
    const uint8_t* a1 = a0 + ((size_t)(end - a0) & ~(size_t)15);
    // head / tail bytes
    uint32_t headN = (uint32_t)(a0 - p), tailN = (uint32_t)(end - a1);
    if (tid < headN) {
      splitHistAddAt(wh, (uint32_t)p[tid] << kSplitSymShift);
      ck ^= p[tid];
    }
    if (tid >= 32 && tid - 32 < tailN) {
      splitHistAddAt(wh, (uint32_t)a1[tid - 32] << kSplitSymShift);
      ck ^= a1[tid - 32];
    }
    const uint4* v = (const uint4*)a0;
    uint32_t nvec = (uint32_t)(a1 - a0) >> 4;
    uint32_t i = tid;
    for (; i + 768 < nvec; i += 1024) {
      uint4 x0 = v[i], x1 = v[i + 256], x2 = v[i + 512], x3 = v[i + 768];
      splitHistAdd4(wh, x0.x); splitHistAdd4(wh, x0.y); splitHistAdd4(wh, x0.z); splitHistAdd4(wh, x0.w);
      splitHistAdd4(wh, x1.x); splitHistAdd4(wh, x1.y); splitHistAdd4(wh, x1.z); splitHistAdd4(wh, x1.w);
      splitHistAdd4(wh, x2.x); splitHistAdd4(wh, x2.y); splitHistAdd4(wh, x2.z); splitHistAdd4(wh, x2.w);
      splitHistAdd4(wh, x3.x); splitHistAdd4(wh, x3.y); splitHistAdd4(wh, x3.z); splitHistAdd4(wh, x3.w);
      ck ^= x0.x ^ x0.y ^ x0.z ^ x0.w ^ x1.x ^ x1.y ^ x1.z ^ x1.w ^ x2.x ^ x2.y ^ x2.z ^ x2.w ^ x3.x ^ x3.y ^ x3.z ^ x3.w;
    }
    for (; i < nvec; i += 256) {
      uint4 x0 = v[i];
      splitHistAdd4(wh, x0.x); splitHistAdd4(wh, x0.y); splitHistAdd4(wh, x0.z); splitHistAdd4(wh, x0.w);
      ck ^= x0.x ^ x0.y ^ x0.z ^ x0.w;
    }
  }
  if (cur != 0xffffffffu) finishElement();
}

int launchHistogram(const Elem* elems, uint32_t numElems, uint32_t totalTiles, uint32_t* hist, cudaStream_t s,
                    uint32_t* normTilesDone, uint2* normTable, int normProbBits, uint32_t* checksumAcc) {
  if (!totalTiles) return kOk;
  uint32_t ctas = min(totalTiles, (uint32_t)smCount() * 4u);
  uint32_t per = divUp(totalTiles, ctas);
  ctas = divUp(totalTiles, per);
  {
    ProfScope ps__("k_histBytes", s);
    k_histBytes<<<ctas, 256, kHistBytesSmem, s>>>(elems, numElems, totalTiles, per, hist, normTilesDone, normTable, normProbBits,
                                      checksumAcc);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

// ---------------------------------------------------------------------------------------------
// Checksum: XOR of the first (n * bytesPerUnit) bytes, folded to 8 bits
// (dietgpu/ans/GpuChecksum.cuh:26-133). One CTA per element (only used when the caller asks for
// checksums, which none of the timed configurations do).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_checksum(const Elem* __restrict__ elems, uint32_t bytesPerUnit,
                                                  const uint32_t* __restrict__ nDev, int useOut,
                                                  uint32_t* __restrict__ checksum) {
  uint32_t e = blockIdx.y;
  const uint8_t* p = (const uint8_t*)(useOut ? elems[e].out : elems[e].in);
  uint32_t units = nDev ? nDev[e] : elems[e].n;
  if (useOut && units > elems[e].n) units = elems[e].n; // never read past the output capacity
  size_t n = (size_t)units * bytesPerUnit;
  uint32_t acc = 0;
  size_t chunk = roundUpSz((n + gridDim.x - 1) / gridDim.x, 16);
  size_t b0 = (size_t)blockIdx.x * chunk;
  size_t b1 = b0 + chunk < n ? b0 + chunk : n;
  if (b0 < b1) {
    const uint8_t* q = p + b0;
    const uint8_t* end = p + b1;
    const uint8_t* a0 = (const uint8_t*)(((uintptr_t)q + 15) & ~(uintptr_t)15);
    if (a0 > end) a0 = end;
    const uint8_t* a1 = a0 + ((size_t)(end - a0) & ~(size_t)15);
    for (const uint8_t* c = q + threadIdx.x; c < a0; c += 256) acc ^= *c;
    for (const uint8_t* c = a1 + threadIdx.x; c < end; c += 256) acc ^= *c;
    const uint4* v = (const uint4*)a0;
    size_t nvec = (size_t)(a1 - a0) >> 4;
    for (size_t i = threadIdx.x; i < nvec; i += 256) {
      uint4 x = v[i];
      acc ^= x.x ^ x.y ^ x.z ^ x.w;
    }
  }
  acc ^= acc >> 16;
  acc ^= acc >> 8;
  acc &= 0xffu;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc ^= __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0 && acc) atomicXor(&checksum[e], acc);
}

int launchChecksum(const Elem* elems, uint32_t numElems, uint32_t maxUnits, uint32_t bytesPerUnit,
                   const uint32_t* nDev, int useOut, uint32_t* checksum, cudaStream_t s) {
  if (!numElems) return kOk;
  DGB_CUDA(cudaMemsetAsync(checksum, 0, sizeof(uint32_t) * numElems, s));
  size_t maxBytes = (size_t)maxUnits * bytesPerUnit;
  uint32_t perElem = (uint32_t)min((size_t)divUp((uint32_t)smCount() * 8u, numElems),
                                   (maxBytes + 65535) / 65536);
  if (perElem < 1) perElem = 1;
  {
    ProfScope ps__("k_checksum", s);
    k_checksum<<<dim3(perElem, numElems), 256, 0, s>>>(elems, bytesPerUnit, nDev, useOut, checksum);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

// Normalisation kernel: one 256-thread CTA per element (see normalizeElement in ans_encode.cuh).
__global__ void __launch_bounds__(256) k_normalize(const Elem* __restrict__ elems,
                                                   const uint32_t* __restrict__ nDev,
                                                   const uint32_t* __restrict__ hist, int probBits,
                                                   uint2* __restrict__ table, uint32_t elemBegin) {
  __shared__ uint32_t sm[256 + 16];
  const uint32_t e = elemBegin + blockIdx.x;
  const uint32_t total = nDev ? nDev[e] : elems[e].n;
  normalizeElement<256>(hist + (size_t)e * 256, total, probBits, table + (size_t)e * 256, sm);
}

// Stand-alone block encoder (see encodeRun in ans_encode.cuh). A warp's first run is static, further
// runs are handed out in order by an atomic counter (fetched one run ahead), so the resident warps
// always stream one contiguous window of the input and finish together.
// COMPACT: the warps also do k_coalesce's job once they are out of runs, as in the fused kernel (tickets for 8-block
// tiles in order, each waiting until the tile and its predecessors in the element are encoded): one launch less and
// the first tiles are compacted while the last blocks are still being encoded. Every run is owned by a running warp
// (all of them are claimed from the counter in this mode) and encoding warps never wait, so the waits end whether or
// not the whole grid is resident.
template <int PB, bool COMPACT>
__global__ void __launch_bounds__(kEncWarps * 32, kEncCtasPerSm) k_encode(const EncParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t wi = threadIdx.x >> 5;
  EncWarp w;
  w.init(smem + wi * kEncWarpSmem);
  const uint32_t numRuns = (p.tileEnd - p.tileBegin) << p.runShift;
  uint32_t run = blockIdx.x * kEncWarps + wi;
  if constexpr (COMPACT) {
    // every run is claimed, the first one too (p.totalWarps is 0): a run must belong to a RUNNING warp, because
    // warps wait for runs below; a statically assigned run of a CTA that is not resident yet (another stream's
    // kernel may hold its place) could be waited for by every resident warp
    if (w.lane == 0) run = atomicAdd(p.runCounter, 1u);
    run = __shfl_sync(0xffffffffu, run, 0);
  }
  while (run < numRuns) {
    uint32_t nextRun = 0;
    if (w.lane == 0) nextRun = p.totalWarps + atomicAdd(p.runCounter, 1u);
    encodeRun<PB, COMPACT>(p, w, run);
    run = __shfl_sync(0xffffffffu, nextRun, 0);
  }
  if constexpr (COMPACT) {
    const uint32_t tiles = p.tileEnd - p.tileBegin;
    uint32_t guess = 0;
    for (;;) {
      uint32_t tk = 0;
      if (w.lane == 0) tk = atomicAdd(p.coalClaim, 1u);
      tk = __shfl_sync(0xffffffffu, tk, 0);
      if (tk >= tiles) break;
      tk += p.tileBegin;
      const uint32_t ce = findElement(&p.a.elems[0].tileBase, sizeof(Elem) / 4, p.a.numElems, tk, guess);
      guess = ce;
      (void)waitTilesEncoded(p, ce, tk, w.lane, true);
      coalesceTileWarp(p, PB, ce, tk, w.lane, w.tabSmem, kEncWarpSmem);
    }
  }
}

// Compaction: tile = 8 consecutive blocks of one element, one warp per block, one CTA per tile.
// Every size is known when this kernel starts, so a tile's exclusive prefix is a plain sum: the
// per-group totals (groups of kGroupBlocks blocks, accumulated by k_encode with one atomicAdd per
// block) of the groups before it, plus the blocks of its own group before it (<= 2 * 256 loads).
__global__ void __launch_bounds__(256) k_coalesce(const EncParams p, int probBits) {
  __shared__ uint32_t sWarp[8];
  __shared__ uint32_t sWords[kTileBlocks];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const EncodeArgs& a = p.a;
  const uint32_t tk = p.tileBegin + blockIdx.x;
  const uint32_t e = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tk, 0);
  const Elem el = a.elems[e];
  const uint32_t t = tk - el.tileBase;
  const uint32_t n = a.nDev ? a.nDev[e] : el.n;
  const uint32_t nb = divUp(n, kBlockSyms);
  const uint32_t numTiles = max(1u, divUp(nb, kTileBlocks));
  if (t >= numTiles) return; // device-side size is smaller than the host bound
  uint8_t* ans = (uint8_t*)el.out;
  uint8_t* fbase = nullptr;
  if (a.floatType != kFtNone) {
    fbase = (uint8_t*)el.out + el.sizeBias;
    ans = fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
    if (a.round) ans += *(const uint32_t*)(fbase + 16);
  }
  // ---- exclusive prefix of this tile ----
  const uint32_t firstBlk = t * kTileBlocks;
  const uint32_t grp = firstBlk / kGroupBlocks;
  const uint32_t* groupSum = p.groupWords + el.groupBase;
  uint32_t part = 0;
  for (uint32_t g = tid; g < grp; g += 256) part += groupSum[g] & kGroupSizeMask; // (count bits: fused kernel only)
  for (uint32_t b = grp * kGroupBlocks + tid; b < firstBlk; b += 256)
    part += roundUp(p.blockWords[el.blockBase + b], 8);
  part = warpSum(part);
  if (lane == 0) sWarp[w] = part;
  const uint32_t blk = firstBlk + w;
  const uint32_t gb = el.blockBase + blk;
  uint32_t blkLen = 0, words = 0;
  if (blk < nb) {
    blkLen = min(kBlockSyms, n - blk * kBlockSyms);
    words = p.blockWords[gb];
  }
  if (lane == 0) sWords[w] = roundUp(words, 8);
  __syncthreads();
  uint32_t excl = 0, myOff = 0, tileWords = 0;
#pragma unroll
  for (uint32_t j = 0; j < 8; ++j) {
    excl += sWarp[j];
    uint32_t x = sWords[j];
    myOff += j < w ? x : 0u;
    tileWords += x;
  }
  if (blkLen) { // index entry and payload (dietgpu/ans/GpuANSEncode.cuh:591-623)
    const uint32_t startWord = excl + myOff;
    if (lane == 0) {
      uint2* bw = (uint2*)(ans + kAnsFixedBytes + 128u * nb);
      bw[blk] = make_uint2((blkLen << 16) | words, startWord);
    }
    uint4* dst = (uint4*)(ans + ansOverhead(nb) + 2 * (size_t)startWord);
    const uint4* src = (const uint4*)(p.scratch + (size_t)gb * kSlotBytes);
    const uint32_t nv = roundUp(words, 8) >> 3;
    for (uint32_t i = lane; i < nv; i += 32) dst[i] = src[i];
  }
  if (t == 0) { // pdf table (:568-573); zero for absent symbols and for an empty element
    const uint2 en = p.table[(size_t)e * 256 + tid];
    const uint32_t pdf = (n && en.x) ? (1u << probBits) - (en.y & 0xfffu) : 0u;
    ((uint16_t*)(ans + kAnsHeaderBytes))[tid] = (uint16_t)pdf;
  }
  if (t == numTiles - 1 && tid == 0) {
    const uint32_t totalWords = excl + tileWords;
    uint32_t cs = (a.useChecksum && a.checksum) ? a.checksum[e] : 0u;
    uint4* h = (uint4*)ans; // :533-565; fields the reference leaves undefined are written as 0
    h[0] = make_uint4(kAnsMagic, nb, n, totalWords);
    h[1] = make_uint4((uint32_t)probBits | (a.useChecksum ? 16u : 0u), cs, 0u, 0u);
    if (nb & 1u) ((uint2*)(ans + kAnsFixedBytes + 128u * nb))[nb] = make_uint2(0u, 0u);
    const uint32_t total = ansOverhead(nb) + 2u * totalWords;
    if (a.floatType != kFtNone && a.round == 0)
      *(uint32_t*)(fbase + 16) = total; // GpuFloatHeader2.firstCompSegmentBytes
    if (a.outSize) {
      if (a.round) a.outSize[e] += total;
      else
        a.outSize[e] = el.sizeBias + total +
                       (a.floatType != kFtNone ? kFloatHeaderBytes + floatRawBytes(a.floatType, n) : 0u);
    }
  }
}

int allocEncodeScratch(Temp& tmp, const EncodeArgs& a, uint32_t totalGroups, EncodeScratch& es, cudaStream_t s) {
  es.table = tmp.alloc<uint2>((size_t)a.numElems * 256);
  es.blockWords = tmp.alloc<uint32_t>(a.totalTiles);
  // one zeroed control array: group sums | run counters (one per possible launch) | fused-kernel control
  const size_t ctrlWords = (size_t)totalGroups + a.numElems + 1 + kFusedCtrlFixed + 3 * (size_t)a.numElems +
                           divUp(a.totalTiles, kTileBlocks) + a.numElems; // + fused kernel's per-tile sums (tiles <= blocks / 8 + elements)
  es.groupWords = tmp.alloc<uint32_t>(ctrlWords);
  es.scratch = (uint8_t*)tmp.alloc((size_t)a.totalTiles * kSlotBytes);
  if (tmp.failed()) return kErrCuda;
  es.runCounters = es.groupWords + totalGroups;
  es.fusedCtrl = es.runCounters + a.numElems + 1;
  DGB_CUDA(cudaMemsetAsync(es.groupWords, 0, ctrlWords * 4, s));
  return kOk;
}

void fillEncParams(EncParams& p, const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r) {
  p.a = a;
  p.table = es.table;
  p.scratch = es.scratch;
  p.blockWords = es.blockWords;
  p.groupWords = es.groupWords;
  p.runCounter = es.runCounters + r.e0;
  p.ready = nullptr;
  p.tileWords = nullptr;
  p.coalClaim = nullptr;
  p.tileBegin = r.tile0;
  p.tileEnd = r.tile1;
  p.totalWarps = 0;
  p.runShift = 2;
}

int launchAnsEncode(const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r, cudaStream_t s,
                    bool tablesReady) {
  if (r.e1 <= r.e0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  if (!tablesReady) {
    ProfScope ps__("k_normalize", s);
    k_normalize<<<r.e1 - r.e0, 256, 0, s>>>(a.elems, a.nDev, a.hist, a.probBits, es.table, r.e0);
  }
  DGB_CUDA(cudaGetLastError());

  EncParams p;
  fillEncParams(p, a, es, r);
  p.runShift = chooseRunShift(r.tile1 - r.tile0, (uint32_t)smCount() * kEncCtasPerSm * kEncWarps);
  const uint32_t runs = (r.tile1 - r.tile0) << p.runShift;
  uint32_t perSm = kEncCtasPerSm;
  if (DGB_KNOB("DGB_ENC_CTAS", 0) > 0) perSm = (uint32_t)DGB_KNOB("DGB_ENC_CTAS", 0); // experiment knob
  if (perSm > kEncCtasPerSm) perSm = kEncCtasPerSm; // never more CTAs than are resident at once: with in-kernel compaction warps wait for runs other CTAs hold
  const uint32_t grid = min(divUp(runs, kEncWarps), (uint32_t)smCount() * perSm);
  p.totalWarps = grid * kEncWarps;
  const uint32_t smemBytes = kEncWarps * kEncWarpSmem;
  // Compaction inside the encoder launch when this launch covers the whole work list (the grouped schedule's
  // ranges share one ticket counter and keep k_coalesce). Needs the per-tile sums the fused kernel uses.
  // Off by default (DGB_ENC_COMPACT=1 turns it on): a tile costs ONE warp about 10 us of dependent loads here against
  // a CTA of eight in k_coalesce, which pays only when later waves of blocks hide it (B200, separate / in-kernel:
  // bf16 1 x 64 MiB = 1024 tiles 0.0999 / 0.0949 ms, but 4 x 8 MiB = 512 tiles 0.0534 / 0.0591, ANS 16 x 1 MiB
  // 57.0 / 60.4 us, fp64 16 x 8 MiB 0.1408 / 0.1458, sparse 0.2106 / 0.2155) - and that gain needed each warp's
  // first run to be assigned statically, which is not safe when the grid is only partly resident (a second stream's
  // kernel on the same GPU): with every run claimed from the counter, as it has to be, 1 x 64 MiB takes 0.1048 ms.
  const bool whole = r.e0 == 0 && r.e1 == a.numElems && r.tile0 == 0;
  const bool compact = whole && es.fusedCtrl && DGB_KNOB("DGB_ENC_COMPACT", 0) != 0; // experiment knob
  if (compact) {
    p.tileWords = es.fusedCtrl + kFusedCtrlFixed + 3 * (size_t)a.numElems;
    p.coalClaim = es.fusedCtrl + kFusedCtrlCoal;
    p.totalWarps = 0; // all runs come from the counter
  }
  auto launch = [&](auto kern) -> int {
    static DeviceOnce onceByPb[6]; // the instantiations share one function-pointer type
    DeviceOnce& once = onceByPb[(a.probBits - 9) * 2 + (compact ? 1 : 0)];
    const int dev = currentDevice();
    if (once.first(dev)) {
      DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
      once.done(dev);
    }
    {
      ProfScope ps__("k_encode", s);
      kern<<<grid, kEncWarps * 32, smemBytes, s>>>(p);
    }
    DGB_CUDA(cudaGetLastError());
    return kOk;
  };
  int rc;
  switch (a.probBits) {
    case 9: rc = compact ? launch(k_encode<9, true>) : launch(k_encode<9, false>); break;
    case 10: rc = compact ? launch(k_encode<10, true>) : launch(k_encode<10, false>); break;
    default: rc = compact ? launch(k_encode<11, true>) : launch(k_encode<11, false>); break;
  }
  if (rc) return rc;
  return compact ? kOk : launchCoalesce(a, es, r, s);
}

int launchCoalesce(const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r, cudaStream_t s) {
  if (r.tile1 <= r.tile0) return kOk;
  EncParams p;
  fillEncParams(p, a, es, r);
  {
    ProfScope ps__("k_coalesce", s);
    k_coalesce<<<r.tile1 - r.tile0, 256, 0, s>>>(p, a.probBits);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

} // namespace dgb
