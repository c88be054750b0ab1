// Fused compress for sm_100a: float split (or, for the raw ANS codec, the byte histogram), table normalisation,
// rANS block encoding and compaction of the encoded blocks into the archive in ONE persistent launch.
//
// Why: measured on B200 the split is bound by shared-memory atomic throughput and the
// encoder by instruction issue; run back to back they add up, run on the same SMs at the same time they
// overlap. Stream-level overlap of separate launches does not work (the grids take turns); so the roles live in
// one grid:
//   * every CTA draws a per-SM ticket; the first `splitCtasPerSm` CTAs of each SM take the split role, the rest
//     the encode role, so every SM runs both kinds of work side by side;
//   * split CTAs claim work in order from an atomic counter (element-major, so elements complete one after the
//     other): two 32768-word tiles per claim for batches of many elements, 4096-word slices for a few huge
//     elements (all CTAs then sweep one contiguous front through memory, and all but one CTA per SM split,
//     because nothing can be encoded before the first table exists). They write raw bytes into the archive and
//     exponent bytes into the plane; the histogram stays in the CTA's shared-memory bins while its claims stay
//     inside one element and is added to the element's global bins when the CTA moves on; the CTA whose slice
//     count completes an element normalises its table and publishes the element's ready flag (release);
//   * encode warps claim runs of 2 blocks (1 for small batches) in the same order, wait (acquire) for the run's element, and encode
//     exactly as the stand-alone kernel does (encodeRun in ans_encode.cuh);
//   * a split CTA that finds no tile left becomes an encode CTA;
//   * compaction (what k_coalesce does for the stand-alone encoder) is the warps' last job: every finished block
//     adds its size and a count of one to its tile's and its group's sum (release); a warp that finds no run left
//     takes tickets for 8-block compaction tiles in order, waits until the tile and all tiles before it in the
//     element are encoded (their sums are then final) and copies the tile's blocks from their L2-resident
//     scratch slots to their final place (every run is by then owned by a running warp).
// Split CTAs never wait and encode warps wait only for split CTAs, so every flag arrives; the grid is sized to be
// co-resident.
#include <cstdio>
#include <cstdlib>
#include <string>
#include <vector>

#include "ans_encode.cuh"
#include "float_split.cuh"
#include "internal.h"
#include "prof.h"

namespace dgb {

struct FusedParams {
  SplitParams sp;
  EncParams ep;
  uint32_t* splitCounter; // zeroed
  uint32_t* smTicket;     // [512] zeroed
  uint32_t* tilesDone;    // [numElems] zeroed
  uint32_t* ready;        // [numElems] zeroed
  uint32_t* runsDone;     // [numElems] zeroed: finished encode runs per element
  uint32_t* coalClaim;    // zeroed: next compaction tile to hand out (nullptr: k_coalesce runs as a second launch)
  uint32_t totalTiles;
  uint32_t splitCtasPerSm;
  uint32_t sliceShift; // log2(slices per tile) the split role claims work in
  uint32_t splitClaim; // slices per claim beyond the fine prefix
  uint32_t fineSlices; // the first fineSlices slices are claimed one at a time (multiple of splitClaim)
  uint32_t* splitCounterB; // zeroed: claims beyond the fine prefix
  uint32_t tailSlices;     // the last tailSlices slices are claimed tailClaim at a time (short stragglers)
  uint32_t tailClaim;
  // nullable (DGB_FUSED_TRACE): per CTA {start, split end, first run end, end, smid/role, encode end}. The timestamps
  // stay in the product build on purpose: compiled out, ptxas allocates the split role differently and fp32
  // 1 x 1 GiB runs 0.706 instead of 0.666 ms, bf16 128 x 4 MiB 0.417 instead of 0.410 (B200, A/B of the two builds)
  unsigned long long* trace;
};

constexpr uint32_t kFusedThreads = kEncWarps * 32; // 128
constexpr uint32_t kFusedHistBytes = splitHistWords<kFusedThreads>() * 4;
constexpr uint32_t kFusedNormBytes = (256 + 16) * 4;
constexpr uint32_t kFusedSplitSmem = kSplitHistAlign + kFusedHistBytes + kFusedNormBytes; // bins aligned inside the block
constexpr uint32_t kFusedSmem = kEncWarps * kEncWarpSmem > kFusedSplitSmem ? kEncWarps * kEncWarpSmem : kFusedSplitSmem;

// Raw ANS (FT == kFtNone): the "split" role only counts bytes; the encode role reads the caller's input.
// Byte histogram of one 32768-byte tile into the CTA's per-warp bins.
template <int NT>
__device__ __forceinline__ void histTileBytes(const uint8_t* in, uint32_t n, uint32_t off, uint32_t sliceBytes,
                                              uint32_t* sh, uint32_t* ckAcc) {
  const uint32_t tid = threadIdx.x;
  const uint32_t wh = smemAddr(sh) + (tid >> 5) * kSplitHistAlign + (tid & (kSplitRep - 1)) * 4u; // see splitHistAddAt
  if (off >= n) return;
  const uint8_t* q = in + off;
  const uint8_t* end = q + min(sliceBytes, n - off);
  const uint8_t* a0 = (const uint8_t*)(((uintptr_t)q + 15) & ~(uintptr_t)15);
  if (a0 > end) a0 = end;
  const uint8_t* a1 = a0 + ((size_t)(end - a0) & ~(size_t)15);
  uint32_t ck = 0; // raw ANS: every byte counts
  if (tid < (uint32_t)(a0 - q)) {
    splitHistAddAt(wh, (uint32_t)q[tid] << kSplitSymShift);
    ck ^= q[tid];
  }
  if (tid >= 32 && tid - 32 < (uint32_t)(end - a1)) {
    splitHistAddAt(wh, (uint32_t)a1[tid - 32] << kSplitSymShift);
    ck ^= a1[tid - 32];
  }
  const uint4* v = (const uint4*)a0;
  const uint32_t nvec = (uint32_t)(a1 - a0) >> 4;
  auto addWord = [&](uint32_t x) { splitHistAdd4(wh, x); };
  uint32_t i = tid;
  for (; i + 3 * NT < nvec; i += 4 * NT) {
    uint4 x[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) x[k] = __ldg(v + i + NT * k);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      addWord(x[k].x); addWord(x[k].y); addWord(x[k].z); addWord(x[k].w);
      ck ^= x[k].x ^ x[k].y ^ x[k].z ^ x[k].w;
    }
  }
  for (; i < nvec; i += NT) {
    const uint4 x = __ldg(v + i);
    addWord(x.x); addWord(x.y); addWord(x.z); addWord(x.w);
    ck ^= x.x ^ x.y ^ x.z ^ x.w;
  }
  if (ckAcc) *ckAcc ^= ck;
}

template <int FT, int PB>
__global__ void __launch_bounds__(kFusedThreads, kEncCtasPerSm) k_compressFused(const __grid_constant__ FusedParams p) {
  if constexpr (FT != kFtNone) static_assert(FloatTraits<FT>::kPlanes == 1, "single exponent plane only");
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ uint32_t sBcast[3];
  const uint32_t tid = threadIdx.x;
  unsigned long long t0 = 0, t1 = 0, t2 = 0, tEnc = 0;
  uint32_t smidTrace = 0;
  if (p.trace && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));

  if (tid == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    sBcast[0] = atomicAdd(&p.smTicket[smid & 511u], 1u) < p.splitCtasPerSm ? 1u : 0u;
    smidTrace = smid;
  }
  __syncthreads();
  const bool splitRole = sBcast[0] != 0;
  __syncthreads();

  if (splitRole) {
    uint32_t* sh = (uint32_t*)splitHistBase(smem);
    uint32_t* normSm = sh + kFusedHistBytes / 4;
    for (uint32_t i = tid; i < kFusedHistBytes / 4; i += kFusedThreads) sh[i] = 0;
    // Work is claimed in order, in slices of a tile (p.sliceShift: log2 slices per tile). Counts stay in the CTA's
    // bins while its claims stay inside one element; they go to the element's global bins, followed by the
    // CTA's slice count, when the CTA moves on. The CTA whose count completes the element owns its table.
    const uint32_t sliceShift = p.sliceShift, sliceWords = kTileSyms >> sliceShift;
    const uint32_t totalSlices = p.totalTiles << sliceShift;
    uint32_t guess = 0, accElem = 0xffffffffu, accSlices = 0;
    uint32_t ck = 0; // this thread's share of the element's checksum when the pass is folded (SplitParams::checksumAcc)
    auto publish = [&]() { // all threads
      __syncthreads();
      splitHistFlush<kFusedThreads>(sh, p.sp.hist0 + (size_t)accElem * 256, tid);
      if (p.sp.checksumAcc) ckFlushWarp(ck, p.sp.checksumAcc + accElem);
      __threadfence(); // plane, raw bytes, histogram and checksum share of these slices before the slice count
      __syncthreads();
      if (tid == 0) {
        const uint32_t want = (p.sp.elems[accElem + 1].tileBase - p.sp.elems[accElem].tileBase) << sliceShift;
        const uint32_t done = atomicAdd(&p.tilesDone[accElem], accSlices) + accSlices;
        sBcast[1] = done == want ? 1u : 0u;
      }
      __syncthreads();
      if (sBcast[1]) { // this CTA completed the element: its histogram is final
        __threadfence();
        if (FT != kFtNone && p.sp.checksumAcc && tid == 0) { // ... and so is its checksum: into the float header
          uint8_t* fbase = (uint8_t*)p.sp.elems[accElem].out + p.sp.elems[accElem].sizeBias;
          ((uint32_t*)fbase)[3] = __ldcg(p.sp.checksumAcc + accElem);
        }
        normalizeElement<kFusedThreads>(p.sp.hist0 + (size_t)accElem * 256, p.sp.elems[accElem].n, PB,
                                        p.ep.table + (size_t)accElem * 256, normSm);
        __threadfence();
        __syncthreads();
        if (tid == 0) {
          DGB_TRACE_MAX(accElem, 0);
          stRelease(p.ready + accElem, 1u);
        }
      }
      __syncthreads(); // sBcast is rewritten after this point
      accSlices = 0;
    };
    // Claims: single slices over the fine prefix (the first elements complete early and in quick succession, so
    // the encode warps do not wait for a whole round of coarse claims), splitClaim slices at a time after it.
    // The next claim is requested before the current one is processed (thread 0 holds the ticket in a register).
    const uint32_t fineSlices = p.fineSlices, perTile = 1u << sliceShift;
    bool fine = fineSlices != 0;
    // Towards the end the claims shrink (decided from a possibly stale look at the counter: any size is valid), so
    // that no CTA is left with 65536 words to split while everybody else already encodes: encode warps take
    // runs in order and would all queue up behind that one element.
    uint32_t ticketLen = 1;
    auto claim = [&]() -> uint32_t { // thread 0 only
      if (fine) {
        const uint32_t t = atomicAdd(p.splitCounter, 1u);
        ticketLen = 1;
        if (t < fineSlices) return t;
        fine = false;
      }
      const uint32_t seen = fineSlices + ldRelaxed(p.splitCounterB);
      ticketLen = seen + p.tailSlices >= totalSlices ? p.tailClaim : p.splitClaim;
      return fineSlices + atomicAdd(p.splitCounterB, ticketLen);
    };
    uint32_t ticket = 0;
    if (tid == 0) ticket = claim();
    for (;;) {
      if (tid == 0) {
        sBcast[0] = ticket;
        sBcast[2] = ticketLen;
      }
      __syncthreads();
      const uint32_t first = sBcast[0];
      const uint32_t last = min(first + sBcast[2], totalSlices);
      __syncthreads();
      if (first >= totalSlices) break;
      if (tid == 0) ticket = claim();
      uint32_t sl = first;
      while (sl < last) {
        const uint32_t tile = sl >> sliceShift;
        const uint32_t e = findElement(&p.sp.elems[0].tileBase, sizeof(Elem) / 4, p.sp.numElems, tile, guess);
        guess = e;
        if (e != accElem) {
          if (accSlices) publish();
          accElem = e;
        }
        const Elem el = p.sp.elems[e];
        const uint32_t sub = sl & (perTile - 1u);
        const bool whole = sub == 0 && sl + perTile <= last; // a whole tile of this claim in one go
        const uint32_t words = whole ? kTileSyms : sliceWords;
        if constexpr (FT == kFtNone) {
          const uint32_t off = (tile - el.tileBase) * kTileSyms + sub * sliceWords;
          histTileBytes<kFusedThreads>((const uint8_t*)el.in, el.n, off, words, sh, p.sp.checksumAcc ? &ck : nullptr);
        } else {
          splitTile<FT, kFusedThreads>(p.sp, e, el, el.n, tile, sh, words, whole ? 0u : sub, &ck);
        }
        const uint32_t step = whole ? perTile : 1u;
        accSlices += step;
        sl += step;
      }
    }
    if (accSlices) publish();
    __syncthreads();
  }

  if (p.trace && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  // ---- encode role ----
  EncWarp w;
  w.init(smem + (tid >> 5) * kEncWarpSmem);
  const uint32_t numRuns = p.totalTiles << p.ep.runShift;
  uint32_t run = 0;
  if (w.lane == 0) run = atomicAdd(p.ep.runCounter, 1u);
  run = __shfl_sync(0xffffffffu, run, 0);
  while (run < numRuns) {
    uint32_t nextRun = 0;
    if (w.lane == 0) nextRun = atomicAdd(p.ep.runCounter, 1u);
    encodeRun<PB, true>(p.ep, w, run);
    if (p.trace && tid == 0 && t2 == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t2)); // end of first run
    run = __shfl_sync(0xffffffffu, nextRun, 0);
  }
  if (p.trace && tid == 0) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(tEnc));
  // Compaction: warps that ran out of runs take tickets for 8-block tiles in order and wait until the tile's
  // element is completely encoded. Every run is by now owned by a running warp that waits for nothing but split
  // CTAs, so the wait cannot deadlock; elements finish in order, so early finishers find work immediately.
  if (p.coalClaim) {
    uint32_t guess = 0;
    for (;;) {
      uint32_t tk = 0;
      if (w.lane == 0) tk = atomicAdd(p.coalClaim, 1u);
      tk = __shfl_sync(0xffffffffu, tk, 0);
      if (tk >= p.totalTiles) break;
      const uint32_t ce = findElement(&p.sp.elems[0].tileBase, sizeof(Elem) / 4, p.sp.numElems, tk, guess);
      guess = ce;
      (void)waitTilesEncoded(p.ep, ce, tk, w.lane, true); // the tile and everything before it in its element
      coalesceTileWarp(p.ep, PB, ce, tk, w.lane, w.tabSmem, kEncWarpSmem);
    }
  }
  if (p.trace && tid == 0) {
    unsigned long long t3;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t3));
    unsigned long long* r = p.trace + 6ull * blockIdx.x;
    r[0] = t0; r[1] = t1; r[2] = t2; r[3] = t3; r[4] = ((unsigned long long)smidTrace << 1) | (splitRole ? 1u : 0u);
    r[5] = tEnc;
  }
}

template <int FT, int PB>
static int launchFusedT(FusedParams& p, cudaStream_t s) {
  auto kern = k_compressFused<FT, PB>;
  // attribute + occupancy query once per (kernel, device)
  static DeviceOnce once;
  static int perSmByDev[kMaxDevices] = {};
  const int dev = currentDevice();
  if (once.first(dev)) {
    DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kFusedSmem));
    int q = 0;
    DGB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&q, kern, (int)kFusedThreads, kFusedSmem));
    __atomic_store_n(&perSmByDev[dev], q, __ATOMIC_RELAXED);
    once.done(dev);
  }
  int perSm = __atomic_load_n(&perSmByDev[dev], __ATOMIC_RELAXED);
  if (perSm < 2) return kErrInternal;
  if (perSm > (int)kEncCtasPerSm) perSm = (int)kEncCtasPerSm;
  uint32_t splitPerSm = 3; // measured on B200 (bf16, 128 x 4 MiB): 2: 0.507, 3: 0.424, 4: 0.442, 5: 0.443 ms
  const int knobSplitCtas = DGB_KNOB("DGB_FUSED_SPLIT_CTAS", 0); // experiment knob
  if (knobSplitCtas > 0) splitPerSm = (uint32_t)knobSplitCtas;
  if (splitPerSm < 1) splitPerSm = 1;
  if (splitPerSm >= (uint32_t)perSm) splitPerSm = (uint32_t)perSm - 1;
  // (few elements: making every CTA split first was measured slower - fp32 1 x 1 GiB 0.81 ms vs 0.72 ms)
  p.splitCtasPerSm = splitPerSm;
  // Granularity of the split claims. Many elements: two whole tiles per claim, so that a CTA's counts are flushed
  // once per 65536 words. Elements so large that every split CTA gets at least a tile's worth of each anyway:
  // 4096-word slices in order, so that the CTAs sweep one contiguous front through memory.
  p.sliceShift = 3;
  p.splitClaim = 16;
  p.fineSlices = 0;
  {
    const uint64_t splitCtas = (uint64_t)splitPerSm * (uint64_t)smCount();
    if (DGB_KNOB("DGB_SPLIT_FINE_TILES", 0) > 0) p.fineSlices = (uint32_t)DGB_KNOB("DGB_SPLIT_FINE_TILES", 0) << p.sliceShift; // experiment knob
    if ((uint64_t)p.totalTiles >= splitCtas * (uint64_t)p.sp.numElems) {
      p.fineSlices = 0xfffffff0u; // everything
      p.splitClaim = 16;
      // nothing can be encoded before the first table exists and there are few tables to wait for: all CTAs
      // but one per SM split (fp32 1 x 1 GiB on B200: 0.73 ms with 3 split CTAs per SM and whole tiles,
      // 0.69 ms with 9 and slices; 9 with whole tiles thrash DRAM: 0.81 ms)
      if (knobSplitCtas <= 0) p.splitCtasPerSm = (uint32_t)perSm - 1;
    }
    {
      uint32_t rounds = 2, tail = 8; // B200: single tiles over the last two rounds: fp32 64 x 16 MiB 0.557 -> 0.542 ms, bf16 equal; 2- and 4-slice claims are slower
      if (DGB_KNOB("DGB_SPLIT_TAIL_ROUNDS", -1) >= 0) rounds = (uint32_t)DGB_KNOB("DGB_SPLIT_TAIL_ROUNDS", -1); // experiment knobs
      if (DGB_KNOB("DGB_SPLIT_TAIL_CLAIM", 0) > 0) tail = (uint32_t)DGB_KNOB("DGB_SPLIT_TAIL_CLAIM", 0);
      p.tailSlices = rounds * (uint32_t)splitCtas * p.splitClaim;
      p.tailClaim = tail;
    }
    if (p.fineSlices > (p.totalTiles << p.sliceShift)) p.fineSlices = p.totalTiles << p.sliceShift;
    p.fineSlices = p.fineSlices / p.splitClaim * p.splitClaim;
  }
  // enough CTAs for the work, never more than fit at once (encode warps wait on split CTAs)
  p.ep.runShift = chooseRunShift(p.totalTiles, (uint32_t)smCount() * kEncCtasPerSm * kEncWarps);
  const uint32_t want = divUp(p.totalTiles << p.ep.runShift, kEncWarps) + p.totalTiles;
  const uint32_t grid = min(want, (uint32_t)smCount() * (uint32_t)perSm);
  static const char* tracePath = getenv("DGB_FUSED_TRACE"); // debugging aid: per-CTA timeline of the last launch
  p.trace = nullptr;
  if (tracePath) DGB_CUDA(cudaMalloc(&p.trace, 6ull * 8 * grid));
#ifdef DGB_TRACE_ELEMS
  {
    void* et = nullptr;
    DGB_CUDA(cudaGetSymbolAddress(&et, g_elemTrace));
    DGB_CUDA(cudaMemsetAsync(et, 0, sizeof(unsigned long long) * 6 * 1024, s));
  }
#endif
  {
    ProfScope ps__("k_compressFused", s);
    kern<<<grid, kFusedThreads, kFusedSmem, s>>>(p);
  }
  DGB_CUDA(cudaGetLastError());
  if (tracePath) {
    std::vector<unsigned long long> h(6ull * grid);
    DGB_CUDA(cudaMemcpyAsync(h.data(), p.trace, h.size() * 8, cudaMemcpyDeviceToHost, s));
    DGB_CUDA(cudaStreamSynchronize(s));
    cudaFree(p.trace);
    if (FILE* f = fopen(tracePath, "w")) {
      for (uint32_t i = 0; i < grid; ++i)
        fprintf(f, "%u %llu %llu %llu %llu %llu %llu\n", i, h[6 * i], h[6 * i + 1], h[6 * i + 2], h[6 * i + 3], h[6 * i + 4],
                h[6 * i + 5]);
      fclose(f);
    }
#ifdef DGB_TRACE_ELEMS
    std::vector<unsigned long long> he(6 * 1024);
    DGB_CUDA(cudaMemcpyFromSymbol(he.data(), g_elemTrace, he.size() * 8));
    if (FILE* f = fopen((std::string(tracePath) + ".elems").c_str(), "w")) {
      for (uint32_t e = 0; e < p.sp.numElems && e < 1024u; ++e)
        fprintf(f, "%u %llu %llu %llu %llu %llu %llu\n", e, he[6 * e], ~he[6 * e + 1], he[6 * e + 2], ~he[6 * e + 3],
                he[6 * e + 4], he[6 * e + 5]);
      fclose(f);
    }
#endif
  }
  return kOk;
}

template <int FT>
static int launchFusedPb(FusedParams& p, int pb, cudaStream_t s) {
  switch (pb) {
    case 9: return launchFusedT<FT, 9>(p, s);
    case 10: return launchFusedT<FT, 10>(p, s);
    default: return launchFusedT<FT, 11>(p, s);
  }
}

int launchFusedCompress(int ft, const EncodeArgs& a, const EncodeScratch& es, uint32_t totalTiles,
                             const uint32_t* checksum, int useChecksum, cudaStream_t s, bool foldChecksum) {
  if (a.numElems == 0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  if (a.nDev || ft == kFtF64) return kErrInvalidArg;
  const WorkRange all{0, a.numElems, 0, totalTiles, 0, a.totalTiles};
  FusedParams p;
  fillEncParams(p.ep, a, es, all);
  p.sp.elems = a.elems;
  p.sp.numElems = a.numElems;
  p.sp.tileBegin = 0;
  p.sp.tileEnd = totalTiles;
  p.sp.tilesPerCta = 1;
  p.sp.unitShift = 0;
  p.sp.roundRobin = 0;
  p.sp.minElemTiles = 0;
  p.sp.nDev = nullptr;
  p.sp.denseIn = nullptr;
  p.sp.plane0 = (uint8_t*)a.planeBase;
  p.sp.plane1 = nullptr;
  p.sp.hist0 = (uint32_t*)a.hist;
  p.sp.hist1 = nullptr;
  p.sp.checksum = checksum;
  p.sp.useChecksum = useChecksum;
  p.sp.checksumAcc = foldChecksum ? (uint32_t*)checksum : nullptr;
  p.sp.normTilesDone = nullptr;
  p.sp.normTable0 = p.sp.normTable1 = nullptr;
  p.sp.normProbBits = 0;
  p.splitCounter = es.fusedCtrl;
  p.splitCounterB = es.fusedCtrl + kFusedCtrlSplitB;
  p.smTicket = es.fusedCtrl + 1;
  p.tilesDone = es.fusedCtrl + kFusedCtrlFixed;
  p.ready = p.tilesDone + a.numElems;
  p.runsDone = p.ready + a.numElems;
  p.ep.tileWords = p.runsDone + a.numElems;
  const bool separateCoalesce = DGB_KNOB("DGB_FUSED_COALESCE", 1) == 0;
  p.coalClaim = separateCoalesce ? nullptr : es.fusedCtrl + kFusedCtrlCoal; // experiment knob
  p.ep.ready = p.ready;
  p.totalTiles = totalTiles;
  int rc;
  switch (ft) {
    case kFtNone: rc = launchFusedPb<kFtNone>(p, a.probBits, s); break;
    case kFtF16: rc = launchFusedPb<kFtF16>(p, a.probBits, s); break;
    case kFtBF16: rc = launchFusedPb<kFtBF16>(p, a.probBits, s); break;
    case kFtF32: rc = launchFusedPb<kFtF32>(p, a.probBits, s); break;
    default: return kErrInvalidArg;
  }
  if (rc) return rc;
  return p.coalClaim ? kOk : launchCoalesce(a, es, all, s);
}

} // namespace dgb
