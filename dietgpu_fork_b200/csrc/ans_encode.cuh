// Device-side building blocks of the ANS encoder, shared by the stand-alone kernels (ans_encode.cu) and the
// fused split + normalise + encode kernel (fused.cu): table normalisation, the rANS row step and the
// warp-level "encode one run of blocks" routine.
#pragma once
#include <cstdlib>

#include "common.cuh"
#include "internal.h"
#include "knobs.h"

namespace dgb {

// ---------------------------------------------------------------------------------------------
// Normalisation: counts -> pdf summing to 2^probBits, bit-exact with the reference
// (dietgpu/ans/GpuANSStatistics.cuh:178-367). One CTA of NT threads per element; thread t owns the
// symbols t, t + NT, ... throughout. Instead of physically sorting we compute each symbol's rank in
// the descending order of (q << 16 | s): the reference's "positions" are these ranks.
//
// Encode-table entry (ours, 8 bytes): x = magic m, y = (2^probBits - pdf) | shift << 12 | cdf' << 17,
// such that for every state v the encoder can hold when it divides (v < pdf << (31 - probBits), i.e.
// after renormalisation)
//     floor(v / pdf) == umulhi(v, m) >> shift
// and the update is v' = q * (2^probBits - pdf) + v + cdf'. For pdf >= 2: l = ceil(log2 pdf),
// m = ceil(2^(31 + l) / pdf) <= 2^32 - 1 (2^31 for a power of two), shift = l - 1; the error term
// e = m * pdf - 2^(31 + l) < pdf gives v * e < pdf^2 << (31 - probBits) <= 2^(31 + l), so the quotient
// is exact. pdf == 1: m = 2^32 - 1 yields q = v - 1 (v >= 32 here), compensated by
// cdf' = cdf + 2^probBits - 1; otherwise cdf' = cdf. Storing K - pdf in the low bits serves both the
// renormalisation test (v >= pdf << (31 - probBits)  <=>  bit 31 of v + (y << (31 - probBits)), the
// multiply drops every other field) and the update; cdf' sits in the high bits, where a single
// shift-right-add (LEA.HI) extracts and adds it. Absent symbols (pdf 0) get an all-zero entry.
//
// sm: at least 256 + 16 words of shared memory. All NT threads of the CTA must call it.
// ---------------------------------------------------------------------------------------------
template <int NT>
__device__ __forceinline__ uint32_t normBlockSum(uint32_t v, uint32_t* sRed, uint32_t lane, uint32_t w) {
  constexpr int NW = NT / 32;
  v = warpSum(v);
  __syncthreads(); // previous readers of sRed are done
  if (lane == 0) sRed[w] = v;
  __syncthreads();
  uint32_t sum = 0;
#pragma unroll
  for (int i = 0; i < NW; ++i) sum += sRed[i];
  return sum;
}

template <int NT>
__device__ void normalizeElement(const uint32_t* hist, uint32_t total, int probBits, uint2* table, uint32_t* sm) {
  constexpr int S = 256 / NT;
  constexpr int NW = NT / 32;
  uint32_t* sKey = sm;
  uint32_t* sRed = sm + 256;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  if (total == 0) { // reference leaves the table untouched (:192-195); we define it as zero
#pragma unroll
    for (int k = 0; k < S; ++k) table[tid + k * NT] = make_uint2(0, 0);
    return;
  }
  const uint32_t weight = 1u << probBits;
  uint32_t q[S], key[S], pos[S];
  uint32_t mySum = 0;
  __syncthreads(); // sm may still be in use by the caller
#pragma unroll
  for (int k = 0; k < S; ++k) {
    const uint32_t s = tid + k * NT;
    const uint32_t count = __ldcg(hist + s);
    // :205-221 — fp32 divide (RN) then multiply (RN), truncation
    const float ratio = __fdiv_rn((float)count, (float)total);
    q[k] = (uint32_t)__fmul_rn((float)weight, ratio);
    if (count > 0 && q[k] == 0) q[k] = 1;
    key[k] = (q[k] << 16) | s;
    sKey[s] = key[k];
    mySum += q[k];
  }
  const int sum = (int)normBlockSum<NT>(mySum, sRed, lane, w); // also publishes sKey
  // rank in descending key order == position after the reference's SortDescending (:229-252)
#pragma unroll
  for (int k = 0; k < S; ++k) {
    uint32_t c = 0;
#pragma unroll 16
    for (int j = 0; j < 256; ++j) c += (sKey[j] > key[k]);
    pos[k] = c;
  }
  int diff = (int)weight - sum;
  if (diff > 0) {
    // :258-273 — the test is on the symbol id, not on the rank
    while (diff > 0) {
      const int kk = diff < 256 ? diff : 256;
#pragma unroll
      for (int k = 0; k < S; ++k)
        if ((int)(tid + k * NT) < kk) q[k] += 1;
      diff -= kk;
    }
  } else if (diff < 0) {
    // :274-315 — take one from the smallest entries that are still > 1, by position
    diff = -diff;
    while (diff > 0) {
      uint32_t c = 0;
#pragma unroll
      for (int k = 0; k < S; ++k) c += q[k] > 1;
      const int g = (int)normBlockSum<NT>(c, sRed, lane, w);
      const int kk = diff < g ? diff : g;
      if (kk <= 0) break;
#pragma unroll
      for (int k = 0; k < S; ++k)
        if ((int)pos[k] >= g - kk && (int)pos[k] < g) q[k] -= 1;
      diff -= kk;
    }
  }
  // exclusive scan over symbols -> cdf (:317-341); symbols of group k precede those of group k + 1
  uint32_t base = 0;
#pragma unroll
  for (int k = 0; k < S; ++k) {
    uint32_t inc = q[k];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (uint32_t)o) inc += t;
    }
    __syncthreads(); // sRed reuse
    if (lane == 31) sRed[w] = inc;
    __syncthreads();
    uint32_t cdf = base + inc - q[k];
    uint32_t groupTotal = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
      const uint32_t x = sRed[i];
      cdf += (uint32_t)i < w ? x : 0u;
      groupTotal += x;
    }
    base += groupTotal;

    uint2 entry = make_uint2(0, 0);
    if (q[k]) {
      const uint32_t l = 32 - __clz(q[k] - 1); // q == 1 -> clz(0) == 32 -> 0
      uint32_t m = 0xffffffffu, shift = 0, cdfAdj = cdf + weight - 1;
      if (q[k] > 1) {
        m = (uint32_t)(((1ull << (31 + l)) + q[k] - 1) / q[k]);
        shift = l - 1;
        cdfAdj = cdf;
      }
      entry = make_uint2(m, (weight - q[k]) | (shift << 12) | (cdfAdj << 17));
    }
    table[tid + k * NT] = entry;
  }
}

// ---------------------------------------------------------------------------------------------
// Block encoder.
// ---------------------------------------------------------------------------------------------
struct EncParams {
  EncodeArgs a;
  uint2* table;         // [numElems][256]
  uint8_t* scratch;     // [totalBlocks][kSlotBytes]
  uint32_t* blockWords; // [totalBlocks] words emitted per block
  uint32_t* groupWords; // [totalGroups] sum of roundUp(words, 8) per group of 256 blocks, zeroed
  uint32_t* tileWords;  // fused kernel: [totalTiles] the same sums per 8-block tile, zeroed (nullptr otherwise)
  uint32_t* runCounter; // zeroed; hands out runs beyond each warp's first one
  uint32_t totalWarps;
  uint32_t tileBegin, tileEnd; // range of 8-block tiles handled by this launch
  uint32_t runShift;           // log2(runs per tile): runs are 8 >> runShift consecutive blocks
  // fused kernel only: per-element flag, non-zero once the element's exponent plane and table are complete
  const uint32_t* ready;
  // stand-alone encoder with in-kernel compaction (k_encode<PB, true>): zeroed ticket counter for the launch's tiles;
  // tileWords is then set as in the fused kernel, ready stays null (the tables exist before the launch)
  uint32_t* coalClaim;
};

void fillEncParams(EncParams& p, const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r);

constexpr uint32_t kEncWarps = 4;
constexpr uint32_t kEncCtasPerSm = 10;
constexpr uint32_t kGroupBlocks = 256;           // compaction: blocks per size group
// fused kernel: the per-group / per-tile size sums carry a count of finished blocks above the size bits
// (a group holds < 2^20 words, a tile < 2^24)
constexpr uint32_t kGroupCountShift = 20, kGroupCountOne = 1u << kGroupCountShift, kGroupSizeMask = kGroupCountOne - 1u;
constexpr uint32_t kTileCountShift = 24, kTileCountOne = 1u << kTileCountShift, kTileSizeMask = kTileCountOne - 1u;
constexpr uint32_t kEncChunkRows = 16;            // rows per staged symbol chunk (512 symbols, one 16-byte load per lane)
constexpr uint32_t kEncChunk = kEncChunkRows * 32;
// Run = consecutive blocks one warp encodes as one stream: 2 when there is enough work to fill the GPU with such
// runs, 1 for small batches. (4-block runs amortise the per-run start-up slightly better but leave a long
// tail: a run then takes ~100 us of wall time with 40 warps per SM; fused kernel 0.42 ms vs 0.386 ms.)
inline uint32_t chooseRunShift(uint32_t totalTiles, uint32_t warpsWanted) {
  uint32_t shift = 2;
  if (DGB_KNOB("DGB_RUN_SHIFT", -1) >= 0) return (uint32_t)DGB_KNOB("DGB_RUN_SHIFT", -1); // experiment knob
  // single blocks until every warp gets about four runs: with fewer, the last round of 2-block runs leaves most
  // warps idle (B200, bf16, fused, 2- vs 1-block runs: 8 x 16 MiB 0.175 / 0.160 ms, 16 x 8 MiB 0.167 / 0.155,
  // 32 x 8 MiB 0.253 / 0.241, 4 x 64 MiB 0.266 / 0.273, 128 x 4 MiB 0.422 / 0.428)
  while (shift < 3 && ((uint64_t)totalTiles << shift) < 4ull * warpsWanted) ++shift;
  return shift;
}
constexpr uint32_t kBlockChunks = kBlockSyms / kEncChunk;
constexpr uint32_t kEncTableBytes = 256 * 8;     // per-warp copy of the element's encode table
constexpr uint32_t kEncStageBytes = 2304;
constexpr uint32_t kEncFlushAt = kEncStageBytes - kEncChunkRows * 64; // room for one more chunk, worst case
constexpr uint32_t kEncRingBytes = 2 * kEncChunk; // double-buffered symbol chunks
constexpr uint32_t kSlotBytes = kStageWords * 2; // scratch bytes per block (worst case, probBits 11)
constexpr uint32_t kEncWarpSmem = kEncTableBytes + kEncStageBytes + kEncRingBytes;

// One rANS step for 32 symbols (dietgpu/ans/GpuANSEncode.cuh:49-90). e = table entry
// {m, (K - pdf) | shift << 12 | cdf' << 17}; outPtr = shared-memory byte address of the next free
// staging word. Written in PTX to pin the instruction selection.
template <int PB>
__device__ __forceinline__ void encStep(uint32_t& state, uint2 e, uint32_t& outPtr, uint32_t ltMask) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b32 sum, vote, pre, addr, t, q, c, k;\n"
      "mad.lo.u32 sum, %3, %5, %0;\n"          // ((K - pdf) << (31 - PB)) + state: bit 31 <=> renormalise
      "setp.lt.s32 p, sum, 0;\n"
      "vote.sync.ballot.b32 vote, p, 0xffffffff;\n"
      "and.b32 pre, vote, %4;\n"
      "popc.b32 pre, pre;\n"
      "mad.lo.u32 addr, pre, 2, %1;\n"
      "@p st.shared.u16 [addr], %0;\n"
      "@p shr.u32 %0, %0, 16;\n"
      "popc.b32 t, vote;\n"
      "mad.lo.u32 %1, t, 2, %1;\n"
      "mul.hi.u32 q, %0, %2;\n"
      "shr.u32 c, %3, 12;\n"
      "shf.r.wrap.b32 q, q, 0, c;\n"           // floor(state / pdf)
      "and.b32 k, %3, 0xfff;\n"
      "shr.u32 c, %3, 17;\n"
      "add.u32 c, c, %0;\n"                    // state + cdf'
      "mad.lo.u32 %0, q, k, c;\n"              // q * (K - pdf) + state + cdf'
      "}\n"
      : "+r"(state), "+r"(outPtr)
      : "r"(e.x), "r"(e.y), "r"(ltMask), "n"(1u << (31 - PB))
      : "memory");
}

// partial last row: lanes >= rem neither emit nor update (:94-136)
template <int PB>
__device__ __forceinline__ void encStepPartial(uint32_t& state, uint2 e, uint32_t& outPtr, uint32_t ltMask,
                                               bool valid) {
  const bool wr = valid && (int32_t)((e.y << (31 - PB)) + state) < 0;
  const uint32_t vote = __ballot_sync(0xffffffffu, wr);
  if (wr) {
    uint32_t addr = outPtr + 2u * __popc(vote & ltMask);
    asm volatile("st.shared.u16 [%0], %1;" ::"r"(addr), "r"(state) : "memory");
    state >>= 16;
  }
  outPtr += 2u * __popc(vote);
  const uint32_t q = __funnelshift_r(__umulhi(state, e.x), 0u, e.y >> 12);
  const uint32_t ns = q * (e.y & 0xfffu) + state + (e.y >> 17);
  state = valid ? ns : state;
}

// 16 bytes of symbols per lane = one 512-symbol chunk per warp. The data is used once: the load bypasses L1
// (.cg), which also makes it safe in the fused kernel where another SM wrote the plane moments ago.
__device__ __forceinline__ uint4 ldChunkVec(const uint8_t* p) {
  uint4 v;
  asm volatile("ld.global.cg.L2::256B.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ uint32_t ldSymByte(const uint8_t* p) {
  uint32_t v;
  asm volatile("ld.global.cg.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
// Lane's 16 bytes [16 * lane, 16 * lane + 16) of the chunk starting at `base` (chunk-relative validity:
// `avail` bytes). Fast path: aligned and fully inside; otherwise byte loads, zero beyond the end.
__device__ __forceinline__ uint4 ldChunk(const uint8_t* base, uint32_t avail, uint32_t lane, bool aligned) {
  const uint8_t* q = base + lane * 16;
  if (aligned && lane * 16 + 16 <= avail) return ldChunkVec(q);
  uint32_t w[4] = {0, 0, 0, 0};
#pragma unroll
  for (int k = 0; k < 16; ++k)
    if (lane * 16 + k < avail) w[k >> 2] |= ldSymByte(q + k) << (8 * (k & 3));
  return make_uint4(w[0], w[1], w[2], w[3]);
}
__device__ __forceinline__ uint32_t ldsSym(uint32_t addr) {
  uint32_t v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}

// table entry of symbol `sym` from the warp's shared-memory copy
__device__ __forceinline__ uint2 ldEnt(uint32_t tabAddr, uint32_t sym) {
  uint2 e;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(tabAddr + sym * 8));
  return e;
}

__device__ __forceinline__ uint32_t ldAcquire(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ldRelaxed(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void redRelease(uint32_t* p, uint32_t v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void redRelaxed(uint32_t* p, uint32_t v) {
  asm volatile("red.relaxed.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void stRelease(uint32_t* p, uint32_t v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---------------------------------------------------------------------------------------------
#ifdef DGB_TRACE_ELEMS
// Debugging aid (make EXTRA=-DDGB_TRACE_ELEMS, never in the shipped build): per element of the fused kernel's last
// launch {ready, ~first run start, last block end, ~first compaction start, last compaction end, summed wait ns}
// in globaltimer ns; the "first" slots hold the complement so that one zero-initialised atomicMax serves all.
static __device__ unsigned long long g_elemTrace[6 * 1024];
__device__ __forceinline__ unsigned long long traceNow() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define DGB_TRACE_MAX(e, k) do { if ((e) < 1024u) atomicMax(&g_elemTrace[6 * (e) + (k)], traceNow()); } while (0)
#define DGB_TRACE_MIN(e, k) do { if ((e) < 1024u) atomicMax(&g_elemTrace[6 * (e) + (k)], ~traceNow()); } while (0)
#else
#define DGB_TRACE_MAX(e, k) do { } while (0)
#define DGB_TRACE_MIN(e, k) do { } while (0)
#endif

// Warp-level encoder. Work unit = run of 8 >> runShift consecutive blocks of one element (half of an
// 8-block tile for large batches): one element lookup and one uninterrupted symbol stream per run.
//   symbols: one byte load per lane and symbol row straight from global memory, requested 8 rows
//            (one batch) before the dependent table load
//   table:   the element's 2 KB encode table in a per-warp shared-memory copy, read 2 rows ahead
//   rows:    rANS step per 32 symbols, ballot/popc compaction into a 2.5 KB per-warp staging buffer
//   flush:   the staged words go, as 16-byte vectors, to the block's slot of a scratch area when the
//            block ends (typical data: once per block) or when the buffer could overflow within the
//            next batch; final lane states go straight into the archive, the word count into a
//            per-block size array
// ---------------------------------------------------------------------------------------------
struct EncWarp {
  uint8_t* tabSmem;
  uint8_t* stage;
  uint32_t tabAddr, stageAddr, ringAddr, ltMask, lane;
  // element cache
  uint32_t curElem, n, nb, blockBase, tileBase, nextTileBase, guess;
  uint32_t* groupSum;
  const uint8_t* elemSrc;
  uint8_t* ans;

  __device__ __forceinline__ void init(uint8_t* warpSmem) {
    tabSmem = warpSmem;
    stage = warpSmem + kEncTableBytes;
    tabAddr = smemAddr(tabSmem);
    stageAddr = smemAddr(stage);
    ringAddr = smemAddr(warpSmem + kEncTableBytes + kEncStageBytes);
    ltMask = laneMaskLt();
    lane = laneId();
    curElem = 0xffffffffu;
    n = nb = blockBase = tileBase = nextTileBase = guess = 0;
    groupSum = nullptr;
    elemSrc = nullptr;
    ans = nullptr;
  }
};

// FUSED: the element's plane and table are produced by other CTAs of the same launch; wait for its flag.
template <int PB, bool FUSED>
__device__ __forceinline__ void encodeRun(const EncParams& p, EncWarp& w, uint32_t run) {
  const EncodeArgs& a = p.a;
  const uint32_t lane = w.lane;
  uint8_t* const stage = w.stage;
  const uint32_t stageAddr = w.stageAddr, tabAddr = w.tabAddr, ltMask = w.ltMask;
  const uint32_t runBlocks = kTileBlocks >> p.runShift;
  const uint32_t tile = p.tileBegin + (run >> p.runShift);
  if (w.curElem == 0xffffffffu || tile >= w.nextTileBase || tile < w.tileBase) {
    w.curElem = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tile, w.guess);
    w.guess = w.curElem;
    const Elem el = a.elems[w.curElem];
    w.nextTileBase = a.elems[w.curElem + 1].tileBase;
    w.tileBase = el.tileBase;
    w.blockBase = el.blockBase;
    w.groupSum = p.groupWords + el.groupBase;
    w.elemSrc = a.planeBase ? a.planeBase + el.planeOff : (const uint8_t*)el.in;
    if (FUSED && p.ready) {
      if (lane == 0) { // relaxed polling with growing back-off (an acquire load per poll would flush L1 each time)
#ifdef DGB_TRACE_ELEMS
        const unsigned long long tw0 = traceNow();
#endif
        uint32_t ns = 200;
        while (ldRelaxed(p.ready + w.curElem) == 0) {
          __nanosleep(ns);
          if (ns < 3200) ns *= 2;
        }
        (void)ldAcquire(p.ready + w.curElem);
#ifdef DGB_TRACE_ELEMS
        if (w.curElem < 1024u) atomicAdd(&g_elemTrace[6 * w.curElem + 5], traceNow() - tw0);
#endif
      }
      __syncwarp();
    }
    { // the element's 2 KB table -> this warp's shared-memory copy
      const uint4* g = (const uint4*)(p.table + (size_t)w.curElem * 256);
      __syncwarp();
#pragma unroll
      for (int i = 0; i < 4; ++i) ((uint4*)w.tabSmem)[lane + 32 * i] = __ldcg(g + lane + 32 * i);
      __syncwarp();
    }
    w.n = a.nDev ? a.nDev[w.curElem] : el.n;
    w.nb = divUp(w.n, kBlockSyms);
    w.ans = (uint8_t*)el.out;
    if (a.floatType != kFtNone) {
      const uint8_t* fbase = (const uint8_t*)el.out + el.sizeBias;
      w.ans = (uint8_t*)fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, w.n);
      if (a.round) w.ans += *(const uint32_t*)(fbase + 16); // firstCompSegmentBytes
    }
  }
  uint32_t blk = (tile - w.tileBase) * kTileBlocks + (run & ((1u << p.runShift) - 1u)) * runBlocks;
  if (blk >= w.nb) return; // empty element, or the device-side size is smaller than the host bound
  if (FUSED && lane == 0) DGB_TRACE_MIN(w.curElem, 1);
  const uint32_t n = w.n;
  const uint32_t runSyms = min(n, (blk + runBlocks) * kBlockSyms) - blk * kBlockSyms;
  const uint8_t* src = w.elemSrc + (size_t)blk * kBlockSyms;

  uint32_t state = kStateMin, outPtr = stageAddr, flushedVecs = 0;
  // staged words -> scratch slot of block `blk`. final: pad with zero words to a multiple of 8 words,
  // publish the sizes and the lane states, reset for the next block; otherwise whole 16-byte vectors
  // go out and the remainder moves to the front of the buffer
  auto flush = [&](bool final) {
    const uint32_t gb = w.blockBase + blk;
    uint4* slot = (uint4*)(p.scratch + (size_t)gb * kSlotBytes);
    uint32_t bytes = outPtr - stageAddr;
    uint32_t words = 0, paddedWords = 0;
    __syncwarp();
    if (final) {
      const uint32_t padded = roundUp(bytes, 16);
      if (bytes + 2 * lane < padded) *(uint16_t*)(stage + bytes + 2 * lane) = 0;
      __syncwarp();
      words = flushedVecs * 8 + (bytes >> 1);
      paddedWords = flushedVecs * 8 + (padded >> 1);
      bytes = padded;
    }
    const uint32_t nv = bytes >> 4;
    for (uint32_t i = lane; i < nv; i += 32) slot[flushedVecs + i] = ((const uint4*)stage)[i];
    if (final) {
      // final lane states go straight to their place in the archive (:584-589)
      ((uint32_t*)(w.ans + kAnsFixedBytes))[blk * 32 + lane] = state;
      __syncwarp(); // the staging buffer is reused by the next block; the slot is complete
      if (lane == 0) {
        p.blockWords[gb] = words;
        if (FUSED) {
          DGB_TRACE_MAX(w.curElem, 2);
          // sizes and, in the high bits, the number of finished blocks (release: slot and size first); the
          // in-kernel compaction of a tile starts as soon as these counts say its predecessors are encoded
          // (one release fence for both counters: two red.release cost two waits for the block's stores)
          asm volatile("fence.acq_rel.gpu;" ::: "memory");
          redRelaxed(&w.groupSum[blk / kGroupBlocks], paddedWords + kGroupCountOne);
          redRelaxed(p.tileWords + w.tileBase + blk / kTileBlocks, paddedWords + kTileCountOne);
        } else {
          atomicAdd(&w.groupSum[blk / kGroupBlocks], paddedWords);
        }
      }
      state = kStateMin;
      outPtr = stageAddr;
      flushedVecs = 0;
      ++blk;
    } else {
      const uint32_t remBytes = bytes & 15u;
      uint32_t carry = 0;
      if (2 * lane < remBytes) carry = *(const uint16_t*)(stage + nv * 16 + 2 * lane);
      __syncwarp();
      if (2 * lane < remBytes) *(uint16_t*)(stage + 2 * lane) = (uint16_t)carry;
      __syncwarp();
      flushedVecs += nv;
      outPtr = stageAddr + remBytes;
    }
  };

  // Symbols: 512-symbol chunks, one 16-byte global load per lane, requested two chunks ahead of use and parked
  // in a register; at the start of chunk c the data of chunk c + 1 goes into the other half of a 1 KB
  // per-warp ring, so the row loop only sees shared-memory loads: symbol bytes 4 rows ahead, table entries
  // 2 rows ahead of their step. The stream runs through block boundaries; only the state and the staging
  // buffer are reset there.
  const bool aligned = (((uintptr_t)src) & 15u) == 0;
  const uint32_t numChunks = divUp(runSyms, kEncChunk);
  const uint32_t ringAddr = w.ringAddr;
  uint4 x = ldChunk(src, runSyms, lane, aligned);
  __syncwarp(); // ring free (previous run)
  asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(ringAddr + lane * 16), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
  if (numChunks > 1) x = ldChunk(src + kEncChunk, runSyms - kEncChunk, lane, aligned);
  for (uint32_t c = 0; c < numChunks; ++c) {
    const uint32_t cur = ringAddr + (c & 1u) * kEncChunk + lane;
    const uint32_t nxt = ringAddr + ((c & 1u) ^ 1u) * kEncChunk + lane;
    __syncwarp(); // every lane is done with the rows of chunk c - 1 (the half being overwritten)
    if (c + 1 < numChunks) {
      asm volatile("st.shared.v4.u32 [%0], {%1, %2, %3, %4};" ::"r"(nxt - lane + lane * 16), "r"(x.x), "r"(x.y), "r"(x.z), "r"(x.w) : "memory");
      if (c + 2 < numChunks) x = ldChunk(src + (size_t)(c + 2) * kEncChunk, runSyms - (c + 2) * kEncChunk, lane, aligned);
    }
    __syncwarp();
    const uint32_t len = min(kEncChunk, runSyms - c * kEncChunk);
    if (len == kEncChunk) {
      uint32_t sym[4];
      uint2 ent[2];
      ent[0] = ldEnt(tabAddr, ldsSym(cur));
      ent[1] = ldEnt(tabAddr, ldsSym(cur + 32));
      sym[2] = ldsSym(cur + 64);
      sym[3] = ldsSym(cur + 96);
      sym[0] = ldsSym(cur + 128);
      sym[1] = ldsSym(cur + 160);
#pragma unroll
      for (int r = 0; r < 16; ++r) {
        encStep<PB>(state, ent[r & 1], outPtr, ltMask);
        if (r + 2 < 16) ent[r & 1] = ldEnt(tabAddr, sym[(r + 2) & 3]);
        if (r + 6 < 16) sym[(r + 2) & 3] = ldsSym(cur + (r + 6) * 32);
      }
      if ((c & (kBlockChunks - 1)) == kBlockChunks - 1) flush(true);
      else if (outPtr - stageAddr >= kEncFlushAt) flush(false);
    } else { // last chunk of an element: leftover rows and the partial row
      const uint32_t rows = len >> 5, rem = len & 31u;
      for (uint32_t r = 0; r < rows; ++r) encStep<PB>(state, ldEnt(tabAddr, ldsSym(cur + r * 32)), outPtr, ltMask);
      if (rem) {
        const bool valid = lane < rem;
        const uint32_t sy = valid ? ldsSym(cur + rows * 32) : 0u;
        encStepPartial<PB>(state, ldEnt(tabAddr, sy), outPtr, ltMask, valid);
      }
    }
  }
  if (runSyms & (kBlockSyms - 1)) flush(true); // the partial block
}

// ---------------------------------------------------------------------------------------------
// Compaction of one 8-block tile by ONE warp (the fused kernel's in-launch version of k_coalesce,
// ans_encode.cu): exclusive prefix of the tile from the per-group and per-tile sums before it, payload copy
// scratch slot -> archive, block index entries; tile 0 also writes the pdf table, the last tile the header,
// firstCompSegmentBytes and outSize (dietgpu/ans/GpuANSEncode.cuh:533-623). Every size of the element must be
// final and visible (the caller acquired the element's completion count); everything another SM produced is
// read through L2 (.cg). The payload goes through the warp's own shared memory (`stage`, kEncWarpSmem bytes,
// free once the warp has no run left): all 16-byte cp.async loads of as many blocks as fit are in flight at
// once, and since consecutive blocks are contiguous in the archive they leave as one contiguous copy.
// Deliberately not inlined: it runs once per 8 blocks and must not take registers from the row loop (the caller
// passes its __grid_constant__ parameter block).
// ---------------------------------------------------------------------------------------------
// Fused kernel: blocks until every block of tiles [0, t] of element e is encoded (t = tk - tileBase), i.e. until the
// sizes coalesceTileWarp sums up are final and the tile's own slots are complete. Whole groups before the tile's
// group are checked through the group counters, the tiles of its own group one per lane. Runs are handed out in
// order, so in steady state the first look succeeds.
// `wait` false: one look, returns whether the tile is ready (for callers that must not block).
static __device__ __noinline__ bool waitTilesEncoded(const EncParams& p, uint32_t e, uint32_t tk, uint32_t lane,
                                                     bool wait) {
  const Elem el = p.a.elems[e];
  const uint32_t t = tk - el.tileBase;
  const uint32_t nb = divUp(p.a.nDev ? p.a.nDev[e] : el.n, kBlockSyms); // device-side sizes: blocks beyond them never come
  if (t * kTileBlocks >= nb) {
    // No blocks in this tile (empty element or padding tile), so no finished block vouches for the element's split:
    // wait for its ready flag instead. The split role writes the float header (zeroing firstCompSegmentBytes) and
    // the compaction that follows this call fills that field in - an idle warp that took an empty element's tile at
    // the start of the kernel used to get there first, and the header write then wiped the size (seen as an
    // occasional wrong byte 16 of an EMPTY element's archive in batches small enough to leave warps without runs).
    if (p.ready) {
      uint32_t ok = 1;
      if (lane == 0) {
        uint32_t ns = 100;
        while ((ok = ldRelaxed(p.ready + e)) == 0 && wait) {
          __nanosleep(ns);
          if (ns < 1600) ns *= 2;
        }
      }
      ok = __shfl_sync(0xffffffffu, ok, 0);
      if (!ok) return false;
      asm volatile("fence.acq_rel.gpu;" ::: "memory");
    }
    return true;
  }
  constexpr uint32_t kGroupTiles = kGroupBlocks / kTileBlocks;
  const uint32_t grp = t / kGroupTiles;
  const uint32_t* groupSum = p.groupWords + el.groupBase;
  const uint32_t* tiles = p.tileWords + el.tileBase;
  uint32_t ns = 100;
  for (;;) {
    bool ok = true;
    for (uint32_t g = lane; g < grp; g += 32) ok = ok && (ldRelaxed(groupSum + g) >> kGroupCountShift) == kGroupBlocks;
    const uint32_t j = grp * kGroupTiles + lane;
    if (j <= t) ok = ok && (ldRelaxed(tiles + j) >> kTileCountShift) == min(kTileBlocks, nb - j * kTileBlocks);
    if (__all_sync(0xffffffffu, ok)) break;
    if (!wait) return false;
    __nanosleep(ns);
    if (ns < 1600) ns *= 2;
  }
  asm volatile("fence.acq_rel.gpu;" ::: "memory");
  return true;
}

static __device__ __noinline__ void coalesceTileWarp(const EncParams& p, int probBits, uint32_t e, uint32_t tk,
                                                     uint32_t lane, uint8_t* stage, uint32_t stageBytes) {
  const EncodeArgs& a = p.a;
  const Elem el = a.elems[e];
  const uint32_t t = tk - el.tileBase;
  const uint32_t n = a.nDev ? a.nDev[e] : el.n;
  const uint32_t nb = divUp(n, kBlockSyms);
  const uint32_t numTiles = max(1u, divUp(nb, kTileBlocks));
  if (t >= numTiles) return;
  if (lane == 0) DGB_TRACE_MIN(e, 3);
  uint8_t* ans = (uint8_t*)el.out;
  uint8_t* fbase = nullptr;
  if (a.floatType != kFtNone) {
    fbase = (uint8_t*)el.out + el.sizeBias;
    ans = fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
    if (a.round) ans += *(const uint32_t*)(fbase + 16);
  }
  constexpr uint32_t kGroupTiles = kGroupBlocks / kTileBlocks; // 32
  const uint32_t firstBlk = t * kTileBlocks;
  const uint32_t grp = t / kGroupTiles;
  const uint32_t* groupSum = p.groupWords + el.groupBase;
  const uint32_t* bwords = p.blockWords + el.blockBase;
  // lane j < 8 owns block firstBlk + j
  const uint32_t myBlk = firstBlk + lane;
  uint32_t words = 0, blkLen = 0;
  if (lane < kTileBlocks && myBlk < nb) {
    words = __ldcg(bwords + myBlk);
    blkLen = min(kBlockSyms, n - myBlk * kBlockSyms);
  }
  uint32_t part = 0;
  if (grp * kGroupTiles + lane < t) part = __ldcg(p.tileWords + tk - t + grp * kGroupTiles + lane) & kTileSizeMask;
  for (uint32_t g = lane; g < grp; g += 128) { // 4 loads in flight per lane (a 4 GiB element has 4096 groups)
    const uint32_t x0 = __ldcg(groupSum + g) & kGroupSizeMask;
    const uint32_t x1 = g + 32 < grp ? __ldcg(groupSum + g + 32) & kGroupSizeMask : 0u;
    const uint32_t x2 = g + 64 < grp ? __ldcg(groupSum + g + 64) & kGroupSizeMask : 0u;
    const uint32_t x3 = g + 96 < grp ? __ldcg(groupSum + g + 96) & kGroupSizeMask : 0u;
    part += x0 + x1 + x2 + x3;
  }
  const uint32_t excl = warpSum(part);
  const uint32_t padded = roundUp(words, 8);
  uint32_t incl = padded;
#pragma unroll
  for (int o = 1; o < (int)kTileBlocks; o <<= 1) {
    const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= (uint32_t)o) incl += v;
  }
  const uint32_t startWord = excl + incl - padded;
  const uint32_t tileWords = __shfl_sync(0xffffffffu, incl, kTileBlocks - 1);
  if (blkLen) ((uint2*)(ans + kAnsFixedBytes + 128u * nb))[myBlk] = make_uint2((blkLen << 16) | words, startWord);

  // ---- payload ----
  const uint32_t kCap = stageBytes / 16; // staging capacity in 16-byte vectors
  const uint32_t stageAddr = smemAddr(stage);
  uint4* dst = (uint4*)(ans + ansOverhead(nb) + 2 * (size_t)excl); // the tile's first vector in the archive
  const uint8_t* slots = p.scratch + (size_t)(el.blockBase + firstBlk) * kSlotBytes;
  uint32_t fill = 0;
  auto drain = [&]() {
    asm volatile("cp.async.wait_all;" ::: "memory");
    __syncwarp();
    for (uint32_t i = lane; i < fill; i += 32) dst[i] = ((const uint4*)stage)[i];
    __syncwarp(); // staging is rewritten next
    dst += fill;
    fill = 0;
  };
#pragma unroll 1
  for (uint32_t j = 0; j < kTileBlocks; ++j) {
    const uint32_t nv = __shfl_sync(0xffffffffu, padded, j) >> 3;
    if (nv == 0) continue;
    const uint4* src = (const uint4*)(slots + (size_t)j * kSlotBytes);
    if (fill + nv > kCap) drain();
    if (nv > kCap) { // a block larger than the staging area (incompressible data): straight through registers
      for (uint32_t i = lane; i < nv; i += 32) dst[i] = __ldcg(src + i);
      dst += nv;
      continue;
    }
    for (uint32_t i = lane; i < nv; i += 32)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageAddr + (fill + i) * 16), "l"(src + i) : "memory");
    fill += nv;
  }
  if (fill) drain();
  if (lane == 0) DGB_TRACE_MAX(e, 4);

  if (t == 0) { // pdf table; zero for absent symbols and for an empty element
    const uint2* tab = p.table + (size_t)e * 256;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint2 en = __ldcg(tab + lane + 32 * k);
      const uint32_t pdf = (n && en.x) ? (1u << probBits) - (en.y & 0xfffu) : 0u;
      ((uint16_t*)(ans + kAnsHeaderBytes))[lane + 32 * k] = (uint16_t)pdf;
    }
  }
  if (t == numTiles - 1 && lane == 0) {
    const uint32_t totalWords = excl + tileWords;
    const uint32_t cs = (a.useChecksum && a.checksum) ? __ldcg(a.checksum + e) : 0u;
    uint4* h = (uint4*)ans;
    h[0] = make_uint4(kAnsMagic, nb, n, totalWords);
    h[1] = make_uint4((uint32_t)probBits | (a.useChecksum ? 16u : 0u), cs, 0u, 0u);
    if (nb & 1u) ((uint2*)(ans + kAnsFixedBytes + 128u * nb))[nb] = make_uint2(0u, 0u);
    const uint32_t total = ansOverhead(nb) + 2u * totalWords;
    if (a.floatType != kFtNone && a.round == 0) *(uint32_t*)(fbase + 16) = total;
    if (a.outSize) {
      if (a.round) a.outSize[e] += total;
      else
        a.outSize[e] = el.sizeBias + total +
                       (a.floatType != kFtNone ? kFloatHeaderBytes + floatRawBytes(a.floatType, n) : 0u);
    }
  }
}

} // namespace dgb
