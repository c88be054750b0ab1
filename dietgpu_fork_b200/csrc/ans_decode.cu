// ANS decode side for sm_100a, with the float join fused behind it.
//
// What is computed follows the reference (citations relative to /root/reference/):
//   decode LUT           dietgpu/ans/GpuANSDecode.cuh:34-41, :405-476
//   per-warp rANS decode dietgpu/ans/GpuANSDecode.cuh:55-217
//   capacity / status    dietgpu/ans/GpuANSDecode.cuh:325-341
//   float join           dietgpu/float/GpuFloatDecompress.cuh:39-592, :709-805
// How: one prep kernel parses headers and expands the LUTs; one persistent kernel takes tiles of
// 8 blocks, stages each block's compressed words in shared memory (so the renormalisation read in
// the dependent chain is an LDS, not a global load), decodes into a shared-memory symbol tile and
// then writes the tile out with 16-byte vectors: plain bytes, or float words joined with the raw
// section. fp64 decodes both exponent planes of the tile before the join, so it is single pass
// too (the reference always takes two passes for fp64, GpuFloatDecompress.cuh:932-935).
#include <cstdlib>
#include <cstring>

#include "../../include/dietgpu_b200.h"

#include "float_math.cuh"
#include "internal.h"
#include "knobs.h"
#include "prof.h"

namespace dgb {

struct DecParams {
  DecodeArgs a;
  const DecInfo* info;  // [planes][numElems]
  const uint32_t* lut;  // [planes][numElems][1 << probBits]
  uint32_t claimTiles;  // consecutive 8-block tiles a CTA takes per claim
  uint32_t* tileCounter; // zeroed by k_decodePrep; claims beyond each CTA's first one
};

// ---------------------------------------------------------------------------------------------
// Prep: one CTA per (element, plane). Locates the ANS archive, checks capacity, reports
// success / size, and expands pdf[256] into the 2^probBits-entry LUT
//   entry = (slot - cdf[sym]) << 20 | pdf[sym] << 8 | sym.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_decodePrep(const DecodeArgs a, DecInfo* __restrict__ info,
                                                    uint32_t* __restrict__ lut, uint32_t* __restrict__ tileCounter) {
  __shared__ uint32_t sCdf[257];
  __shared__ uint32_t sWarp[8];
  const uint32_t e = blockIdx.x, plane = blockIdx.y, s = threadIdx.x, lane = s & 31, w = s >> 5;
  // programmatic dependent launch: k_decode (launched with the stream-serialization attribute) may become resident
  // and run its prologue now; it waits (griddepcontrol.wait) for this grid to complete before it reads anything
  // written here. Without the attribute on the next launch this is a no-op.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  if (e == 0 && plane == 0 && s == 0) *tileCounter = 0; // k_decode's dynamic tile claims start here
  const Elem el = a.elems[e];
  const uint8_t* arch = (const uint8_t*)el.in;
  uint32_t n, reportN;
  bool ok = true;
  const uint8_t* ans;
  if (a.floatType != kFtNone) {
    if (a.sparseNested) {
      uint32_t total = *(const uint32_t*)arch; // GpuSparseFloatHeader.size
      arch += 16 + sparseBitmapBytes(total);
    }
    const uint32_t* fh = (const uint32_t*)arch;
    ok = fh[0] == kFloatMagic && (int)(fh[2] & 0xfu) == a.floatType;
    n = fh[1];
    reportN = n;
    ans = arch + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
    if (plane) ans += fh[4]; // firstCompSegmentBytes
  } else {
    ans = arch;
    n = ((const uint32_t*)ans)[2];
    reportN = n;
  }
  const uint32_t* ah = (const uint32_t*)ans;
  if (a.floatType == kFtNone) ok = ah[0] == kAnsMagic;
  ok = ok && el.n >= n;
  if (ok && a.floatType != kFtNone) ok = ah[0] == kAnsMagic && ah[2] == n;
  const uint32_t nb = ok ? ah[1] : 0;
  const uint32_t pb = ok ? (ah[4] & 0xfu) : (uint32_t)a.probBits;
  if (ok && (int)pb != a.probBits) ok = false;
  // the block count is implied by the symbol count: a header that disagrees would make k_decode index past the
  // caller's capacity (the reference only asserts, dietgpu/ans/GpuANSDecode.cuh:325-341)
  if (ok && nb != divUp(n, kBlockSyms)) ok = false;
  if (s == 0) {
    DecInfo d;
    d.ans = ans;
    d.fbase = a.floatType != kFtNone ? arch : nullptr;
    d.nb = ok ? nb : 0;
    d.n = ok ? n : 0;
    d.ok = ok;
    d.totalWords = ok ? ah[3] : 0;
    info[plane * a.numElems + e] = d;
    if (plane == 0) {
      if (a.outSuccess) a.outSuccess[e] = ok ? 1 : 0;
      if (a.outSize) a.outSize[e] = reportN;
    }
  }
  if (!ok || n == 0) return;

  const uint32_t pdf = ((const uint16_t*)(ans + kAnsHeaderBytes))[s];
  uint32_t inc = pdf;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (uint32_t)o) inc += t;
  }
  if (lane == 31) sWarp[w] = inc;
  __syncthreads();
  uint32_t cdf = inc - pdf;
  for (uint32_t i = 0; i < w; ++i) cdf += sWarp[i];
  sCdf[s] = cdf;
  if (s == 255) sCdf[256] = cdf + pdf;
  __syncthreads();
  // every thread fills a strided share of the slots; symbol found by binary search on the cdf
  const uint32_t slots = 1u << a.probBits;
  uint32_t* out = lut + ((size_t)plane * a.numElems + e) * slots;
  for (uint32_t j = s; j < slots; j += 256) {
    uint32_t lo = 0, hi = 256; // cdf[lo] <= j < cdf[hi]
    while (hi - lo > 1) {
      uint32_t mid = (lo + hi) >> 1;
      if (sCdf[mid] <= j) lo = mid; else hi = mid;
    }
    // skip zero-width symbols that share the same cdf (binary search lands on the last of them,
    // which is the one with non-zero width when j < cdf[hi])
    uint32_t p = sCdf[lo + 1] - sCdf[lo];
    out[j] = ((j - sCdf[lo]) << 20) | (p << 8) | lo;
  }
}

// ---------------------------------------------------------------------------------------------
// Block decoder: 8 warps per CTA share the element's LUT(s) in shared memory; apart from the
// barrier pair at an element change the warps are autonomous, one ANS block each.
//
//   compressed words: the decoder walks a block's words from the end. They are pulled into a
//       2 KB per-warp shared-memory ring in 1 KB pieces by cp.async.bulk (TMA 1-D) + mbarrier,
//       always keeping the piece under the read pointer and the one below it resident, so the
//       renormalisation read inside the dependent chain is an LDS.
//   symbols: 16 rows (512 symbols) at a time into a double-buffered 512-byte transposer, then
//       written out by the same warp with 16-byte vectors: plain bytes, or float words joined
//       with the raw section (both planes first for fp64).
// The small footprint (~3 KB per warp) is what buys occupancy: the per-row dependent chain holds
// two LDS and a POPC, so throughput comes from many resident warps.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kDecWarps = 8;
constexpr uint32_t kPieceWords = 512;                 // 1 KB
constexpr uint32_t kDecRingBytes = 2 * kPieceWords * 2; // 2 pieces
constexpr uint32_t kDecChunk = 512;                   // symbols per transposer buffer (16 rows)

// Shared memory of a CTA, from a base aligned to the LUT size (so that "(x & mask) | base" forms both the
// LUT and the ring addresses in one LOP3):
//   LUT[planes] (4 << PB each) | ring[warps][planes] (2 KB each) | transposer[warps][planes] (512 B) |
//   raw stage[warps] | mbarriers[warps][planes] (2 x 8 B)
// raw-section bytes of one 512-word chunk staged per warp by cp.async (single-plane float types)
template <int FT> __host__ __device__ constexpr uint32_t decRawStage() {
  return (FT == kFtF16 || FT == kFtBF16) ? kDecChunk : (FT == kFtF32 ? 3 * kDecChunk : 0);
}
template <int FT> __host__ __device__ constexpr uint32_t decWarpSmem() {
  return (FT == kFtF64 ? 2 : 1) * (kDecRingBytes + kDecChunk + 16) + decRawStage<FT>();
}
template <int PB, int FT> __host__ __device__ constexpr uint32_t decCtaSmem() {
  return (4u << PB) /* alignment slack */ + (FT == kFtF64 ? 2 : 1) * (4u << PB) + kDecWarps * decWarpSmem<FT>();
}

// per-plane decode context of one warp
struct DecCtx {
  uint32_t state;
  uint32_t ptrBytes;   // 2 * (index one past the next word to read)
  int cur;             // 1 KB piece holding the next word to read (-1: the block has no words)
  uint32_t pend;       // bit s: ring slot s has a copy in flight that has not been waited for
  const uint8_t* src;  // this block's first payload word
  uint32_t totalBytes; // roundUp(words, 8) * 2
  uint32_t ringAddr, bar0, lutAddr;
};

// Ring protocol: pieces `cur` and `cur - 1` are resident or in flight (slot = piece & 1). A 16-row chunk reads
// at most 1 KB, so it crosses at most one piece boundary; when it has, the slot of the piece left behind takes
// piece `cur - 1` (new cur), which is waited for at the start of the next chunk.
__device__ __forceinline__ void decRequest(DecCtx& c, int piece, uint32_t lane) {
  const uint32_t slot = (uint32_t)piece & 1u;
  if (lane == 0) {
    const uint32_t off = (uint32_t)piece * kPieceWords * 2;
    const uint32_t bytes = min(kPieceWords * 2, c.totalBytes - off);
    mbarExpectTx(c.bar0 + 8 * slot, bytes);
    bulkLoad(c.ringAddr + slot * kPieceWords * 2, c.src + off, bytes, c.bar0 + 8 * slot);
  }
  c.pend |= 1u << slot;
}

// One decode step of a full symbol row (dietgpu/ans/GpuANSDecode.cuh:55-90, :174-217): LUT lookup, state
// update, ballot of the lanes that must renormalise, their 16-bit reads from the end of the block's words in
// descending lane order. lutAddr is aligned to the LUT size and ringAddr to the ring size. Returns the LUT
// entry (symbol in the low byte). PTX to pin the instruction selection (18 instructions).
template <int PB>
__device__ __forceinline__ uint32_t decStep(uint32_t& state, uint32_t lutAddr, uint32_t ringAddr, uint32_t& ptrBytes,
                                            uint32_t geMask) {
  uint32_t e;
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      ".reg .b32 a, hi, t, b, ns, v, pre, ra, w, c;\n"
      "shl.b32 a, %0, 2;\n"
      "lop3.b32 a, a, %6, %3, 0xEA;\n"        // ((state << 2) & mask) | lutAddr
      "ld.shared.u32 %2, [a];\n"
      "shr.u32 hi, %0, %7;\n"
      "mul.lo.u32 t, %2, 4096;\n"             // pdf = (e << 12) >> 20: the shift left runs on the FMA pipe
      "shr.u32 t, t, 20;\n"                   //   (IMAD.SHL), one ALU instruction less than shr + and
      "shr.u32 b, %2, 20;\n"                  // slot - cdf
      "mad.lo.u32 ns, t, hi, b;\n"
      "setp.lt.u32 p, ns, 32768;\n"
      "vote.sync.ballot.b32 v, p, 0xffffffff;\n"
      "and.b32 pre, v, %5;\n"
      "popc.b32 pre, pre;\n"
      "mad.lo.u32 ra, pre, 0xfffffffe, %1;\n" // ptr - 2 * (readers at or above this lane)
      "lop3.b32 ra, ra, %8, %4, 0xEA;\n"      // (ra & ringMask) | ringAddr
      "ld.shared.u16 w, [ra];\n"              // every lane loads (a valid ring slot); only readers use it
      "@p mad.lo.u32 ns, ns, 65536, w;\n"     // (ns << 16) | w on the FMA pipe (w < 65536, ns < 32768)
      "popc.b32 c, v;\n"
      "mad.lo.u32 %1, c, 0xfffffffe, %1;\n"
      "mov.b32 %0, ns;\n"
      "}\n"
      : "+r"(state), "+r"(ptrBytes), "=&r"(e)
      : "r"(lutAddr), "r"(ringAddr), "r"(geMask), "n"(((1u << PB) - 1u) << 2), "n"(PB), "n"(kDecRingBytes - 1)
      : "memory");
  return e;
}

// partial row: lanes >= rem keep their state and read nothing
template <int PB>
__device__ __forceinline__ uint32_t decStepPartial(uint32_t& state, uint32_t lutAddr, uint32_t ringAddr,
                                                   uint32_t& ptrBytes, uint32_t geMask, bool valid) {
  uint32_t e;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(lutAddr + ((state & ((1u << PB) - 1u)) << 2)));
  uint32_t ns = ((e >> 8) & 0xfffu) * (state >> PB) + (e >> 20);
  ns = valid ? ns : state;
  const bool rd = valid && ns < kStateMin;
  const uint32_t vote = __ballot_sync(0xffffffffu, rd);
  if (rd) {
    uint32_t w;
    const uint32_t a = ringAddr + ((ptrBytes - 2u * __popc(vote & geMask)) & (kDecRingBytes - 1));
    asm volatile("ld.shared.u16 %0, [%1];" : "=r"(w) : "r"(a));
    ns = (ns << 16) | w;
  }
  state = ns;
  ptrBytes -= 2u * __popc(vote);
  return e;
}

// Pulls the raw-section bytes of one block (the ones writeChunk will join with the block's symbols) into L2
// when the block starts: one 128-byte line per lane and request, so the joins later wait an L2 hit instead of
// a DRAM round trip.
template <int FT>
__device__ __forceinline__ void prefetchRawBlock(const DecInfo& d, uint32_t blk, uint32_t blkLen, uint32_t lane) {
  if constexpr (FT != kFtNone) {
    const RawSections rs = rawSections(FT, (uint8_t*)d.fbase + kFloatHeaderBytes, d.n);
    constexpr uint32_t aBytes = FT == kFtF64 ? 4 : (FT == kFtF32 ? 2 : 1); // per word in section a
    constexpr uint32_t bBytes = FT == kFtF64 ? 2 : (FT == kFtF32 ? 1 : 0);
    const uint8_t* pa = (const uint8_t*)rs.a + (size_t)blk * kBlockSyms * aBytes;
    for (uint32_t off = lane * 128; off < blkLen * aBytes; off += 32 * 128)
      asm volatile("prefetch.global.L2 [%0];" ::"l"(pa + off));
    if constexpr (bBytes != 0) {
      const uint8_t* pb = (const uint8_t*)rs.b + (size_t)blk * kBlockSyms * bBytes;
      for (uint32_t off = lane * 128; off < blkLen * bBytes; off += 32 * 128)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(pb + off));
    }
  }
}

// Starts the asynchronous copy of the chunk's raw bytes into the warp's stage (16 bytes per lane and request)
// and returns raw-section pointers biased so that the global indices used by the join land in the stage.
template <int FT>
__device__ __forceinline__ RawSections stageRaw(const RawSections& g, uint32_t wordOff, uint32_t len, uint8_t* stage,
                                                uint32_t lane) {
  constexpr uint32_t aBytes = FT == kFtF32 ? 2 : 1;
  const uint32_t stageAddr = smemAddr(stage);
  const uint8_t* srcA = g.a + (size_t)wordOff * aBytes;
  for (uint32_t off = lane * 16; off < len * aBytes; off += 512)
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageAddr + off), "l"(srcA + off) : "memory");
  RawSections r;
  r.a = stage - (size_t)wordOff * aBytes;
  r.b = r.a;
  if constexpr (FT == kFtF32) {
    const uint8_t* srcB = g.b + wordOff;
    for (uint32_t off = lane * 16; off < len; off += 512)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageAddr + 2 * kDecChunk + off), "l"(srcB + off)
                   : "memory");
    r.b = stage + 2 * kDecChunk - wordOff;
  }
  asm volatile("cp.async.commit_group;" ::: "memory");
  return r;
}

template <int FT>
__device__ __forceinline__ void writeChunk(const RawSections& rs, uint8_t* outBase, uint32_t wordOff, uint32_t len,
                                           const uint8_t* sym0, const uint8_t* sym1, uint32_t lane) {
  if constexpr (FT == kFtNone) {
    uint8_t* out = outBase + wordOff;
    if ((((uintptr_t)out) & 15u) == 0 && len == kDecChunk) {
      ((uint4*)out)[lane] = ((const uint4*)sym0)[lane];
    } else {
      for (uint32_t i = lane; i < len; i += 32) out[i] = sym0[i];
    }
  } else {
    constexpr uint32_t G = FloatTraits<FT>::kGroup;
    uint32_t vecWords = 0;
    if ((((uintptr_t)outBase) & 15u) == 0) {
      const uint32_t ng = len / G, g0 = wordOff / G;
      uint4* o = (uint4*)outBase;
#pragma unroll 2
      for (uint32_t i = lane; i < ng; i += 32) o[g0 + i] = joinGroup<FT>(g0 + i, sym0 + i * G, sym1 + i * G, rs);
      vecWords = ng * G;
    }
    for (uint32_t i = vecWords + lane; i < len; i += 32)
      joinScalar<FT>(outBase, wordOff + i, sym0[i], sym1[i], rs);
  }
}

// per-halfword rotate-right-by-one of two packed 16-bit words in 3 instructions: shift, shift (IMAD.SHL, FMA pipe),
// bit select (one LOP3)
__device__ __forceinline__ uint32_t rotr1x2sel(uint32_t y) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, 0x7fff7fff, 0xE4;" : "=r"(r) : "r"(y >> 1), "r"(y << 15)); // (a & c) | (b & ~c), c = the mask
  return r;
}

// Full 512-word chunk, 16-byte aligned output, raw bytes staged in shared memory: fixed trip counts and
// shared-memory addresses only (the general writeChunk handles everything else).
template <int FT>
__device__ __forceinline__ void writeChunkFull(uint32_t symAddr, uint32_t rawAddr, uint4* out, uint32_t lane) {
  if constexpr (FT == kFtF16 || FT == kFtBF16) {
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const uint32_t i = lane + 32 * k;
      uint2 e, r;
      asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(symAddr + i * 8));
      asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "r"(rawAddr + i * 8));
      uint32_t y0 = __byte_perm(r.x, e.x, 0x5140), y1 = __byte_perm(r.x, e.x, 0x7362);
      uint32_t y2 = __byte_perm(r.y, e.y, 0x5140), y3 = __byte_perm(r.y, e.y, 0x7362);
      if constexpr (FT == kFtBF16) {
        y0 = rotr1x2sel(y0); y1 = rotr1x2sel(y1); y2 = rotr1x2sel(y2); y3 = rotr1x2sel(y3);
      }
      out[i] = make_uint4(y0, y1, y2, y3);
    }
  } else { // kFtF32: 128 groups of 4 words; lo = 2 bytes per word at rawAddr, hi = 1 byte per word after 1 KB
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const uint32_t i = lane + 32 * k;
      uint32_t e, hi;
      uint2 lo;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(e) : "r"(symAddr + i * 4));
      asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(lo.x), "=r"(lo.y) : "r"(rawAddr + i * 8));
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(hi) : "r"(rawAddr + 2 * kDecChunk + i * 4));
      const uint32_t t01 = __byte_perm(hi, e, 0x5140), t23 = __byte_perm(hi, e, 0x7362);
      const uint32_t v0 = __byte_perm(lo.x, t01, 0x5410), v1 = __byte_perm(lo.x, t01, 0x7632);
      const uint32_t v2 = __byte_perm(lo.y, t23, 0x5410), v3 = __byte_perm(lo.y, t23, 0x7632);
      out[i] = make_uint4(__funnelshift_r(v0, v0, 1), __funnelshift_r(v1, v1, 1), __funnelshift_r(v2, v2, 1),
                          __funnelshift_r(v3, v3, 1));
    }
  }
}


// ---------------------------------------------------------------------------------------------
// The common block: 4096 symbols, one plane, 16-byte aligned archive and output. Same ring / transposer / join
// as the general chunk loop in k_decode, minus everything a partial or misaligned block needs: fixed trip
// counts, global pointers that step down one chunk per iteration, no generic-address arithmetic, no partial
// row. (ncu, round 1: 145 of 449 warp instructions per 16-row chunk were bookkeeping around the 16 row steps.)
// ---------------------------------------------------------------------------------------------
template <int PB, int FT>
__device__ __forceinline__ void decodeBlockFast(DecCtx& cx, uint32_t& phaseBits, uint32_t geMask, uint32_t lane,
                                                uint32_t transAddr, uint32_t rawStageAddr, const RawSections& rs,
                                                uint32_t blk, uint8_t* outBase) {
  constexpr uint32_t kChunks = kBlockSyms / kDecChunk; // 8
  uint32_t state = cx.state, ptrBytes = cx.ptrBytes;
  const uint32_t symAddr = transAddr + lane;
  const uint32_t lastWord = blk * kBlockSyms + (kChunks - 1) * kDecChunk; // first word of the block's last chunk
  // this lane's global addresses in the chunk being decoded (the block is walked from its end)
  const uint8_t* rawA = nullptr;
  const uint8_t* rawB = nullptr;
  uint4* out;
  if constexpr (FT == kFtNone) {
    out = (uint4*)outBase + (kChunks - 1) * (kDecChunk / 16); // outBase is the block's first byte
  } else if constexpr (FT == kFtF32) {
    rawA = rs.a + (size_t)lastWord * 2 + lane * 16;
    rawB = rs.b + (size_t)lastWord + lane * 16;
    out = (uint4*)outBase + lastWord / 4;
  } else {
    rawA = rs.a + (size_t)lastWord + lane * 16;
    out = (uint4*)outBase + lastWord / 8;
  }
  const uint32_t stageLane = rawStageAddr + lane * 16;
#pragma unroll 1
  for (uint32_t c = 0; c < kChunks; ++c) {
    if constexpr (FT == kFtF16 || FT == kFtBF16) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageLane), "l"(rawA) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
      rawA -= kDecChunk;
    } else if constexpr (FT == kFtF32) {
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageLane), "l"(rawA) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageLane + 512), "l"(rawA + 512) : "memory");
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(stageLane + 2 * kDecChunk), "l"(rawB) : "memory");
      asm volatile("cp.async.commit_group;" ::: "memory");
      rawA -= 2 * kDecChunk;
      rawB -= kDecChunk;
    }
    if (cx.pend) { // copies requested since the last chunk must have landed
#pragma unroll
      for (uint32_t slot = 0; slot < 2; ++slot) {
        if (cx.pend & (1u << slot)) {
          mbarWait(cx.bar0 + 8 * slot, (phaseBits >> slot) & 1u);
          phaseBits ^= 1u << slot;
        }
      }
      cx.pend = 0;
    }
#pragma unroll
    for (int r = 15; r >= 0; --r) {
      const uint32_t ev = decStep<PB>(state, cx.lutAddr, cx.ringAddr, ptrBytes, geMask);
      asm volatile("st.shared.u8 [%0], %1;" ::"r"(symAddr + r * 32), "r"(ev) : "memory");
    }
    // crossed into the piece below: the slot left behind takes the next lower piece
    if (cx.cur >= 1 && ptrBytes <= (uint32_t)cx.cur * (kPieceWords * 2)) {
      cx.cur -= 1;
      if (cx.cur >= 1) {
        __syncwarp(); // all lanes are past their reads of the slot being refilled
        decRequest(cx, cx.cur - 1, lane);
      }
    }
    if constexpr (FT != kFtNone) asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();
    if constexpr (FT == kFtNone) {
      uint4 v;
      asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(transAddr + lane * 16));
      out[lane] = v;
      out -= kDecChunk / 16;
    } else {
      writeChunkFull<FT>(transAddr, rawStageAddr, out, lane);
      out -= kDecChunk / FloatTraits<FT == kFtNone ? kFtF16 : FT>::kGroup;
    }
    __syncwarp(); // transposer and raw stage are reused by the next chunk
  }
  cx.state = state;
  cx.ptrBytes = ptrBytes;
}

template <int PB, int FT>
__global__ void __launch_bounds__(kDecWarps * 32, FT == kFtF64 ? 4 : 6) k_decode(const DecParams p) {
  constexpr int kPlanes = FT == kFtF64 ? 2 : 1;
  constexpr uint32_t kLut = 1u << PB;
  extern __shared__ __align__(128) uint8_t smemRaw[];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  // base aligned to the LUT size (see decCtaSmem)
  const uint32_t rawAddr = smemAddr(smemRaw);
  uint8_t* smem = smemRaw + (((rawAddr + kLut * 4 - 1) & ~(kLut * 4 - 1)) - rawAddr);
  uint32_t* sLut = (uint32_t*)smem; // kPlanes * kLut
  uint8_t* rings = smem + kPlanes * kLut * 4;
  uint8_t* transBase = rings + kDecWarps * kPlanes * kDecRingBytes;
  uint8_t* rawStage = transBase + kDecWarps * kPlanes * kDecChunk + w * decRawStage<FT>();
  uint8_t* bars = transBase + kDecWarps * kPlanes * kDecChunk + kDecWarps * decRawStage<FT>();
  uint8_t* trans[kPlanes];
  DecCtx ctx[kPlanes];
#pragma unroll
  for (int pl = 0; pl < kPlanes; ++pl) {
    const uint32_t slot = w * kPlanes + pl;
    ctx[pl].ringAddr = smemAddr(rings + slot * kDecRingBytes);
    trans[pl] = transBase + slot * kDecChunk;
    ctx[pl].bar0 = smemAddr(bars + slot * 16);
    ctx[pl].lutAddr = smemAddr(sLut + pl * kLut);
    if (lane == 0) {
      mbarInit(ctx[pl].bar0, 1);
      mbarInit(ctx[pl].bar0 + 8, 1);
    }
  }
  if (lane == 0) asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  __syncwarp();
  uint32_t phaseBits = 0; // bit (2 * plane + slot)
  const uint32_t geMask = laneMaskGe();
  const DecodeArgs& a = p.a;

  // everything above touched only this CTA's shared memory; what k_decodePrep wrote (info, LUTs, the claim counter)
  // is read from here on (no-op unless this grid was launched as a programmatic dependent)
  asm volatile("griddepcontrol.wait;" ::: "memory");
  // Tiles are claimed in order, `claimTiles` at a time (first claim static), so the CTAs finish together.
  __shared__ uint32_t sNext;
  uint32_t cur = 0xffffffffu;
  uint32_t claim = blockIdx.x;
  for (;;) {
    const uint32_t t0 = claim * p.claimTiles;
    if (t0 >= a.totalTiles) break;
    const uint32_t t1 = min(t0 + p.claimTiles, a.totalTiles);
    if (tid == 0) sNext = gridDim.x + atomicAdd(p.tileCounter, 1u);
  for (uint32_t tile = t0; tile < t1; ++tile) {
    const uint32_t e = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tile, cur);
    if (e != cur) { // CTA-uniform: every warp of the CTA walks the same tile sequence
      __syncthreads();
      if (p.info[e].ok && p.info[e].n) {
#pragma unroll
        for (int pl = 0; pl < kPlanes; ++pl) {
          const uint4* src = (const uint4*)(p.lut + ((size_t)pl * a.numElems + e) * kLut);
          for (uint32_t i = tid; i < kLut / 4; i += kDecWarps * 32) ((uint4*)(sLut + pl * kLut))[i] = src[i];
        }
      }
      __syncthreads();
      cur = e;
    }
    const DecInfo d0 = p.info[e];
    const Elem el = a.elems[e];
    const uint32_t blk = (tile - el.tileBase) * kTileBlocks + w;
    if (!d0.ok || blk >= d0.nb) continue;
    if constexpr (kPlanes == 2) {
      if (!p.info[a.numElems + e].ok) { // the second exponent plane's header is bad: nothing of this element decodes
        if (lane == 0 && blk == 0 && a.outSuccess) a.outSuccess[e] = 0;
        continue;
      }
    }
    const uint32_t nb = d0.nb;

    // Block index entries come from the archive: an entry that disagrees with what the header implies (symbols of
    // the block, 16-byte aligned start inside the payload) would turn into a misaligned or out-of-bounds bulk
    // copy; such a block is skipped and the element reported as failed.
    const uint32_t blkLen = min(kBlockSyms, d0.n - blk * kBlockSyms);
    bool sane = true;
#pragma unroll
    for (int pl = 0; pl < kPlanes; ++pl) {
      const DecInfo dp = pl == 0 ? d0 : p.info[a.numElems + e];
      const uint2 bw = ((const uint2*)(dp.ans + kAnsFixedBytes + 128u * nb))[blk];
      const uint32_t words = bw.x & 0xffffu;
      sane = sane && (bw.x >> 16) == blkLen && words <= kStageWords && (bw.y & 7u) == 0 &&
             (uint64_t)bw.y + roundUp(words, 8) <= dp.totalWords;
    }
    if (!sane) {
      if (lane == 0 && a.outSuccess) a.outSuccess[e] = 0;
      continue;
    }
#pragma unroll
    for (int pl = 0; pl < kPlanes; ++pl) {
      const uint8_t* ans = pl == 0 ? d0.ans : p.info[a.numElems + e].ans;
      DecCtx& c = ctx[pl];
      c.state = ((const uint32_t*)(ans + kAnsFixedBytes))[blk * 32 + lane];
      const uint2 bw = ((const uint2*)(ans + kAnsFixedBytes + 128u * nb))[blk];
      const uint32_t words = bw.x & 0xffffu;
      c.src = ans + ansOverhead(nb) + 2 * (size_t)bw.y;
      c.totalBytes = roundUp(words, 8) * 2;
      c.ptrBytes = words * 2;
      c.cur = words ? (int)((words - 1) / kPieceWords) : -1;
      c.pend = 0;
      __syncwarp(); // every lane is done with the ring of the previous block
      if (c.cur >= 0) decRequest(c, c.cur, lane);
      if (c.cur >= 1) decRequest(c, c.cur - 1, lane);
    }
    uint8_t* outBase;
    if constexpr (FT == kFtNone) outBase = (uint8_t*)el.out + (size_t)blk * kBlockSyms;
    else
      outBase = a.denseBase ? a.denseBase + el.planeOff * sizeof(typename FloatTraits<FT == kFtNone ? kFtF16 : FT>::Word)
                            : (uint8_t*)el.out;

    RawSections rsGlobal;
    rsGlobal.a = rsGlobal.b = nullptr;
    bool rawAligned = false;
    const bool outAligned = (((uintptr_t)outBase) & 15u) == 0;
    if constexpr (FT != kFtNone) {
      rsGlobal = rawSections(FT, (uint8_t*)d0.fbase + kFloatHeaderBytes, d0.n);
      rawAligned = (((uintptr_t)d0.fbase) & 15u) == 0;
      if (!rawAligned || decRawStage<FT>() == 0) prefetchRawBlock<FT>(d0, blk, blkLen, lane);
    }
    if constexpr (kPlanes == 1) {
      if (blkLen == kBlockSyms && outAligned && (FT == kFtNone || rawAligned)) {
        decodeBlockFast<PB, FT>(ctx[0], phaseBits, geMask, lane, smemAddr(trans[0]), smemAddr(rawStage), rsGlobal, blk,
                                outBase);
        continue;
      }
    }
    const uint32_t numChunks = divUp(blkLen, kDecChunk);
    for (int c = (int)numChunks - 1; c >= 0; --c) {
      const uint32_t len = min(kDecChunk, blkLen - (uint32_t)c * kDecChunk);
      const uint32_t rows = len >> 5, rem = len & 31u;
      const uint32_t wordOff = FT == kFtNone ? (uint32_t)c * kDecChunk : blk * kBlockSyms + (uint32_t)c * kDecChunk;
      // raw bytes of this chunk: asynchronous copy into the warp's stage while the chunk is decoded
      RawSections rs = rsGlobal;
      if constexpr (decRawStage<FT>() != 0) {
        if (rawAligned) rs = stageRaw<FT>(rsGlobal, wordOff, len, rawStage, lane);
      }
#pragma unroll
      for (int pl = 0; pl < kPlanes; ++pl) {
        DecCtx& cx = ctx[pl];
        if (cx.pend) { // copies requested since the last chunk must have landed
#pragma unroll
          for (uint32_t slot = 0; slot < 2; ++slot) {
            if (cx.pend & (1u << slot)) {
              mbarWait(cx.bar0 + 8 * slot, (phaseBits >> (2 * pl + slot)) & 1u);
              phaseBits ^= 1u << (2 * pl + slot);
            }
          }
          cx.pend = 0;
        }
        uint32_t state = cx.state, ptrBytes = cx.ptrBytes;
        const uint32_t symAddr = smemAddr(trans[pl]) + lane;
        if (rem) {
          const bool valid = lane < rem;
          const uint32_t ev = decStepPartial<PB>(state, cx.lutAddr, cx.ringAddr, ptrBytes, geMask, valid);
          if (valid) asm volatile("st.shared.u8 [%0], %1;" ::"r"(symAddr + rows * 32), "r"(ev) : "memory");
        }
        if (rows == 16) {
#pragma unroll
          for (int r = 15; r >= 0; --r) {
            const uint32_t ev = decStep<PB>(state, cx.lutAddr, cx.ringAddr, ptrBytes, geMask);
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(symAddr + r * 32), "r"(ev) : "memory");
          }
        } else {
          for (int r = (int)rows - 1; r >= 0; --r) {
            const uint32_t ev = decStep<PB>(state, cx.lutAddr, cx.ringAddr, ptrBytes, geMask);
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(symAddr + r * 32), "r"(ev) : "memory");
          }
        }
        cx.state = state;
        cx.ptrBytes = ptrBytes;
        // crossed into the piece below: the slot left behind takes the next lower piece
        if (cx.cur >= 1 && ptrBytes <= (uint32_t)cx.cur * (kPieceWords * 2)) {
          cx.cur -= 1;
          if (cx.cur >= 1) {
            __syncwarp(); // all lanes are past their reads of the slot being refilled
            decRequest(cx, cx.cur - 1, lane);
          }
        }
      }
      if constexpr (decRawStage<FT>() != 0) asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncwarp();
      bool done = false;
      if constexpr (decRawStage<FT>() != 0) {
        if (len == kDecChunk && rawAligned && outAligned) {
          writeChunkFull<FT>(smemAddr(trans[0]), smemAddr(rawStage),
                             (uint4*)outBase + wordOff / FloatTraits<FT == kFtNone ? kFtF16 : FT>::kGroup, lane);
          done = true;
        }
      }
      if (!done) writeChunk<FT>(rs, outBase, wordOff, len, trans[0], trans[kPlanes - 1], lane);
      __syncwarp(); // transposer and raw stage are reused by the next chunk
    }
  }
    __syncthreads();
    claim = sNext;
    __syncthreads();
  }
}

template <int PB, int FT>
static int launchDecodeKernel(DecParams& p, cudaStream_t s) {
  constexpr uint32_t smemBytes = decCtaSmem<PB, FT>();
  {
    static DeviceOnce once;
    const int dev = currentDevice();
    if (once.first(dev)) {
      DGB_CUDA(cudaFuncSetAttribute(k_decode<PB, FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
      once.done(dev);
    }
  }
  const uint32_t perSm = FT == kFtF64 ? 4u : 6u;
  uint32_t ctas = min(p.a.totalTiles, (uint32_t)smCount() * perSm);
  p.claimTiles = 1; // measured on B200 (bf16, 128 x 4 MiB): 1: 0.300, 2: 0.307, 3: 0.320, 4: 0.327 ms; static ranges: 0.359 ms
  if (DGB_KNOB("DGB_DEC_CLAIM", 0) > 0) p.claimTiles = (uint32_t)DGB_KNOB("DGB_DEC_CLAIM", 0); // experiment knob
  if (p.a.totalTiles < ctas * p.claimTiles) p.claimTiles = 1;
  ctas = min(ctas, divUp(p.a.totalTiles, p.claimTiles));
  {
    ProfScope ps__("k_decode", s);
    // Programmatic dependent launch: this grid becomes resident during k_decodePrep and runs its prologue; it waits
    // (griddepcontrol.wait) before it reads what the prep kernel wrote. Measured on B200 (same box, DGB_DEC_PDL 0 / 1):
    // bf16 128 x 4 MiB decompress 0.2719 -> 0.2697 ms, 16 x 8 MiB 0.0843 -> 0.0808; but at one wave of tiles it costs:
    // 1 x 64 MiB 0.058 -> 0.068 ms, 32 x 2 MiB 0.061 -> 0.072. So: from two waves up.
    if (DGB_KNOB("DGB_DEC_PDL", 1) != 0 && p.a.totalTiles >= 2048) {
      cudaLaunchConfig_t cfg = {};
      cfg.gridDim = dim3(ctas);
      cfg.blockDim = dim3(kDecWarps * 32);
      cfg.dynamicSmemBytes = smemBytes;
      cfg.stream = s;
      cudaLaunchAttribute at[1];
      at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
      at[0].val.programmaticStreamSerializationAllowed = 1;
      cfg.attrs = at;
      cfg.numAttrs = 1;
      DGB_CUDA(cudaLaunchKernelEx(&cfg, k_decode<PB, FT>, p));
    } else {
      k_decode<PB, FT><<<ctas, kDecWarps * 32, smemBytes, s>>>(p);
    }
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

template <int PB>
static int launchDecodeFt(DecParams& p, cudaStream_t s) {
  switch (p.a.floatType) {
    case kFtNone: return launchDecodeKernel<PB, kFtNone>(p, s);
    case kFtF16: return launchDecodeKernel<PB, kFtF16>(p, s);
    case kFtBF16: return launchDecodeKernel<PB, kFtBF16>(p, s);
    case kFtF32: return launchDecodeKernel<PB, kFtF32>(p, s);
    case kFtF64: return launchDecodeKernel<PB, kFtF64>(p, s);
    default: return kErrInvalidArg;
  }
}

int launchDecode(Temp& tmp, const DecodeArgs& a, cudaStream_t s) {
  if (a.numElems == 0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  const int planes = a.floatType == kFtF64 ? 2 : 1;
  DecInfo* info = tmp.alloc<DecInfo>((size_t)planes * a.numElems);
  uint32_t* lut = tmp.alloc<uint32_t>(((size_t)planes * a.numElems) << a.probBits);
  uint32_t* counter = tmp.alloc<uint32_t>(4);
  if (tmp.failed()) return kErrCuda;
  {
    ProfScope ps__("k_decodePrep", s);
    k_decodePrep<<<dim3(a.numElems, planes), 256, 0, s>>>(a, info, lut, counter);
  }
  DGB_CUDA(cudaGetLastError());
  DecParams p;
  p.a = a;
  p.info = info;
  p.lut = lut;
  p.claimTiles = 1;
  p.tileCounter = counter;
  if (a.totalTiles == 0) return kOk;
  switch (a.probBits) {
    case 9: return launchDecodeFt<9>(p, s);
    case 10: return launchDecodeFt<10>(p, s);
    default: return launchDecodeFt<11>(p, s);
  }
}


// ---------------------------------------------------------------------------------------------
// No-allocation decode (dgb_decode_batch_noalloc, include/dietgpu_b200.h). No counterpart in the reference: added
// for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
// serving loops, harnesses that forbid it). It is launchDecode above with the decoder's scratch - descriptor table,
// DecInfo records, expanded LUTs and the tile-claim counter - placed in ONE file-scope device array of fixed size
// instead of the caller's temp region / the stream-ordered pool, and with the descriptor table shipped as a kernel
// argument instead of a host-to-device copy. Nothing on this path calls cudaMalloc, cudaFree, cudaMallocAsync,
// cudaMallocHost or an allocating library routine.
//
// Limit: kNoAllocMaxBatch batch elements per call (the scratch does not depend on the elements' sizes).
// The array is resident on every GPU that loads the library until the process exits (about 0.5 MiB).
// Every call on a device shares the array and it keeps the previous call's contents: each call rewrites every record
// it later reads (descriptors by k_noAllocLoadElems, DecInfo + LUT + the claim counter by k_decodePrep), so nothing
// needs a separate clear - but two calls that OVERLAP on the GPU (different streams, or different host threads
// without ordering) race on it. Calls on one stream are ordered and safe.
// ---------------------------------------------------------------------------------------------
constexpr uint32_t kNoAllocMaxBatch = DGB_NOALLOC_MAX_BATCH; // 32 (include/dietgpu_b200.h)
constexpr size_t noAllocRound(size_t v) { return (v + 255) / 256 * 256; }
constexpr size_t kNoAllocElemBytes = noAllocRound((kNoAllocMaxBatch + 1) * sizeof(Elem));
constexpr size_t kNoAllocInfoBytes = noAllocRound(2 * kNoAllocMaxBatch * sizeof(DecInfo));
constexpr size_t kNoAllocCounterBytes = 256;
constexpr size_t kNoAllocLutBytes = (size_t)2 * kNoAllocMaxBatch * (4u << 11); // two planes (fp64), probBits <= 11
constexpr size_t kNoAllocScratchBytes = kNoAllocElemBytes + kNoAllocInfoBytes + kNoAllocCounterBytes + kNoAllocLutBytes;

// In host code this name is only a stand-in for the symbol, NOT a device address: its address for the current device
// comes from cudaGetSymbolAddress (noAllocScratch below). Device code could name it directly.
static __device__ __align__(256) uint8_t g_noAllocScratch[kNoAllocScratchBytes];

struct NoAllocElems {
  Elem e[kNoAllocMaxBatch + 1];
};

// descriptor table: kernel argument -> scratch (replaces the H2D copy of the ordinary path)
__global__ void __launch_bounds__(64) k_noAllocLoadElems(const __grid_constant__ NoAllocElems t, Elem* __restrict__ dst,
                                                         uint32_t count) {
  const uint32_t* src = (const uint32_t*)t.e;
  uint32_t* d = (uint32_t*)dst;
  for (uint32_t i = threadIdx.x; i < count * (uint32_t)(sizeof(Elem) / 4); i += blockDim.x) d[i] = src[i];
}

// device address of the scratch array on the CURRENT device; looked up once per device, with that device current
static int noAllocScratch(uint8_t** out) {
  static uint8_t* ptrByDev[kMaxDevices] = {};
  static DeviceOnce once;
  const int dev = currentDevice();
  if (once.first(dev)) {
    uint8_t* ptr = nullptr;
    DGB_CUDA(cudaGetSymbolAddress((void**)&ptr, g_noAllocScratch));
    if (!ptr) return kErrCuda;
    __atomic_store_n(&ptrByDev[dev], ptr, __ATOMIC_RELEASE);
    once.done(dev);
  }
  *out = __atomic_load_n(&ptrByDev[dev], __ATOMIC_ACQUIRE);
  return *out ? kOk : kErrCuda;
}

int decodeNoAlloc(int ft, int pb, uint32_t num, const void* const* in, void* const* out, const uint32_t* cap,
                  uint8_t* successDev, uint32_t* sizeDev, cudaStream_t s) {
  if (num == 0) return kOk;
  if (!in || !out || !cap || pb < 9 || pb > 11 || ft < kFtNone || ft > kFtF64) return kErrInvalidArg;
  if (num > kNoAllocMaxBatch) {
    fprintf(stderr, "dietgpu_b200: dgb_decode_batch_noalloc: %u batch elements exceed the fixed scratch (maximum %u "
            "per call); split the batch or use the ordinary entry points\n", num, kNoAllocMaxBatch);
    return kErrTooLarge;
  }
  const uint32_t wb = ft == kFtNone ? 1u : floatWordBytes(ft);
  NoAllocElems tbl;
  memset(&tbl, 0, sizeof(tbl));
  uint64_t tiles = 0, plane = 0;
  for (uint32_t i = 0; i < num; ++i) {
    if (((uintptr_t)in[i]) & 15u) return kErrAlignment;
    if (((uintptr_t)out[i]) & (wb - 1)) return kErrAlignment;
    Elem& e = tbl.e[i];
    e.in = in[i];
    e.out = out[i];
    e.n = cap[i];
    e.tileBase = (uint32_t)tiles;
    e.planeOff = plane;
    const uint32_t t = divUp(cap[i], kTileSyms);
    tiles += t ? t : 1;
    plane += roundUpSz(cap[i], 128);
    if (tiles > 0x7fffffffull) return kErrInvalidArg;
  }
  tbl.e[num].tileBase = (uint32_t)tiles; // sentinel

  uint8_t* base = nullptr;
  int rc = noAllocScratch(&base);
  if (rc) return rc;
  Elem* elemsDev = (Elem*)base;
  DecInfo* info = (DecInfo*)(base + kNoAllocElemBytes);
  uint32_t* counter = (uint32_t*)(base + kNoAllocElemBytes + kNoAllocInfoBytes);
  uint32_t* lut = (uint32_t*)(base + kNoAllocElemBytes + kNoAllocInfoBytes + kNoAllocCounterBytes);
  const int planes = ft == kFtF64 ? 2 : 1;
  {
    ProfScope ps__("k_noAllocLoadElems", s);
    k_noAllocLoadElems<<<1, 64, 0, s>>>(tbl, elemsDev, num + 1);
  }
  DGB_CUDA(cudaGetLastError());
  DecodeArgs a;
  a.elems = elemsDev;
  a.numElems = num;
  a.totalTiles = (uint32_t)tiles;
  a.floatType = ft;
  a.sparseNested = 0;
  a.probBits = pb;
  a.outSuccess = successDev;
  a.outSize = sizeDev;
  a.denseBase = nullptr;
  {
    ProfScope ps__("k_decodePrep", s);
    k_decodePrep<<<dim3(num, planes), 256, 0, s>>>(a, info, lut, counter); // also zeroes the claim counter
  }
  DGB_CUDA(cudaGetLastError());
  DecParams p;
  p.a = a;
  p.info = info;
  p.lut = lut;
  p.claimTiles = 1;
  p.tileCounter = counter;
  switch (pb) {
    case 9: return launchDecodeFt<9>(p, s);
    case 10: return launchDecodeFt<10>(p, s);
    default: return launchDecodeFt<11>(p, s);
  }
}

} // namespace dgb

extern "C" int dgb_decode_batch_noalloc(int float_type, int prob_bits, uint32_t num_in_batch, const void* const* in,
                                        void* const* out, const uint32_t* out_capacity, uint8_t* out_success_dev,
                                        uint32_t* out_size_dev, void* stream) {
  return dgb::decodeNoAlloc(float_type, prob_bits, num_in_batch, in, out, out_capacity, out_success_dev, out_size_dev,
                            (cudaStream_t)stream);
}
