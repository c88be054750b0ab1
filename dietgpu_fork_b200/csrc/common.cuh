// Shared device/host helpers and the stream-format constants of the DietGPU archives.
//
// Format facts restated from the reference (citations relative to /root/reference/):
//   ANS archive      dietgpu/ans/GpuANSUtils.cuh:17-229, dietgpu/ans/GpuANSEncode.cuh:511-624
//   float archive    dietgpu/float/GpuFloatUtils.cuh:19-96, :173-391
//   sparse archive   dietgpu/float/GpuFloatUtils.cuh:98-128
// Everything here is written for sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace dgb {

// ---- format constants -------------------------------------------------------------------
constexpr uint32_t kBlockSyms = 4096;      // symbols per ANS block (GpuANSUtils.cuh:37)
constexpr uint32_t kLanes = 32;            // interleaved rANS states per block (DeviceDefs.cuh:14)
constexpr uint32_t kStateMin = 1u << 15;   // kANSStartState == kANSMinState (GpuANSUtils.cuh:41-49)
constexpr uint32_t kAnsMagic = 0xD00D0001u;
constexpr uint32_t kFloatMagic = 0xF00F0001u;
constexpr uint32_t kAnsHeaderBytes = 32;
constexpr uint32_t kAnsPdfBytes = 512;
constexpr uint32_t kAnsFixedBytes = kAnsHeaderBytes + kAnsPdfBytes; // 544
constexpr uint32_t kFloatHeaderBytes = 32; // GpuFloatHeader + GpuFloatHeader2

// ---- tiling of the work (ours, not the format's) ------------------------------------------
// One CTA of 8 warps works on a "tile" of 8 consecutive ANS blocks of one batch element.
constexpr uint32_t kTileBlocks = 8;
constexpr uint32_t kTileSyms = kTileBlocks * kBlockSyms; // 32768
constexpr uint32_t kThreads = kTileBlocks * kLanes;      // 256
// Worst-case words one block can emit: each lane codes 128 symbols of <= probBits bits and
// ends holding >= 15 bits -> <= floor(128*probBits/16) words per lane; probBits 11 -> 88.
constexpr uint32_t kStageWords = 88 * 32;                // 2816 u16 = 5632 B per warp

enum FloatType : int { kFtNone = 0, kFtF16 = 1, kFtBF16 = 2, kFtF32 = 3, kFtF64 = 4 };

__host__ __device__ __forceinline__ uint32_t divUp(uint32_t a, uint32_t b) { return a / b + (a % b != 0); } // no wrap near 2^32
__host__ __device__ __forceinline__ uint32_t roundUp(uint32_t a, uint32_t b) { return divUp(a, b) * b; }
__host__ __device__ __forceinline__ size_t roundUpSz(size_t a, size_t b) { return (a + b - 1) / b * b; }

// bytes before the block payloads: header | pdf | states | blockWords (GpuANSUtils.cuh:68-82)
__host__ __device__ __forceinline__ uint32_t ansOverhead(uint32_t nb) {
  return kAnsFixedBytes + 128u * nb + 8u * roundUp(nb, 2);
}

__host__ __device__ __forceinline__ uint32_t floatWordBytes(int ft) {
  return ft == kFtF64 ? 8u : (ft == kFtF32 ? 4u : 2u);
}

// FloatTypeInfo<FT>::getUncompDataSize (GpuFloatUtils.cuh:204-208, :258-262, :303-314, :370-382)
__host__ __device__ __forceinline__ uint32_t floatRawBytes(int ft, uint32_t n) {
  switch (ft) {
    case kFtF16:
    case kFtBF16: return roundUp(n, 16);
    case kFtF32: return 2u * roundUp(n, 8) + roundUp(n, 16);
    case kFtF64: return 4u * roundUp(n, 4) + 2u * roundUp(n, 8);
    default: return 0;
  }
}

__host__ __device__ __forceinline__ uint32_t sparseBitmapBytes(uint32_t n) {
  return roundUp((n + 7) / 8, 16);
}

// ---- device-only helpers ---------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t laneId() {
  uint32_t l;
  asm("mov.u32 %0, %%laneid;" : "=r"(l));
  return l;
}
__device__ __forceinline__ uint32_t laneMaskLt() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
  return m;
}
__device__ __forceinline__ uint32_t laneMaskGe() {
  uint32_t m;
  asm("mov.u32 %0, %%lanemask_ge;" : "=r"(m));
  return m;
}

__device__ __forceinline__ uint32_t warpSum(uint32_t v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- shared-memory addresses, mbarrier and TMA 1-D bulk copies (sm_90+) --------------------------
__device__ __forceinline__ uint32_t smemAddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbarInit(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbarExpectTx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbarWait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra WAIT_DONE;\n"
      "bra WAIT_LOOP;\n"
      "WAIT_DONE:\n"
      "}\n" ::"r"(bar), "r"(parity)
      : "memory");
}
// global -> shared bulk copy (16-byte aligned source, destination and size), completion on `bar`
__device__ __forceinline__ void bulkLoad(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

// ---- checksum folded into the pass that reads the data anyway -------------------------------------------------
// The reference's checksum is the XOR of the covered bytes folded to 8 bits (dietgpu/ans/GpuChecksum.cuh:26-133), so
// any 32-bit XOR accumulation works whatever the alignment, as long as all four byte lanes are folded at the end.
// `covered` = number of leading bytes of the element that count (all of them for raw ANS; the first N BYTES of the
// N-word input for the float codecs: dietgpu/float/GpuFloatCompress.cuh:708-710).
// ckVec: the 16 bytes at byte offset 16 * g of the element.
__device__ __forceinline__ void ckVec(uint32_t& ck, uint4 x, uint32_t g, uint32_t covered) {
  const uint32_t full = covered >> 4;
  if (g < full) {
    ck ^= x.x ^ x.y ^ x.z ^ x.w;
  } else if (g == full) {
    const int k = (int)(covered & 15u); // leading bytes of this vector that count
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kb = k - 4 * j;
      if (kb >= 4) ck ^= w[j];
      else if (kb > 0) ck ^= w[j] & ((1u << (8 * kb)) - 1u);
    }
  }
}
// ckWord: one `bytes`-wide little-endian word at byte offset `off` of the element
__device__ __forceinline__ void ckWord(uint32_t& ck, unsigned long long v, uint32_t bytes, unsigned long long off,
                                       uint32_t covered) {
  if (off >= covered) return;
  const unsigned long long left = covered - off;
  if (left < bytes) v &= (1ull << (8 * left)) - 1ull;
  ck ^= (uint32_t)v ^ (uint32_t)(v >> 32);
}
__device__ __forceinline__ uint32_t ckFold8(uint32_t ck) {
  ck ^= ck >> 16;
  ck ^= ck >> 8;
  return ck & 0xffu;
}
// warp-reduces the lanes' accumulators and adds the folded result to the element's checksum word; clears ck
__device__ __forceinline__ void ckFlushWarp(uint32_t& ck, uint32_t* dst) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) ck ^= __shfl_xor_sync(0xffffffffu, ck, o);
  const uint32_t f = ckFold8(ck);
  if (laneId() == 0 && f) atomicXor(dst, f);
  ck = 0;
}

// ---- decoupled look-back tile state ------------------------------------------------------
// One 64-bit word per tile: [63:62] flag, [31:0] value. A single 8-byte store publishes flag and
// value together, so no fence is needed between them.
constexpr unsigned long long kTsAggregate = 1ull << 62;
constexpr unsigned long long kTsPrefix = 2ull << 62;

__device__ __forceinline__ void tsStore(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long tsLoad(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Executed by one full warp. `mine` is this tile's aggregate, `t` its index inside the element,
// `ts` the element's tile-state row. Returns the exclusive prefix (sum of aggregates of tiles
// 0..t-1) in every lane and publishes this tile's inclusive prefix. Tiles are handed out by a
// global ticket in increasing order, so every tile waited on belongs to a CTA that is already
// running: the spin cannot deadlock. A bounded spin turns a logic error into an error flag
// instead of a hung GPU.
__device__ __forceinline__ uint32_t lookbackExclusive(unsigned long long* ts, uint32_t t, uint32_t mine,
                                                      uint32_t* errFlag) {
  const uint32_t lane = laneId();
  if (t == 0) {
    if (lane == 0) tsStore(ts, kTsPrefix | mine);
    return 0;
  }
  if (lane == 0) tsStore(ts + t, kTsAggregate | mine);
  uint32_t excl = 0;
  int idx = (int)t - 1; // tile examined by lane 0
  uint32_t spins = 0;
  while (true) {
    int my = idx - (int)lane;
    unsigned long long v = kTsPrefix; // tiles before tile 0: prefix 0
    if (my >= 0) v = tsLoad(ts + my);
    // wait until all 32 examined tiles are published
    while (__any_sync(0xffffffffu, (v >> 62) == 0)) {
      if (++spins > (1u << 20)) {
        if (lane == 0 && errFlag) atomicExch(errFlag, 1u);
        break;
      }
      __nanosleep(32);
      if (my >= 0) v = tsLoad(ts + my);
    }
    uint32_t isPrefix = __ballot_sync(0xffffffffu, (v >> 62) == 2);
    uint32_t val = (uint32_t)v;
    if (isPrefix) {
      uint32_t first = __ffs(isPrefix) - 1; // nearest tile that already knows its inclusive prefix
      excl += warpSum(lane <= first ? val : 0u);
      break;
    }
    excl += warpSum(val);
    idx -= 32;
    if (spins > (1u << 20)) break;
  }
  if (lane == 0) tsStore(ts + t, kTsPrefix | (excl + mine));
  return excl;
}

// Work list: element i owns tiles [tb(i), tb(i+1)), where tb(i) = tileBase[i * strideWords] is the
// tileBase field of the i-th descriptor (descriptor numElems is a sentinel holding the total).
// Every element owns at least one tile. Finds the element of a global tile index, starting from a
// cached guess (a CTA visits tiles in increasing order).
__device__ __forceinline__ uint32_t findElement(const uint32_t* __restrict__ tileBase, uint32_t strideWords,
                                                uint32_t numElems, uint32_t tile, uint32_t guess) {
#define DGB_TB(i) tileBase[(size_t)(i) * strideWords]
  uint32_t lo = 0, hi = numElems; // invariant: tb(lo) <= tile < tb(hi)
  if (guess < numElems && DGB_TB(guess) <= tile) {
    if (tile < DGB_TB(guess + 1)) return guess;
    lo = guess + 1;
    if (tile < DGB_TB(lo + 1)) return lo;
  }
  while (hi - lo > 1) {
    uint32_t mid = (lo + hi) >> 1;
    if (DGB_TB(mid) <= tile) lo = mid; else hi = mid;
  }
#undef DGB_TB
  return lo;
}
#endif // __CUDACC__

} // namespace dgb
