// Device-side float split shared by k_floatSplit (float_split.cu) and the fused compress kernel (fused.cu):
// word -> exponent plane(s) + raw section + per-warp shared-memory exponent histogram(s).
//
// What is computed follows the reference's splitFloat (dietgpu/float/GpuFloatCompress.cuh:44-551): the raw
// bits go straight into the final archive, the exponent bytes into a temp plane that the ANS encoder reads
// next, and the float header is written by the element's first tile. Differences in how: work is cut into
// 32768-word tiles, each thread handles one 16-byte input vector per step with warp-contiguous loads and
// stores, histograms live in per-warp shared-memory bins, and all padding the reference
// leaves undefined is written as 0.
#pragma once
#include "float_math.cuh"
#include "internal.h"

namespace dgb {

// Replicas of a warp's bins (1, 2 or 4; lane l counts into replica l % kSplitRep). One: `ATOMS.POPC.INC` merges the
// lanes of an instruction that hit the same counter, and what an increment costs follows the number of DISTINCT
// counters it touches - replicas only multiply those. B200, replicas 4 / 2 / 1: fp16 128 x 4 MiB 0.480 / 0.470 /
// 0.461 ms, fp64 16 x 8 MiB 0.146 / 0.141 / 0.141, ANS 16 x 1 MiB 59.1 / 58.3 / 57.9 us, bf16 and fp32 equal.
#ifndef DGB_SPLIT_REP
#define DGB_SPLIT_REP 1
#endif
constexpr int kSplitRep = DGB_SPLIT_REP;
constexpr uint32_t kSplitSymShift = kSplitRep == 4 ? 4 : kSplitRep == 2 ? 3 : 2; // a symbol's place in a counter's address
static_assert(kSplitRep == 1 || kSplitRep == 2 || kSplitRep == 4, "");
constexpr uint32_t kSplitPrefetchVecs = 2048; // 32 KB ahead of the loads
template <int NT> __host__ __device__ constexpr uint32_t splitHistWords() { return (NT / 32) * 256 * kSplitRep; } // per plane

struct SplitParams {
  const Elem* elems;
  uint32_t numElems;
  uint32_t tileBegin, tileEnd; // range of 32768-word tiles handled by this launch
  uint32_t tilesPerCta;
  // stand-alone kernel, batches of few huge elements: work units of (tile >> unitShift) words dealt out round robin
  // (unit u -> CTA u % grid), so that at any moment the CTAs stream one contiguous window of memory instead of one
  // far-apart range each (DRAM efficiency); 0 / 0: contiguous ranges of whole tiles
  uint32_t unitShift;
  uint32_t roundRobin;   // 1: over the whole work list, 2: inside every element (see k_floatSplit)
  uint32_t minElemTiles; // host hint: tiles of the shortest element (0: unknown)
  const uint32_t* nDev;     // nullable (sparse: nnz)
  const uint8_t* denseIn;   // nullable: input rows at denseIn + planeOff * wordBytes (sparse)
  uint8_t* plane0;
  uint8_t* plane1;
  uint32_t* hist0;          // [numElems][256]
  uint32_t* hist1;
  const uint32_t* checksum; // nullable: checksums computed beforehand (k_checksum)
  int useChecksum;
  // nullable, zeroed [numElems]: the split pass itself accumulates the checksum (one XOR per loaded vector) and the
  // CTA that completes an element writes it into the float header; needs the completion protocol (normTilesDone in
  // k_floatSplit, always in the fused kernel). When set, `checksum` is ignored.
  uint32_t* checksumAcc;
  // stand-alone kernel: when set, the CTA whose tiles complete an element's histogram(s) also builds its encode
  // table(s) - the k_normalize launch disappears from small batches, which are launch-latency bound
  uint32_t* normTilesDone;  // nullable; zeroed [numElems]
  uint2* normTable0;        // [numElems][256]
  uint2* normTable1;        // second exponent plane (fp64)
  int normProbBits;
};

// The CTA's bins start at a SHARED address aligned to one warp's bins (splitHistBase): a warp's 256 x kSplitRep
// counters are 1 KB (x kSplitRep), so a counter's address is  warp base | symbol << kSplitSymShift | replica << 2
// and an increment costs a shift, one LOP3 and the ATOMS (an index the compiler has to add to a base costs one
// instruction more per symbol).
constexpr uint32_t kSplitHistAlign = 256 * kSplitRep * 4;
// first kSplitHistAlign-aligned byte of a dynamic shared-memory block (callers allocate kSplitHistAlign more)
__device__ __forceinline__ uint8_t* splitHistBase(uint8_t* smem) {
  const uint32_t a = smemAddr(smem);
  return smem + ((kSplitHistAlign - (a & (kSplitHistAlign - 1u))) & (kSplitHistAlign - 1u));
}
// whRep: shared address of the warp's bins | replica * 4; symShl: the symbol shifted left by kSplitSymShift, anything
// elsewhere
__device__ __forceinline__ void splitHistAddAt(uint32_t whRep, uint32_t symShl) {
  const uint32_t addr = (symShl & (0xffu << kSplitSymShift)) | whRep;
  asm volatile("red.shared.add.u32 [%0], 1;" ::"r"(addr) : "memory");
}
// the four bytes of x
__device__ __forceinline__ void splitHistAdd4(uint32_t whRep, uint32_t x) {
  splitHistAddAt(whRep, x << kSplitSymShift);
  splitHistAddAt(whRep, x >> (8 - kSplitSymShift));
  splitHistAddAt(whRep, x >> (16 - kSplitSymShift));
  splitHistAddAt(whRep, x >> (24 - kSplitSymShift));
}

// sums the warps x replicas of the bins this thread owns, adds to global, clears. Callers bracket it with
// __syncthreads().
template <int NT>
__device__ __forceinline__ void splitHistFlush(uint32_t* sh, uint32_t* g, uint32_t tid) {
#pragma unroll
  for (int k = 0; k < 256 / NT; ++k) {
    const uint32_t bin = tid + k * NT;
    uint32_t sum = 0;
#pragma unroll
    for (int w = 0; w < NT / 32; ++w) {
#pragma unroll
      for (int r = 0; r < kSplitRep; ++r) {
        const uint32_t idx = w * 256 * kSplitRep + bin * kSplitRep + r;
        sum += sh[idx];
        sh[idx] = 0;
      }
    }
    if (sum) atomicAdd(&g[bin], sum);
  }
}

template <int FT, int G>
__device__ __forceinline__ void splitHistGroup(uint32_t wh0, uint32_t wh1, const uint32_t* ex, const uint32_t* ex1) {
  if constexpr (G == 8) {
    splitHistAdd4(wh0, ex[0]);
    splitHistAdd4(wh0, ex[1]);
  } else if constexpr (G == 4) {
    splitHistAdd4(wh0, ex[0]);
  } else {
    splitHistAddAt(wh0, ex[0] << kSplitSymShift);
    splitHistAddAt(wh0, ex[0] >> (8 - kSplitSymShift));
    splitHistAddAt(wh1, ex1[0] << kSplitSymShift);
    splitHistAddAt(wh1, ex1[0] >> (8 - kSplitSymShift));
  }
}

// One tile of element e (el = elems[e], n = its word count) by the NT threads of the CTA, or the slice
// [slice * sliceWords, (slice + 1) * sliceWords) of it (sliceWords: a power of two dividing the tile). sh: the
// CTA's histogram bins (kPlanes * splitHistWords<NT>() words at a kSplitHistAlign-aligned shared address, zeroed by
// the caller).
template <int FT, int NT, bool PIPE = true>
__device__ __forceinline__ void splitTile(const SplitParams& p, uint32_t e, const Elem& el, uint32_t n, uint32_t tile,
                                          uint32_t* sh, uint32_t sliceWords = kTileSyms, uint32_t slice = 0,
                                          uint32_t* ckAcc = nullptr) {
  using Tr = FloatTraits<FT>;
  using Word = typename Tr::Word;
  constexpr uint32_t G = Tr::kGroup;
  constexpr int kPlanes = Tr::kPlanes;
  const uint32_t tid = threadIdx.x;
  // this warp's bins | this lane's replica (sh is kSplitHistAlign-aligned, see splitHistAddAt)
  const uint32_t wh0 = smemAddr(sh) + (tid >> 5) * kSplitHistAlign + (tid & (kSplitRep - 1)) * 4u;
  const uint32_t wh1 = wh0 + splitHistWords<NT>() * 4u;
  const uint32_t t = tile - el.tileBase;
  const uint32_t sliceOff = slice * sliceWords;
  const uint32_t off = t * kTileSyms + sliceOff;
  uint8_t* fbase = (uint8_t*)el.out + el.sizeBias;
  if (t == 0 && slice == 0 && tid == 0) {
    // GpuFloatHeader + GpuFloatHeader2 (dietgpu/float/GpuFloatUtils.cuh:19-96); the second
    // header's firstCompSegmentBytes is filled in by the encoder
    uint32_t cs = (p.useChecksum && p.checksum && !p.checksumAcc) ? p.checksum[e] : 0u; // folded: patched at completion
    ((uint4*)fbase)[0] = make_uint4(kFloatMagic, n, (uint32_t)FT | (p.useChecksum ? 16u : 0u), cs);
    ((uint4*)fbase)[1] = make_uint4(0u, 0u, 0u, 0u);
  }
  if (off >= n) return;
  const uint32_t len = min(sliceWords, n - off);
  const uint8_t* in = p.denseIn ? p.denseIn + el.planeOff * sizeof(Word) : (const uint8_t*)el.in;
  uint8_t* plane0 = p.plane0 + el.planeOff;
  uint8_t* plane1 = kPlanes == 2 ? p.plane1 + el.planeOff : plane0;
  const RawSections rs = rawSections(FT, fbase + kFloatHeaderBytes, n);
  // checksum folded into this pass: the first n BYTES of the n-word input count (0: off)
  const uint32_t covered = (ckAcc && p.checksumAcc) ? n : 0u;
  uint32_t ck = 0;

  uint32_t vecWords = 0;
  if ((((uintptr_t)in) & 15u) == 0) {
    const uint32_t ng = len / G;
    const uint32_t g0 = off / G;
    const uint4* v = (const uint4*)in + g0;
    uint32_t i = tid;
    if constexpr (!PIPE) { // four independent vectors per thread and step
      for (; i + 3 * NT < ng; i += 4 * NT) {
        uint4 x[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) x[k] = v[i + NT * k];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t ex[2], ex1[2];
          if (covered) ckVec(ck, x[k], g0 + i + NT * k, covered);
          splitGroup<FT>(x[k], g0 + i + NT * k, plane0, plane1, rs, ex, ex1);
          splitHistGroup<FT, G>(wh0, wh1, ex, ex1);
        }
      }
    }
    // two vectors per thread and step, the next pair requested before the current one is processed
    if (PIPE && i + NT < ng) {
      uint4 x0 = v[i], x1 = v[i + NT];
      // input ahead of the register pipeline is pulled into L2 (one request per lane = 512 B per warp): with
      // two 16-byte loads per thread in flight, DRAM latency would otherwise cap the read rate
      const uint4* vEnd = (const uint4*)in + (n * sizeof(Word)) / 16;
      for (;;) {
        const uint32_t j = i + 2 * NT;
        const bool more = j + NT < ng;
        if (v + j + kSplitPrefetchVecs < vEnd) {
          asm volatile("prefetch.global.L2 [%0];" ::"l"(v + j + kSplitPrefetchVecs));
          asm volatile("prefetch.global.L2 [%0];" ::"l"(v + j + NT + kSplitPrefetchVecs));
        }
        uint4 y0, y1;
        if (more) {
          y0 = v[j];
          y1 = v[j + NT];
        }
        uint32_t ex[2], ex1[2];
        if (covered) {
          ckVec(ck, x0, g0 + i, covered);
          ckVec(ck, x1, g0 + i + NT, covered);
        }
        splitGroup<FT>(x0, g0 + i, plane0, plane1, rs, ex, ex1);
        splitHistGroup<FT, G>(wh0, wh1, ex, ex1);
        splitGroup<FT>(x1, g0 + i + NT, plane0, plane1, rs, ex, ex1);
        splitHistGroup<FT, G>(wh0, wh1, ex, ex1);
        i = j;
        if (!more) break;
        x0 = y0;
        x1 = y1;
      }
    }
    for (; i < ng; i += NT) {
      uint32_t ex[2], ex1[2];
      const uint4 xv = v[i];
      if (covered) ckVec(ck, xv, g0 + i, covered);
      splitGroup<FT>(xv, g0 + i, plane0, plane1, rs, ex, ex1);
      splitHistGroup<FT, G>(wh0, wh1, ex, ex1);
    }
    vecWords = ng * G;
  }
  // scalar path: whole tile for inputs that are not 16-byte aligned, else the tail
  for (uint32_t i = vecWords + tid; i < len; i += NT) {
    uint32_t e0, e1;
    if (covered) ckWord(ck, (unsigned long long)((const Word*)in)[off + i], sizeof(Word),
                        (unsigned long long)(off + i) * sizeof(Word), covered);
    splitScalar<FT>(in, off + i, plane0, plane1, rs, e0, e1);
    splitHistAddAt(wh0, e0 << kSplitSymShift);
    if (kPlanes == 2) splitHistAddAt(wh1, e1 << kSplitSymShift);
  }
  if (covered) *ckAcc ^= ck;
  // zero the section padding after the last word (undefined in the reference)
  if (off + len == n && tid < 16) {
    const uint32_t i = n + tid;
    if constexpr (FT == kFtF16 || FT == kFtBF16) {
      if (i < roundUp(n, 16)) rs.a[i] = 0;
    } else if constexpr (FT == kFtF32) {
      if (i < roundUp(n, 8)) ((uint16_t*)rs.a)[i] = 0;
      if (i < roundUp(n, 16)) rs.b[i] = 0;
    } else {
      if (i < roundUp(n, 4)) ((uint32_t*)rs.a)[i] = 0;
      if (i < roundUp(n, 8)) ((uint16_t*)rs.b)[i] = 0;
    }
  }
}

} // namespace dgb
