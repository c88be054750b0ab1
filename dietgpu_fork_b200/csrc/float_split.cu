// Float split for sm_100a: word -> exponent plane(s) + raw section + fused exponent histogram(s).
//
// What is computed follows the reference's splitFloat (dietgpu/float/GpuFloatCompress.cuh:44-551):
// the raw bits go straight into the final archive, the exponent bytes into a temp plane that the
// ANS encoder reads next, and the float header is written by the element's first tile. Differences
// in how: every CTA owns a contiguous range of 32768-word tiles, each thread handles one 16-byte
// input vector per step with warp-contiguous loads and stores, histograms live in per-warp 4-way
// replicated shared-memory bins, and all padding the reference leaves undefined is written as 0.
#include <cstdlib>

#include "float_math.cuh"
#include "internal.h"
#include "prof.h"

namespace dgb {

constexpr int kSplitRep = 4;
constexpr uint32_t kSplitHistWords = 8 * 256 * kSplitRep; // per plane

struct SplitParams {
  const Elem* elems;
  uint32_t numElems;
  uint32_t tileBegin, tileEnd; // range of 32768-word tiles handled by this launch
  uint32_t tilesPerCta;
  const uint32_t* nDev;     // nullable (sparse: nnz)
  const uint8_t* denseIn;   // nullable: input rows at denseIn + planeOff * wordBytes (sparse)
  uint8_t* plane0;
  uint8_t* plane1;
  uint32_t* hist0;          // [numElems][256]
  uint32_t* hist1;
  const uint32_t* checksum; // nullable
  int useChecksum;
};

__device__ __forceinline__ void splitHistAdd(uint32_t* wh, uint32_t sym, uint32_t rep) {
  atomicAdd(&wh[sym * kSplitRep + rep], 1u);
}

__device__ __forceinline__ void splitHistFlush(uint32_t* sh, uint32_t* g, uint32_t tid) {
  uint32_t sum = 0;
#pragma unroll
  for (int w = 0; w < 8; ++w) {
#pragma unroll
    for (int r = 0; r < kSplitRep; ++r) {
      uint32_t idx = w * 256 * kSplitRep + tid * kSplitRep + r;
      sum += sh[idx];
      sh[idx] = 0;
    }
  }
  if (sum) atomicAdd(&g[tid], sum);
}

template <int FT>
__global__ void __launch_bounds__(256) k_floatSplit(const SplitParams p) {
  using Tr = FloatTraits<FT>;
  using Word = typename Tr::Word;
  constexpr uint32_t G = Tr::kGroup;
  constexpr int kPlanes = Tr::kPlanes;
  extern __shared__ __align__(16) uint32_t sh[]; // kPlanes * kSplitHistWords
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < kPlanes * kSplitHistWords; i += 256) sh[i] = 0;
  __syncthreads();
  uint32_t* wh0 = sh + (tid >> 5) * 256 * kSplitRep;
  uint32_t* wh1 = wh0 + kSplitHistWords;
  const uint32_t rep = tid & (kSplitRep - 1);

  const uint32_t t0 = p.tileBegin + blockIdx.x * p.tilesPerCta;
  const uint32_t t1 = min(t0 + p.tilesPerCta, p.tileEnd);
  uint32_t cur = 0xffffffffu;
  for (uint32_t tile = t0; tile < t1; ++tile) {
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tile, cur);
    if (e != cur) {
      if (cur != 0xffffffffu) {
        __syncthreads();
        splitHistFlush(sh, p.hist0 + cur * 256, tid);
        if (kPlanes == 2) splitHistFlush(sh + kSplitHistWords, p.hist1 + cur * 256, tid);
        __syncthreads();
      }
      cur = e;
    }
    const Elem el = p.elems[e];
    const uint32_t n = p.nDev ? p.nDev[e] : el.n;
    const uint32_t t = tile - el.tileBase;
    const uint32_t off = t * kTileSyms;
    uint8_t* fbase = (uint8_t*)el.out + el.sizeBias;
    if (t == 0 && tid == 0) {
      // GpuFloatHeader + GpuFloatHeader2 (dietgpu/float/GpuFloatUtils.cuh:19-96); the second
      // header's firstCompSegmentBytes is filled in by the encoder
      uint32_t cs = (p.useChecksum && p.checksum) ? p.checksum[e] : 0u;
      ((uint4*)fbase)[0] = make_uint4(kFloatMagic, n, (uint32_t)FT | (p.useChecksum ? 16u : 0u), cs);
      ((uint4*)fbase)[1] = make_uint4(0u, 0u, 0u, 0u);
    }
    if (off >= n) continue;
    const uint32_t len = min(kTileSyms, n - off);
    const uint8_t* in = p.denseIn ? p.denseIn + el.planeOff * sizeof(Word) : (const uint8_t*)el.in;
    uint8_t* plane0 = p.plane0 + el.planeOff;
    uint8_t* plane1 = kPlanes == 2 ? p.plane1 + el.planeOff : plane0;
    const RawSections rs = rawSections(FT, fbase + kFloatHeaderBytes, n);

    uint32_t vecWords = 0;
    if ((((uintptr_t)in) & 15u) == 0) {
      const uint32_t ng = len / G;
      const uint32_t g0 = off / G;
      const uint4* v = (const uint4*)in + g0;
      uint32_t i = tid;
      for (; i + 768 < ng; i += 1024) {
        uint4 x[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) x[k] = v[i + 256 * k];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          uint32_t ex[2], ex1[2];
          splitGroup<FT>(x[k], g0 + i + 256 * k, plane0, plane1, rs, ex, ex1);
          if constexpr (G == 8) {
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              splitHistAdd(wh0, (ex[0] >> (8 * b)) & 0xffu, rep);
              splitHistAdd(wh0, (ex[1] >> (8 * b)) & 0xffu, rep);
            }
          } else if constexpr (G == 4) {
#pragma unroll
            for (int b = 0; b < 4; ++b) splitHistAdd(wh0, (ex[0] >> (8 * b)) & 0xffu, rep);
          } else {
            splitHistAdd(wh0, ex[0] & 0xffu, rep);
            splitHistAdd(wh0, ex[0] >> 8, rep);
            splitHistAdd(wh1, ex1[0] & 0xffu, rep);
            splitHistAdd(wh1, ex1[0] >> 8, rep);
          }
        }
      }
      for (; i < ng; i += 256) {
        uint32_t ex[2], ex1[2];
        splitGroup<FT>(v[i], g0 + i, plane0, plane1, rs, ex, ex1);
        if constexpr (G == 8) {
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            splitHistAdd(wh0, (ex[0] >> (8 * b)) & 0xffu, rep);
            splitHistAdd(wh0, (ex[1] >> (8 * b)) & 0xffu, rep);
          }
        } else if constexpr (G == 4) {
#pragma unroll
          for (int b = 0; b < 4; ++b) splitHistAdd(wh0, (ex[0] >> (8 * b)) & 0xffu, rep);
        } else {
          splitHistAdd(wh0, ex[0] & 0xffu, rep);
          splitHistAdd(wh0, ex[0] >> 8, rep);
          splitHistAdd(wh1, ex1[0] & 0xffu, rep);
          splitHistAdd(wh1, ex1[0] >> 8, rep);
        }
      }
      vecWords = ng * G;
    }
    // scalar path: whole tile for inputs that are not 16-byte aligned, else the tail
    for (uint32_t i = vecWords + tid; i < len; i += 256) {
      uint32_t e0, e1;
      splitScalar<FT>(in, off + i, plane0, plane1, rs, e0, e1);
      splitHistAdd(wh0, e0, rep);
      if (kPlanes == 2) splitHistAdd(wh1, e1, rep);
    }
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
  if (cur != 0xffffffffu) {
    __syncthreads();
    splitHistFlush(sh, p.hist0 + cur * 256, tid);
    if (kPlanes == 2) splitHistFlush(sh + kSplitHistWords, p.hist1 + cur * 256, tid);
  }
}

template <int FT>
static int launchSplitT(SplitParams& p, cudaStream_t s) {
  constexpr uint32_t smemBytes = FloatTraits<FT>::kPlanes * kSplitHistWords * 4;
  DGB_CUDA(cudaFuncSetAttribute(k_floatSplit<FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
  const uint32_t tiles = p.tileEnd - p.tileBegin;
  uint32_t perSm = FT == kFtF64 ? 3u : 6u;
  if (const char* e = getenv("DGB_SPLIT_CTAS")) perSm = (uint32_t)atoi(e); // experiment knob
  uint32_t ctas = min(tiles, (uint32_t)smCount() * perSm);
  p.tilesPerCta = divUp(tiles, ctas);
  ctas = divUp(tiles, p.tilesPerCta);
  {
    ProfScope ps__("k_floatSplit", s);
    k_floatSplit<FT><<<ctas, 256, smemBytes, s>>>(p);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

int launchFloatSplit(int ft, const Elem* elems, uint32_t numElems, uint32_t tileBegin, uint32_t tileEnd,
                     const uint32_t* nDev,
                     const uint8_t* denseIn, uint8_t* plane0, uint8_t* plane1, uint32_t* hist0,
                     uint32_t* hist1, const uint32_t* checksum, int useChecksum, cudaStream_t s) {
  if (tileEnd <= tileBegin) return kOk;
  SplitParams p;
  p.elems = elems;
  p.numElems = numElems;
  p.tileBegin = tileBegin;
  p.tileEnd = tileEnd;
  p.tilesPerCta = 1;
  p.nDev = nDev;
  p.denseIn = denseIn;
  p.plane0 = plane0;
  p.plane1 = plane1;
  p.hist0 = hist0;
  p.hist1 = hist1;
  p.checksum = checksum;
  p.useChecksum = useChecksum;
  switch (ft) {
    case kFtF16: return launchSplitT<kFtF16>(p, s);
    case kFtBF16: return launchSplitT<kFtBF16>(p, s);
    case kFtF32: return launchSplitT<kFtF32>(p, s);
    case kFtF64: return launchSplitT<kFtF64>(p, s);
    default: return kErrInvalidArg;
  }
}

} // namespace dgb
