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
#include "float_math.cuh"
#include "internal.h"
#include "prof.h"

namespace dgb {

struct DecParams {
  DecodeArgs a;
  const DecInfo* info;  // [planes][numElems]
  const uint32_t* lut;  // [planes][numElems][1 << probBits]
  uint32_t* ticket;
};

// ---------------------------------------------------------------------------------------------
// Prep: one CTA per (element, plane). Locates the ANS archive, checks capacity, reports
// success / size, and expands pdf[256] into the 2^probBits-entry LUT
//   entry = (slot - cdf[sym]) << 20 | pdf[sym] << 8 | sym.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_decodePrep(const DecodeArgs a, DecInfo* __restrict__ info,
                                                    uint32_t* __restrict__ lut) {
  __shared__ uint32_t sCdf[257];
  __shared__ uint32_t sWarp[8];
  const uint32_t e = blockIdx.x, plane = blockIdx.y, s = threadIdx.x, lane = s & 31, w = s >> 5;
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
  if (s == 0) {
    DecInfo d;
    d.ans = ans;
    d.fbase = a.floatType != kFtNone ? arch : nullptr;
    d.nb = ok ? nb : 0;
    d.n = ok ? n : 0;
    d.ok = ok;
    d.pad = 0;
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
// Tile decoder
// ---------------------------------------------------------------------------------------------
template <int PB>
__device__ __forceinline__ uint32_t decRow(uint32_t& state, const uint32_t* __restrict__ lut,
                                           const uint16_t* __restrict__ stage, uint32_t& ptr,
                                           uint32_t geMask, bool valid) {
  const uint32_t e = lut[state & ((1u << PB) - 1u)];
  uint32_t ns = ((e >> 8) & 0xfffu) * (state >> PB) + (e >> 20);
  state = valid ? ns : state;
  const bool rd = valid && state < kStateMin;
  const uint32_t vote = __ballot_sync(0xffffffffu, rd);
  if (rd) state = (state << 16) + stage[ptr - __popc(vote & geMask)];
  ptr -= __popc(vote);
  return e & 0xffu;
}

template <int FT>
__device__ __forceinline__ void writeTile(const DecInfo& d, const Elem& el, const DecodeArgs& a, uint32_t t,
                                          const uint8_t* sSym0, const uint8_t* sSym1, uint32_t tid) {
  const uint32_t off = t * kTileSyms;
  const uint32_t len = min(kTileSyms, d.n - off);
  if constexpr (FT == kFtNone) {
    uint8_t* out = (uint8_t*)el.out + off;
    if ((((uintptr_t)out) & 15u) == 0) {
      const uint32_t nv = len >> 4;
      for (uint32_t i = tid; i < nv; i += kThreads) ((uint4*)out)[i] = ((const uint4*)sSym0)[i];
      const uint32_t done = nv << 4;
      if (tid < len - done) out[done + tid] = sSym0[done + tid];
    } else {
      for (uint32_t i = tid; i < len; i += kThreads) out[i] = sSym0[i];
    }
  } else {
    using Tr = FloatTraits<FT>;
    constexpr uint32_t G = Tr::kGroup;
    uint8_t* outBase = a.denseBase ? a.denseBase + el.planeOff * sizeof(typename Tr::Word) : (uint8_t*)el.out;
    RawSections rs = rawSections(FT, (uint8_t*)d.fbase + kFloatHeaderBytes, d.n);
    uint32_t vecWords = 0;
    if ((((uintptr_t)outBase) & 15u) == 0) {
      const uint32_t ng = len / G;
      const uint32_t g0 = off / G;
      uint4* o = (uint4*)outBase;
      for (uint32_t i = tid; i < ng; i += kThreads)
        o[g0 + i] = joinGroup<FT>(g0 + i, sSym0 + i * G, sSym1 + i * G, rs);
      vecWords = ng * G;
    }
    for (uint32_t i = vecWords + tid; i < len; i += kThreads)
      joinScalar<FT>(outBase, off + i, sSym0[i], sSym1[i], rs);
  }
}

template <int PB, int FT>
__global__ void __launch_bounds__(256, FT == kFtF64 ? 1 : 2) k_decode(const DecParams p) {
  constexpr int kPlanes = FT == kFtF64 ? 2 : 1;
  constexpr uint32_t kLut = 1u << PB;
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* sSym = smem;                                                 // kPlanes * kTileSyms
  uint16_t* sStage = (uint16_t*)(smem + kPlanes * kTileSyms);          // 8 * kStageWords
  uint32_t* sLut = (uint32_t*)(smem + kPlanes * kTileSyms + kTileBlocks * kStageWords * 2);
  __shared__ uint32_t sTicket, sElem;

  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t geMask = laneMaskGe();
  const DecodeArgs& a = p.a;
  uint32_t guess = 0;

  while (true) {
    if (tid == 0) {
      uint32_t tk = atomicAdd(p.ticket, 1u);
      sTicket = tk;
      if (tk < a.totalTiles) sElem = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tk, guess);
    }
    __syncthreads();
    const uint32_t tk = sTicket;
    if (tk >= a.totalTiles) break;
    const uint32_t e = sElem;
    guess = e;
    const Elem el = a.elems[e];
    const uint32_t t = tk - el.tileBase;
    const DecInfo d0 = p.info[e];
    if (!d0.ok || t >= divUp(d0.nb, kTileBlocks)) {
      __syncthreads();
      continue;
    }
    const uint32_t nb = d0.nb;
    const uint32_t blk = t * kTileBlocks + w;

#pragma unroll
    for (int plane = 0; plane < kPlanes; ++plane) {
      const uint8_t* ans = plane == 0 ? d0.ans : p.info[a.numElems + e].ans;
      {
        const uint4* src = (const uint4*)(p.lut + ((size_t)plane * a.numElems + e) * kLut);
        for (uint32_t i = tid; i < kLut / 4; i += kThreads) ((uint4*)sLut)[i] = src[i];
      }
      uint32_t state = 0, len = 0, words = 0;
      uint16_t* stage = sStage + w * kStageWords;
      if (blk < nb) {
        state = ((const uint32_t*)(ans + kAnsFixedBytes))[blk * 32 + lane];
        const uint2 bw = ((const uint2*)(ans + kAnsFixedBytes + 128u * nb))[blk];
        len = min(bw.x >> 16, kBlockSyms);
        words = min(bw.x & 0xffffu, kStageWords);
        const uint4* src = (const uint4*)(ans + ansOverhead(nb) + 2 * (size_t)bw.y);
        const uint32_t nv = roundUp(words, 8) >> 3;
        for (uint32_t i = lane; i < nv; i += 32) ((uint4*)stage)[i] = src[i];
      }
      __syncthreads();
      if (len) {
        uint8_t* sym = sSym + plane * kTileSyms + w * kBlockSyms + lane;
        const uint32_t rows = len >> 5, rem = len & 31u;
        uint32_t ptr = words;
        if (rem) {
          const bool valid = lane < rem;
          uint32_t sv = decRow<PB>(state, sLut, stage, ptr, geMask, valid);
          if (valid) sym[rows * 32] = (uint8_t)sv;
        }
#pragma unroll 8
        for (int r = (int)rows - 1; r >= 0; --r) sym[r * 32] = (uint8_t)decRow<PB>(state, sLut, stage, ptr, geMask, true);
      }
      __syncthreads();
    }
    writeTile<FT>(d0, el, a, t, sSym, sSym + (kPlanes - 1) * kTileSyms, tid);
  }
}

template <int PB, int FT>
static int launchDecodeKernel(const DecParams& p, uint32_t grid, cudaStream_t s) {
  constexpr int kPlanes = FT == kFtF64 ? 2 : 1;
  constexpr uint32_t smemBytes = kPlanes * kTileSyms + kTileBlocks * kStageWords * 2 + (4u << PB);
  DGB_CUDA(cudaFuncSetAttribute(k_decode<PB, FT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
  {
    ProfScope ps__("k_decode", s);
    k_decode<PB, FT><<<grid, kThreads, smemBytes, s>>>(p);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

template <int PB>
static int launchDecodeFt(const DecParams& p, uint32_t grid, cudaStream_t s) {
  switch (p.a.floatType) {
    case kFtNone: return launchDecodeKernel<PB, kFtNone>(p, grid, s);
    case kFtF16: return launchDecodeKernel<PB, kFtF16>(p, grid, s);
    case kFtBF16: return launchDecodeKernel<PB, kFtBF16>(p, grid, s);
    case kFtF32: return launchDecodeKernel<PB, kFtF32>(p, grid, s);
    case kFtF64: return launchDecodeKernel<PB, kFtF64>(p, grid, s);
    default: return kErrInvalidArg;
  }
}

int launchDecode(Temp& tmp, const DecodeArgs& a, cudaStream_t s) {
  if (a.numElems == 0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  const int planes = a.floatType == kFtF64 ? 2 : 1;
  DecInfo* info = tmp.alloc<DecInfo>((size_t)planes * a.numElems);
  uint32_t* lut = tmp.alloc<uint32_t>(((size_t)planes * a.numElems) << a.probBits);
  uint32_t* ticket = tmp.alloc<uint32_t>(4);
  if (tmp.failed()) return kErrCuda;
  DGB_CUDA(cudaMemsetAsync(ticket, 0, 16, s));
  {
    ProfScope ps__("k_decodePrep", s);
    k_decodePrep<<<dim3(a.numElems, planes), 256, 0, s>>>(a, info, lut);
  }
  DGB_CUDA(cudaGetLastError());
  DecParams p;
  p.a = a;
  p.info = info;
  p.lut = lut;
  p.ticket = ticket;
  uint32_t perSm = a.floatType == kFtF64 ? 1u : 2u;
  uint32_t grid = min(a.totalTiles, (uint32_t)smCount() * perSm);
  if (grid == 0) return kOk;
  switch (a.probBits) {
    case 9: return launchDecodeFt<9>(p, grid, s);
    case 10: return launchDecodeFt<10>(p, grid, s);
    default: return launchDecodeFt<11>(p, grid, s);
  }
}

} // namespace dgb
