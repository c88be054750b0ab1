// Stand-alone float split kernel (sparse path, fp64, and the non-fused fallback): every CTA owns a contiguous
// range of 32768-word tiles. The per-tile work is splitTile in float_split.cuh.
#include <cstdlib>

#include "float_split.cuh"
#include "internal.h"
#include "prof.h"

namespace dgb {

template <int FT, bool PIPE>
__global__ void __launch_bounds__(256) k_floatSplit(const SplitParams p) {
  constexpr int kPlanes = FloatTraits<FT>::kPlanes;
  constexpr uint32_t kHistWords = splitHistWords<256>();
  extern __shared__ __align__(16) uint32_t sh[]; // kPlanes * kHistWords
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < kPlanes * kHistWords; i += 256) sh[i] = 0;
  __syncthreads();

  const uint32_t t0 = p.tileBegin + blockIdx.x * p.tilesPerCta;
  const uint32_t t1 = min(t0 + p.tilesPerCta, p.tileEnd);
  uint32_t cur = 0xffffffffu;
  for (uint32_t tile = t0; tile < t1; ++tile) {
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tile, cur);
    if (e != cur) {
      if (cur != 0xffffffffu) {
        __syncthreads();
        splitHistFlush<256>(sh, p.hist0 + cur * 256, tid);
        if (kPlanes == 2) splitHistFlush<256>(sh + kHistWords, p.hist1 + cur * 256, tid);
        __syncthreads();
      }
      cur = e;
    }
    const Elem el = p.elems[e];
    const uint32_t n = p.nDev ? p.nDev[e] : el.n;
    splitTile<FT, 256, PIPE>(p, e, el, n, tile, sh);
  }
  if (cur != 0xffffffffu) {
    __syncthreads();
    splitHistFlush<256>(sh, p.hist0 + cur * 256, tid);
    if (kPlanes == 2) splitHistFlush<256>(sh + kHistWords, p.hist1 + cur * 256, tid);
  }
}

template <int FT>
static int launchSplitT(SplitParams& p, cudaStream_t s) {
  constexpr uint32_t smemBytes = FloatTraits<FT>::kPlanes * splitHistWords<256>() * 4;
  static const bool pipe = getenv("DGB_SPLIT_PIPE") ? atoi(getenv("DGB_SPLIT_PIPE")) != 0 : true; // experiment knob
  auto kern = pipe ? k_floatSplit<FT, true> : k_floatSplit<FT, false>;
  DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
  const uint32_t tiles = p.tileEnd - p.tileBegin;
  uint32_t perSm = FT == kFtF64 ? 3u : 6u;
  if (const char* e = getenv("DGB_SPLIT_CTAS")) perSm = (uint32_t)atoi(e); // experiment knob
  uint32_t ctas = min(tiles, (uint32_t)smCount() * perSm);
  p.tilesPerCta = divUp(tiles, ctas);
  ctas = divUp(tiles, p.tilesPerCta);
  {
    ProfScope ps__("k_floatSplit", s);
    kern<<<ctas, 256, smemBytes, s>>>(p);
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
