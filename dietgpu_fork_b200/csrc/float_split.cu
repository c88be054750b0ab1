// Stand-alone float split kernel (sparse path, fp64, and the non-fused fallback): every CTA owns a contiguous
// range of 32768-word tiles. The per-tile work is splitTile in float_split.cuh.
#include <cstdlib>

#include "ans_encode.cuh"
#include "float_split.cuh"
#include "internal.h"
#include "prof.h"

namespace dgb {

template <int FT, bool PIPE>
__global__ void __launch_bounds__(256) k_floatSplit(const SplitParams p) {
  constexpr int kPlanes = FloatTraits<FT>::kPlanes;
  constexpr uint32_t kHistWords = splitHistWords<256>();
  extern __shared__ __align__(16) uint8_t smemRaw[]; // kSplitHistAlign + kPlanes * kHistWords * 4
  uint32_t* const sh = (uint32_t*)splitHistBase(smemRaw);
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < kPlanes * kHistWords; i += 256) sh[i] = 0;
  __syncthreads();

  // work units: whole tiles in a contiguous range per CTA, or tile slices dealt out round robin (SplitParams)
  const uint32_t unitShift = p.unitShift, perTile = 1u << unitShift, unitWords = kTileSyms >> unitShift;
  uint32_t unitsTotal = (p.tileEnd - p.tileBegin) << unitShift;
  uint32_t u0 = p.roundRobin ? blockIdx.x : min(blockIdx.x * p.tilesPerCta, unitsTotal);
  uint32_t u1 = p.roundRobin ? unitsTotal : min(u0 + p.tilesPerCta, unitsTotal);
  uint32_t uStep = p.roundRobin ? gridDim.x : 1u;
  if (p.roundRobin == 2) {
    // round robin per element: CTA b starts at tile floor(b * tiles / grid); the CTAs whose start falls into one
    // element share that element (every element is at least tiles / grid tiles long, checked by the host), so a
    // CTA flushes its histogram once and every element is swept by one contiguous front
    const uint32_t tiles = p.tileEnd - p.tileBegin, grid = gridDim.x;
    const uint32_t tb = (uint32_t)((uint64_t)blockIdx.x * tiles / grid);
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, p.tileBegin + tb, 0);
    const uint32_t eb = p.elems[e].tileBase - p.tileBegin, ee = p.elems[e + 1].tileBase - p.tileBegin;
    const uint32_t b0 = (uint32_t)(((uint64_t)eb * grid + tiles - 1) / tiles);
    const uint32_t b1 = min(grid, (uint32_t)(((uint64_t)ee * grid + tiles - 1) / tiles));
    u0 = (eb << unitShift) + (blockIdx.x - b0);
    u1 = ee << unitShift;
    uStep = b1 - b0;
  }
  uint32_t cur = 0xffffffffu, acc = 0; // element whose counts are in the bins, units of it done by this CTA
  uint32_t ck = 0;                     // this thread's share of the element's checksum (folded pass, SplitParams)
  __shared__ uint32_t sLast;
  // this CTA's counts of element `cur` go to the global bins; if they were the element's last, its table(s) follow
  auto finishElement = [&]() {
    __syncthreads();
    splitHistFlush<256>(sh, p.hist0 + cur * 256, tid);
    if (kPlanes == 2) splitHistFlush<256>(sh + kHistWords, p.hist1 + cur * 256, tid);
    if (p.normTilesDone) {
      const uint32_t eb = p.elems[cur].tileBase, ee = p.elems[cur + 1].tileBase;
      if (p.checksumAcc) ckFlushWarp(ck, p.checksumAcc + cur); // before the fence, like the histogram
      __threadfence();
      __syncthreads();
      if (tid == 0) sLast = atomicAdd(&p.normTilesDone[cur], acc) + acc == (ee - eb) << unitShift ? 1u : 0u;
      __syncthreads();
      if (sLast) {
        __threadfence();
        if (p.checksumAcc && tid == 0) { // every CTA's share is in: the float header gets its checksum
          uint8_t* fbase = (uint8_t*)p.elems[cur].out + p.elems[cur].sizeBias;
          ((uint32_t*)fbase)[3] = __ldcg(p.checksumAcc + cur);
        }
        const uint32_t total = p.nDev ? p.nDev[cur] : p.elems[cur].n;
        normalizeElement<256>(p.hist0 + (size_t)cur * 256, total, p.normProbBits, p.normTable0 + (size_t)cur * 256, sh);
        if (kPlanes == 2)
          normalizeElement<256>(p.hist1 + (size_t)cur * 256, total, p.normProbBits, p.normTable1 + (size_t)cur * 256, sh);
        __syncthreads();
        for (uint32_t i = tid; i < 256 + 16; i += 256) sh[i] = 0; // the bins normalizeElement used as scratch
      }
    }
    __syncthreads();
    acc = 0;
  };
  for (uint32_t u = u0; u < u1; u += uStep) {
    const uint32_t tile = p.tileBegin + (u >> unitShift);
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tile, cur);
    if (e != cur) {
      if (cur != 0xffffffffu) finishElement();
      cur = e;
    }
    const Elem el = p.elems[e];
    const uint32_t n = p.nDev ? p.nDev[e] : el.n;
    splitTile<FT, 256, PIPE>(p, e, el, n, tile, sh, unitWords, u & (perTile - 1u), &ck);
    ++acc;
  }
  if (cur != 0xffffffffu) finishElement();
}

template <int FT>
static int launchSplitT(SplitParams& p, cudaStream_t s) {
  constexpr uint32_t smemBytes = kSplitHistAlign + FloatTraits<FT>::kPlanes * splitHistWords<256>() * 4;
  const bool pipe = DGB_KNOB("DGB_SPLIT_PIPE", 1) != 0; // experiment knob
  auto kern = pipe ? k_floatSplit<FT, true> : k_floatSplit<FT, false>;
  {
    static DeviceOnce once[2];
    const int dev = currentDevice();
    if (once[pipe].first(dev)) {
      DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
      once[pipe].done(dev);
    }
  }
  const uint32_t tiles = p.tileEnd - p.tileBegin;
  uint32_t perSm = FT == kFtF64 ? 3u : 6u;
  if (DGB_KNOB("DGB_SPLIT_CTAS", 0) > 0) perSm = (uint32_t)DGB_KNOB("DGB_SPLIT_CTAS", 0); // experiment knob
  uint32_t ctas = min(tiles, (uint32_t)smCount() * perSm);
  p.tilesPerCta = divUp(tiles, ctas);
  ctas = divUp(tiles, p.tilesPerCta);
  p.unitShift = 0;
  p.roundRobin = 0;
  {
    // few huge elements (every CTA gets at least a tile's worth of each element anyway, so the histogram flush
    // per element change stays rare): half tiles round robin over a full grid. B200, bf16 1 x 256 MiB: contiguous
    // ranges 123.7 us, round robin in units of 32768 / 16384 / 8192 / 4096 words 108.5 / 104.3 / 110.7 / 121.1 us
    const uint32_t full = (uint32_t)smCount() * perSm;
    bool rr = (uint64_t)tiles >= (uint64_t)full * p.numElems;
    if (DGB_KNOB("DGB_SPLIT_RR", kKnobUnset) != kKnobUnset) rr = DGB_KNOB("DGB_SPLIT_RR", kKnobUnset) != 0; // experiment knob
    if (rr) {
      p.unitShift = 1;
      if (DGB_KNOB("DGB_SPLIT_RR_SHIFT", -1) >= 0) p.unitShift = (uint32_t)DGB_KNOB("DGB_SPLIT_RR_SHIFT", -1);
      p.roundRobin = 1;
      ctas = min(tiles << p.unitShift, full);
    } else {
      // several elements, each long enough to get CTAs of its own: round robin inside every element
      bool per = p.tileBegin == 0 && p.minElemTiles >= 2 * divUp(tiles, min(tiles, full)) && tiles >= full;
      per = per && DGB_KNOB("DGB_SPLIT_RR_ELEM", 1) != 0; // experiment knob
      if (per) {
        p.unitShift = 1; // at least four units per CTA, so that the CTAs of an element finish together
        while (p.unitShift < 3 && ((uint64_t)tiles << p.unitShift) < 4ull * full) ++p.unitShift;
        if (DGB_KNOB("DGB_SPLIT_RR_SHIFT", -1) >= 0) p.unitShift = (uint32_t)DGB_KNOB("DGB_SPLIT_RR_SHIFT", -1);
        p.roundRobin = 2;
        ctas = full;
      }
    }
  }
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
                     uint32_t* hist1, const uint32_t* checksum, int useChecksum, cudaStream_t s,
                     uint32_t* normTilesDone, uint2* normTable0, uint2* normTable1, int normProbBits,
                     uint32_t minElemTiles, bool foldChecksum) {
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
  p.checksumAcc = (foldChecksum && useChecksum && normTilesDone) ? (uint32_t*)checksum : nullptr;
  p.normTilesDone = normTilesDone;
  p.normTable0 = normTable0;
  p.normTable1 = normTable1;
  p.normProbBits = normProbBits;
  p.minElemTiles = minElemTiles;
  switch (ft) {
    case kFtF16: return launchSplitT<kFtF16>(p, s);
    case kFtBF16: return launchSplitT<kFtBF16>(p, s);
    case kFtF32: return launchSplitT<kFtF32>(p, s);
    case kFtF64: return launchSplitT<kFtF64>(p, s);
    default: return kErrInvalidArg;
  }
}

} // namespace dgb
