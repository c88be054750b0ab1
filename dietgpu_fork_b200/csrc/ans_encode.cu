// ANS encode side for sm_100a: byte histogram, probability normalisation, and the tile encoder
// with single-pass (decoupled look-back) output compaction.
//
// What is computed follows the reference (citations relative to /root/reference/):
//   histogram            dietgpu/ans/GpuANSStatistics.cuh:21-143
//   normalisation        dietgpu/ans/GpuANSStatistics.cuh:178-367
//   per-warp rANS coder  dietgpu/ans/GpuANSEncode.cuh:49-211
//   archive assembly     dietgpu/ans/GpuANSEncode.cuh:511-624
// How it is computed does not: the reference encodes into an uncoalesced temp buffer, runs a
// 1-3 kernel prefix sum and copies every compressed byte a second time. Here one persistent
// kernel stages a tile of 8 blocks in shared memory, encodes it with 8 warps, obtains the tile's
// output offset with a decoupled look-back over tile aggregates and writes states, block index
// and payload straight to their final place; the last tile of an element writes the header.
#include "internal.h"
#include "prof.h"

namespace dgb {

int smCount() {
  static int cached = 0;
  if (!cached) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&cached, cudaDevAttrMultiProcessorCount, dev);
    if (cached <= 0) cached = 148;
  }
  return cached;
}

// ---------------------------------------------------------------------------------------------
// Byte histogram. Each CTA owns a contiguous range of tiles, accumulates into per-warp,
// 4-way replicated shared-memory bins (lanes l and l+4 share a replica) and flushes to global
// memory whenever the element changes.
// ---------------------------------------------------------------------------------------------
constexpr int kHistRep = 4;

__device__ __forceinline__ void histAdd(uint32_t* wh, uint32_t sym, uint32_t rep) {
  atomicAdd(&wh[sym * kHistRep + rep], 1u);
}

__device__ __forceinline__ void histAddWord(uint32_t* wh, uint32_t w, uint32_t rep) {
  histAdd(wh, w & 0xffu, rep);
  histAdd(wh, (w >> 8) & 0xffu, rep);
  histAdd(wh, (w >> 16) & 0xffu, rep);
  histAdd(wh, w >> 24, rep);
}

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

__global__ void __launch_bounds__(256) k_histBytes(const Elem* __restrict__ elems, uint32_t numElems,
                                                   uint32_t totalTiles, uint32_t tilesPerCta,
                                                   uint32_t* __restrict__ hist) {
  __shared__ uint32_t sh[8 * 256 * kHistRep];
  const uint32_t tid = threadIdx.x;
  for (uint32_t i = tid; i < 8 * 256 * kHistRep; i += 256) sh[i] = 0;
  __syncthreads();
  uint32_t* wh = sh + (tid >> 5) * 256 * kHistRep;
  const uint32_t rep = tid & (kHistRep - 1);

  uint32_t t0 = blockIdx.x * tilesPerCta;
  uint32_t t1 = min(t0 + tilesPerCta, totalTiles);
  uint32_t cur = 0xffffffffu;
  for (uint32_t tile = t0; tile < t1; ++tile) {
    uint32_t e = findElement(&elems[0].tileBase, sizeof(Elem) / 4, numElems, tile, cur);
    if (e != cur) {
      if (cur != 0xffffffffu) histFlush(sh, hist + cur * 256, tid);
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
    if (tid < headN) histAdd(wh, p[tid], rep);
    if (tid >= 32 && tid - 32 < tailN) histAdd(wh, a1[tid - 32], rep);
    const uint4* v = (const uint4*)a0;
    uint32_t nvec = (uint32_t)(a1 - a0) >> 4;
    uint32_t i = tid;
    for (; i + 768 < nvec; i += 1024) {
      uint4 x0 = v[i], x1 = v[i + 256], x2 = v[i + 512], x3 = v[i + 768];
      histAddWord(wh, x0.x, rep); histAddWord(wh, x0.y, rep); histAddWord(wh, x0.z, rep); histAddWord(wh, x0.w, rep);
      histAddWord(wh, x1.x, rep); histAddWord(wh, x1.y, rep); histAddWord(wh, x1.z, rep); histAddWord(wh, x1.w, rep);
      histAddWord(wh, x2.x, rep); histAddWord(wh, x2.y, rep); histAddWord(wh, x2.z, rep); histAddWord(wh, x2.w, rep);
      histAddWord(wh, x3.x, rep); histAddWord(wh, x3.y, rep); histAddWord(wh, x3.z, rep); histAddWord(wh, x3.w, rep);
    }
    for (; i < nvec; i += 256) {
      uint4 x0 = v[i];
      histAddWord(wh, x0.x, rep); histAddWord(wh, x0.y, rep); histAddWord(wh, x0.z, rep); histAddWord(wh, x0.w, rep);
    }
  }
  if (cur != 0xffffffffu) histFlush(sh, hist + cur * 256, tid);
}

int launchHistogram(const Elem* elems, uint32_t numElems, uint32_t totalTiles, uint32_t* hist,
                    cudaStream_t s) {
  if (!totalTiles) return kOk;
  uint32_t ctas = min(totalTiles, (uint32_t)smCount() * 4u);
  uint32_t per = divUp(totalTiles, ctas);
  ctas = divUp(totalTiles, per);
  {
    ProfScope ps__("k_histBytes", s);
    k_histBytes<<<ctas, 256, 0, s>>>(elems, numElems, totalTiles, per, hist);
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

// ---------------------------------------------------------------------------------------------
// Normalisation: counts -> pdf summing to 2^probBits, bit-exact with the reference
// (dietgpu/ans/GpuANSStatistics.cuh:178-367). One 256-thread CTA per element; thread s owns
// symbol s throughout. Instead of physically sorting we compute each symbol's rank in the
// descending order of (q << 16 | s): the reference's "positions" are these ranks.
//
// Encode-table entry (ours): x = magic, y = pdf | cdf << 12 | shift << 24 with
// floor(v / pdf) == (umulhi(v, magic) + v) >> shift for v < 2^31.
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_normalize(const Elem* __restrict__ elems,
                                                   const uint32_t* __restrict__ nDev,
                                                   const uint32_t* __restrict__ hist, int probBits,
                                                   uint2* __restrict__ table) {
  __shared__ uint32_t sKey[256];
  __shared__ uint32_t sWarp[8];
  const uint32_t e = blockIdx.x, s = threadIdx.x, lane = s & 31, w = s >> 5;
  const uint32_t total = nDev ? nDev[e] : elems[e].n;
  if (total == 0) { // reference leaves the table untouched (:192-195); we define it as zero
    table[e * 256 + s] = make_uint2(0, 0);
    return;
  }
  const uint32_t count = hist[e * 256 + s];
  const uint32_t weight = 1u << probBits;
  // :205-221 — fp32 divide (RN) then multiply (RN), truncation
  float ratio = __fdiv_rn((float)count, (float)total);
  uint32_t q = (uint32_t)__fmul_rn((float)weight, ratio);
  if (count > 0 && q == 0) q = 1;

  uint32_t ws = warpSum(q);
  if (lane == 0) sWarp[w] = ws;
  const uint32_t key = (q << 16) | s;
  sKey[s] = key;
  __syncthreads();
  int sum = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) sum += (int)sWarp[i];
  // rank in descending key order == position after the reference's SortDescending (:229-252)
  uint32_t pos = 0;
#pragma unroll 16
  for (int j = 0; j < 256; ++j) pos += (sKey[j] > key);

  int diff = (int)weight - sum;
  if (diff > 0) {
    // :258-273 — the test is on the symbol id, not on the rank
    while (diff > 0) {
      int k = diff < 256 ? diff : 256;
      if ((int)s < k) q += 1;
      diff -= k;
    }
  } else if (diff < 0) {
    // :274-315 — take one from the smallest entries that are still > 1, by position
    diff = -diff;
    while (diff > 0) {
      int g = __syncthreads_count(q > 1);
      int k = diff < g ? diff : g;
      if (k <= 0) break;
      if ((int)pos >= g - k && (int)pos < g) q -= 1;
      diff -= k;
    }
  }
  // exclusive scan over symbols -> cdf (:317-341)
  uint32_t inc = q;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= (uint32_t)o) inc += t;
  }
  __syncthreads(); // sWarp reuse
  if (lane == 31) sWarp[w] = inc;
  __syncthreads();
  uint32_t cdf = inc - q;
  for (uint32_t i = 0; i < w; ++i) cdf += sWarp[i];

  uint2 entry = make_uint2(0, 0);
  if (q) {
    uint32_t shift = 32 - __clz(q - 1); // q == 1 -> clz(0) == 32 -> 0
    uint64_t magic = ((1ull << 32) * ((1ull << shift) - q)) / q + 1; // :343-366
    entry.x = (uint32_t)magic;
    entry.y = q | (cdf << 12) | (shift << 24);
  }
  table[e * 256 + s] = entry;
}

// ---------------------------------------------------------------------------------------------
// Tile encoder
// ---------------------------------------------------------------------------------------------
struct EncParams {
  EncodeArgs a;
  const uint2* table;           // [numElems][256]
  unsigned long long* tileState; // [totalTiles], zeroed
  uint32_t* ticket;             // zeroed
  uint32_t* errFlag;
};

constexpr uint32_t kEncSmem = kTileSyms + 16 + kTileBlocks * kStageWords * 2 + 256 * 8;

template <int PB>
__device__ __forceinline__ void encRow(uint32_t& state, uint32_t sym, const uint2* __restrict__ tab,
                                       uint16_t* __restrict__ stage, uint32_t& off, uint32_t ltMask,
                                       bool valid) {
  uint2 e = tab[sym];
  uint32_t pdf = e.y & 0xfffu;
  uint32_t cdf = (e.y >> 12) & 0xfffu;
  uint32_t shift = e.y >> 24;
  bool wr = valid && state >= (pdf << (31 - PB));
  uint32_t vote = __ballot_sync(0xffffffffu, wr);
  if (wr) {
    stage[off + __popc(vote & ltMask)] = (uint16_t)state;
    state >>= 16;
  }
  off += __popc(vote);
  uint32_t q = (__umulhi(state, e.x) + state) >> shift;
  uint32_t ns = q * ((1u << PB) - pdf) + state + cdf; // (q << PB) + state % pdf + cdf
  state = valid ? ns : state;
}

template <int PB>
__global__ void __launch_bounds__(256, 2) k_encode(const EncParams p) {
  extern __shared__ __align__(16) uint8_t smem[];
  uint8_t* sSym = smem;
  uint16_t* sStage = (uint16_t*)(smem + kTileSyms + 16);
  uint2* sTab = (uint2*)(smem + kTileSyms + 16 + kTileBlocks * kStageWords * 2);
  __shared__ uint32_t sTicket, sElem, sExcl;
  __shared__ uint32_t sWords[kTileBlocks];

  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t ltMask = laneMaskLt();
  const EncodeArgs& a = p.a;
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
    const uint32_t n = a.nDev ? a.nDev[e] : el.n;
    const uint32_t nb = divUp(n, kBlockSyms);
    const uint32_t numTiles = max(1u, divUp(nb, kTileBlocks));
    if (t >= numTiles) { // device-side size is smaller than the host bound
      __syncthreads();
      continue;
    }

    // where this element's ANS archive lives
    uint8_t* ans = (uint8_t*)el.out;
    uint8_t* fbase = nullptr; // float archive (behind the sparse header + bitmap when nested)
    if (a.floatType != kFtNone) {
      fbase = (uint8_t*)el.out + el.sizeBias;
      ans = fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
      if (a.round) ans += *(const uint32_t*)(fbase + 16); // firstCompSegmentBytes
    }

    sTab[tid] = p.table[e * 256 + tid];

    // ---- stage this tile's symbols in shared memory (any input alignment) ----
    const uint8_t* elemStart = a.planeBase ? a.planeBase + el.planeOff : (const uint8_t*)el.in;
    const uint8_t* elemEnd = elemStart + n;
    const uint32_t symOff = t * kTileSyms;
    const uint32_t len = n > symOff ? min(kTileSyms, n - symOff) : 0;
    const uint8_t* g = elemStart + symOff;
    const uint32_t mis = (uint32_t)((uintptr_t)g & 15u);
    {
      const uint8_t* gA = g - mis;
      const uint32_t nvec = len ? divUp(mis + len, 16) : 0;
      for (uint32_t v = tid; v < nvec; v += kThreads) {
        const uint8_t* src = gA + 16 * (size_t)v;
        uint4 val;
        if (src >= elemStart && src + 16 <= elemEnd) {
          val = *(const uint4*)src;
        } else {
          uint32_t b[4] = {0, 0, 0, 0};
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const uint8_t* c = src + k;
            if (c >= elemStart && c < elemEnd) b[k >> 2] |= (uint32_t)(*c) << (8 * (k & 3));
          }
          val = make_uint4(b[0], b[1], b[2], b[3]);
        }
        ((uint4*)sSym)[v] = val;
      }
    }
    __syncthreads();

    // ---- encode: warp w owns block t*8 + w ----
    const uint32_t blk = t * kTileBlocks + w;
    uint32_t blkLen = 0;
    if (blk < nb) blkLen = min(kBlockSyms, n - blk * kBlockSyms);
    uint16_t* stage = sStage + w * kStageWords;
    uint32_t state = kStateMin, words = 0;
    if (blkLen) {
      const uint8_t* sym = sSym + mis + w * kBlockSyms + lane;
      const uint32_t rows = blkLen >> 5;
#pragma unroll 8
      for (uint32_t r = 0; r < rows; ++r) encRow<PB>(state, sym[r * 32], sTab, stage, words, ltMask, true);
      const uint32_t rem = blkLen & 31u;
      if (rem) {
        bool valid = lane < rem;
        uint32_t sv = valid ? sym[rows * 32] : 0u;
        encRow<PB>(state, sv, sTab, stage, words, ltMask, valid);
      }
      // zero the padding up to a multiple of 8 words (the reference leaves it undefined)
      uint32_t padded = roundUp(words, 8);
      if (words + lane < padded) stage[words + lane] = 0;
    }
    if (lane == 0) sWords[w] = roundUp(words, 8);
    __syncthreads();

    uint32_t myOff = 0, tileWords = 0;
#pragma unroll
    for (uint32_t j = 0; j < kTileBlocks; ++j) {
      uint32_t x = sWords[j];
      myOff += j < w ? x : 0u;
      tileWords += x;
    }
    if (w == 0) {
      uint32_t ex = lookbackExclusive(p.tileState + el.tileBase, t, tileWords, p.errFlag);
      if (lane == 0) sExcl = ex;
    }
    __syncthreads();
    const uint32_t excl = sExcl;

    // ---- write this block: states, index entry, payload (dietgpu/ans/GpuANSEncode.cuh:584-623) ----
    if (blkLen) {
      const uint32_t startWord = excl + myOff;
      ((uint32_t*)(ans + kAnsFixedBytes))[blk * 32 + lane] = state;
      if (lane == 0) {
        uint2* bw = (uint2*)(ans + kAnsFixedBytes + 128u * nb);
        bw[blk] = make_uint2((blkLen << 16) | words, startWord);
      }
      __syncwarp();
      uint4* dst = (uint4*)(ans + ansOverhead(nb) + 2 * (size_t)startWord);
      const uint4* src = (const uint4*)stage;
      const uint32_t nv = roundUp(words, 8) >> 3;
      for (uint32_t i = lane; i < nv; i += 32) dst[i] = src[i];
    }
    if (t == 0) { // pdf table (:568-573); zero for an empty element
      uint32_t pdf = n ? (sTab[tid].y & 0xfffu) : 0u;
      ((uint16_t*)(ans + kAnsHeaderBytes))[tid] = (uint16_t)pdf;
    }
    if (t == numTiles - 1 && tid == 0) {
      const uint32_t totalWords = excl + tileWords;
      uint32_t cs = (a.useChecksum && a.checksum) ? a.checksum[e] : 0u;
      uint4* h = (uint4*)ans; // :533-565; fields the reference leaves undefined are written as 0
      h[0] = make_uint4(kAnsMagic, nb, n, totalWords);
      h[1] = make_uint4((uint32_t)PB | (a.useChecksum ? 16u : 0u), cs, 0u, 0u);
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
}

int launchAnsEncode(Temp& tmp, const EncodeArgs& a, cudaStream_t s) {
  if (a.numElems == 0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  uint2* table = tmp.alloc<uint2>((size_t)a.numElems * 256);
  // tile states + ticket + error flag share one zeroed region
  size_t zeroBytes = (size_t)a.totalTiles * 8 + 16;
  uint8_t* zero = (uint8_t*)tmp.alloc(zeroBytes);
  if (tmp.failed()) return kErrCuda;
  DGB_CUDA(cudaMemsetAsync(zero, 0, zeroBytes, s));
  {
    ProfScope ps__("k_normalize", s);
    k_normalize<<<a.numElems, 256, 0, s>>>(a.elems, a.nDev, a.hist, a.probBits, table);
  }
  DGB_CUDA(cudaGetLastError());

  EncParams p;
  p.a = a;
  p.table = table;
  p.tileState = (unsigned long long*)zero;
  p.ticket = (uint32_t*)(zero + (size_t)a.totalTiles * 8);
  p.errFlag = p.ticket + 1;
  uint32_t grid = min(a.totalTiles, (uint32_t)smCount() * 2u);
  auto launch = [&](auto kern) -> int {
    DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kEncSmem));
    {
      ProfScope ps__("k_encode", s);
      kern<<<grid, kThreads, kEncSmem, s>>>(p);
    }
    DGB_CUDA(cudaGetLastError());
    return kOk;
  };
  switch (a.probBits) {
    case 9: return launch(k_encode<9>);
    case 10: return launch(k_encode<10>);
    default: return launch(k_encode<11>);
  }
}

} // namespace dgb
