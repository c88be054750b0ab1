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
#include <cstdlib>

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
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_normalize(const Elem* __restrict__ elems,
                                                   const uint32_t* __restrict__ nDev,
                                                   const uint32_t* __restrict__ hist, int probBits,
                                                   uint2* __restrict__ table, uint32_t elemBegin) {
  __shared__ uint32_t sKey[256];
  __shared__ uint32_t sWarp[8];
  const uint32_t e = elemBegin + blockIdx.x, s = threadIdx.x, lane = s & 31, w = s >> 5;
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
    const uint32_t l = 32 - __clz(q - 1); // q == 1 -> clz(0) == 32 -> 0
    uint32_t m = 0xffffffffu, shift = 0, cdfAdj = cdf + weight - 1;
    if (q > 1) {
      m = (uint32_t)(((1ull << (31 + l)) + q - 1) / q);
      shift = l - 1;
      cdfAdj = cdf;
    }
    entry = make_uint2(m, (weight - q) | (shift << 12) | (cdfAdj << 17));
  }
  table[e * 256 + s] = entry;
}

// ---------------------------------------------------------------------------------------------
// Block encoder + compaction (two kernels).
//
// k_encode: one autonomous warp per 4096-symbol ANS block, no CTA-wide barrier, no waiting on
// other warps, no shared-memory input staging.
//   symbols: one byte load per lane and symbol row straight from global memory (a row is one
//            32-byte sector, four rows share an L1 line), issued a whole 8-row batch ahead of use
//   table:   the element's 4 KB encode table, read through L1, four rows ahead of use
//   rows:    rANS step per 32 symbols (13 instructions), ballot/popc compaction into a 2.5 KB
//            per-warp staging buffer
//   flush:   the staged words go, as 16-byte vectors, to the block's slot of a scratch area when
//            the block ends (typical data: once per block) or when the buffer could overflow
//            within the next batch; final lane states go straight into the archive, the word
//            count into a per-block size array
// k_coalesce: every size is known when it starts; per tile of 8 blocks it sums the sizes before
// the tile and copies the payload from the scratch slots (L2-resident: just written) to its final
// place together with the block index, the pdf table and the header.
//
// Why not a single pass: a block's final offset is known only when ALL earlier blocks of the
// element are encoded, i.e. after the slowest of them; measured on B200, warps spent as long
// waiting on that look-back as encoding (profiles/r1_k_encode_v2), and holding a worst-case 5 KB
// block in shared memory while waiting capped occupancy at 24 warps/SM.
// ---------------------------------------------------------------------------------------------
struct EncParams {
  EncodeArgs a;
  const uint2* table;   // [numElems][256]
  uint8_t* scratch;     // [totalBlocks][kSlotBytes]
  uint32_t* blockWords; // [totalBlocks] words emitted per block
  uint32_t* groupWords; // [totalGroups] sum of roundUp(words, 8) per group of 256 blocks, zeroed
  uint32_t* runCounter;  // zeroed; hands out runs beyond each warp's first one
  uint32_t totalWarps;
  uint32_t tileBegin, tileEnd; // range of 8-block tiles handled by this launch
};

constexpr uint32_t kEncWarps = 4;
constexpr uint32_t kEncCtasPerSm = 10;
constexpr uint32_t kGroupBlocks = 256;           // compaction: blocks per size group
constexpr uint32_t kEncBatchRows = 8;
constexpr uint32_t kRunBlocks = 4;              // consecutive blocks one warp encodes as one stream (half a tile)
constexpr uint32_t kBlockBatches = kBlockSyms / (32 * kEncBatchRows);
constexpr uint32_t kEncTableBytes = 256 * 8;     // per-warp copy of the element's encode table
constexpr uint32_t kEncStageBytes = 2560;
constexpr uint32_t kEncFlushAt = kEncStageBytes - kEncBatchRows * 64; // room for one more batch, worst case
constexpr uint32_t kSlotBytes = kStageWords * 2; // scratch bytes per block (worst case, probBits 11)
constexpr uint32_t kEncWarpSmem = kEncTableBytes + kEncStageBytes;

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

__device__ __forceinline__ uint32_t ldSym(const uint8_t* p) {
  uint32_t v;
  asm volatile("ld.global.nc.L2::256B.u8 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}

// table entry of symbol `sym` from the warp's shared-memory copy
__device__ __forceinline__ uint2 ldEnt(uint32_t tabAddr, uint32_t sym) {
  uint2 e;
  asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(e.x), "=r"(e.y) : "r"(tabAddr + sym * 8));
  return e;
}

template <int PB>
__global__ void __launch_bounds__(kEncWarps * 32, kEncCtasPerSm) k_encode(const EncParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint8_t* tabSmem = smem + w * kEncWarpSmem;
  uint8_t* stage = tabSmem + kEncTableBytes;
  const uint32_t tabAddr = smemAddr(tabSmem), stageAddr = smemAddr(stage);
  const uint32_t ltMask = laneMaskLt();
  const EncodeArgs& a = p.a;
  uint32_t guess = 0;

  // Work unit = run of kRunBlocks consecutive blocks of one element (half of an 8-block tile): one
  // element lookup and one uninterrupted symbol stream per run. A warp's first run is static, further
  // runs are handed out in order by an atomic counter (fetched one run ahead), so the resident warps
  // always stream one contiguous window of the input and finish together.
  uint32_t curElem = 0xffffffffu, n = 0, nb = 0, blockBase = 0, tileBase = 0, nextTileBase = 0;
  uint32_t* groupSum = nullptr;
  const uint8_t* elemSrc = nullptr;
  uint8_t* ans = nullptr;
  const uint32_t numRuns = (p.tileEnd - p.tileBegin) * (kTileBlocks / kRunBlocks);
  uint32_t run = blockIdx.x * kEncWarps + w;
  while (run < numRuns) {
    uint32_t nextRun = 0;
    if (lane == 0) nextRun = p.totalWarps + atomicAdd(p.runCounter, 1u);
    const uint32_t tile = p.tileBegin + run / (kTileBlocks / kRunBlocks);
    if (curElem == 0xffffffffu || tile >= nextTileBase || tile < tileBase) {
      curElem = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tile, guess);
      guess = curElem;
      const Elem el = a.elems[curElem];
      nextTileBase = a.elems[curElem + 1].tileBase;
      tileBase = el.tileBase;
      blockBase = el.blockBase;
      groupSum = p.groupWords + el.groupBase;
      elemSrc = (a.planeBase ? a.planeBase + el.planeOff : (const uint8_t*)el.in) + lane;
      { // the element's 2 KB table -> this warp's shared-memory copy
        const uint4* g = (const uint4*)(p.table + (size_t)curElem * 256);
        __syncwarp();
#pragma unroll
        for (int i = 0; i < 4; ++i) ((uint4*)tabSmem)[lane + 32 * i] = g[lane + 32 * i];
        __syncwarp();
      }
      n = a.nDev ? a.nDev[curElem] : el.n;
      nb = divUp(n, kBlockSyms);
      ans = (uint8_t*)el.out;
      if (a.floatType != kFtNone) {
        const uint8_t* fbase = (const uint8_t*)el.out + el.sizeBias;
        ans = (uint8_t*)fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
        if (a.round) ans += *(const uint32_t*)(fbase + 16); // firstCompSegmentBytes
      }
    }
    uint32_t blk = (tile - tileBase) * kTileBlocks + (run % (kTileBlocks / kRunBlocks)) * kRunBlocks;
    if (blk < nb) { // else: empty element, or the device-side size is smaller than the host bound
      const uint32_t runSyms = min(n, (blk + kRunBlocks) * kBlockSyms) - blk * kBlockSyms;
      const uint8_t* src = elemSrc + (size_t)blk * kBlockSyms;

      uint32_t state = kStateMin, outPtr = stageAddr, flushedVecs = 0;
      // staged words -> scratch slot of block `blk`. final: pad with zero words to a multiple of 8 words,
      // publish the sizes and the lane states, reset for the next block; otherwise whole 16-byte vectors
      // go out and the remainder moves to the front of the buffer
      auto flush = [&](bool final) {
        const uint32_t gb = blockBase + blk;
        uint4* slot = (uint4*)(p.scratch + (size_t)gb * kSlotBytes);
        uint32_t bytes = outPtr - stageAddr;
        __syncwarp();
        if (final) {
          const uint32_t padded = roundUp(bytes, 16);
          if (bytes + 2 * lane < padded) *(uint16_t*)(stage + bytes + 2 * lane) = 0;
          __syncwarp();
          if (lane == 0) {
            p.blockWords[gb] = flushedVecs * 8 + (bytes >> 1);
            atomicAdd(&groupSum[blk / kGroupBlocks], flushedVecs * 8 + (padded >> 1));
          }
          bytes = padded;
        }
        const uint32_t nv = bytes >> 4;
        for (uint32_t i = lane; i < nv; i += 32) slot[flushedVecs + i] = ((const uint4*)stage)[i];
        if (final) {
          // final lane states go straight to their place in the archive (:584-589)
          ((uint32_t*)(ans + kAnsFixedBytes))[blk * 32 + lane] = state;
          __syncwarp(); // the staging buffer is reused by the next block
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

      const uint32_t rows = runSyms >> 5, rem = runSyms & 31u;
      const uint32_t batches = rows / kEncBatchRows;
      if (batches) {
        // software pipeline over rows: at row r the step consumes ent[r % 2]; that register then takes the
        // table entry of row r + 2, whose symbol sits in sym[(r + 2) % 8]; that register in turn takes the
        // symbol of row r + 10. Symbols are thus requested 8 rows (>= 150 issue slots x resident warps)
        // before the dependent table load, table entries 2 rows before their step. The stream runs
        // through block boundaries; only the state and the staging buffer are reset there.
        uint32_t sym[8];
        uint2 ent[2];
        ent[0] = ldEnt(tabAddr, ldSym(src));
        ent[1] = ldEnt(tabAddr, ldSym(src + 32));
#pragma unroll
        for (int r = 2; r < 8; ++r) sym[r] = ldSym(src + r * 32);
        {
          const uint8_t* ld = batches > 1 ? src + 8 * 32 : src;
          sym[0] = ldSym(ld);
          sym[1] = ldSym(ld + 32);
        }
        const uint8_t* nx = src + 10 * 32; // row r + 10 for r = first row of the batch
        for (uint32_t b = 0; b < batches; ++b, nx += kEncBatchRows * 32) {
          // rows past the end of the run are replaced by re-reads of its first rows (valid memory, the
          // values are never used): the loads stay unconditional, so nothing waits on them here
          const uint8_t* ld1 = b + 1 < batches ? nx : src;
          const uint8_t* ld2 = b + 2 < batches ? nx : src;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            encStep<PB>(state, ent[i & 1], outPtr, ltMask);
            ent[i & 1] = ldEnt(tabAddr, sym[(i + 2) & 7]);
            sym[(i + 2) & 7] = ldSym((i < 6 ? ld1 : ld2) + i * 32);
          }
          if ((b & (kBlockBatches - 1)) == kBlockBatches - 1) flush(true);
          else if (outPtr - stageAddr >= kEncFlushAt) flush(false);
        }
      }
      // last block of an element: leftover rows and the partial row
      if (rows % kEncBatchRows || rem) {
        if (outPtr - stageAddr >= kEncFlushAt) flush(false);
        const uint8_t* tail = src + (size_t)batches * (kEncBatchRows * 32);
        for (uint32_t r = 0; r < rows % kEncBatchRows; ++r)
          encStep<PB>(state, ldEnt(tabAddr, ldSym(tail + r * 32)), outPtr, ltMask);
        if (rem) {
          const bool valid = lane < rem;
          const uint32_t sym = valid ? ldSym(tail + (rows % kEncBatchRows) * 32) : 0u;
          encStepPartial<PB>(state, ldEnt(tabAddr, sym), outPtr, ltMask, valid);
        }
      }
      if (runSyms & (kBlockSyms - 1)) flush(true); // the partial block
    }
    run = __shfl_sync(0xffffffffu, nextRun, 0);
  }
}

// Compaction: tile = 8 consecutive blocks of one element, one warp per block, one CTA per tile.
// Every size is known when this kernel starts, so a tile's exclusive prefix is a plain sum: the
// per-group totals (groups of kGroupBlocks blocks, accumulated by k_encode with one atomicAdd per
// block) of the groups before it, plus the blocks of its own group before it (<= 2 * 256 loads).
__global__ void __launch_bounds__(256) k_coalesce(const EncParams p, int probBits) {
  __shared__ uint32_t sWarp[8];
  __shared__ uint32_t sWords[kTileBlocks];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const EncodeArgs& a = p.a;
  const uint32_t tk = p.tileBegin + blockIdx.x;
  const uint32_t e = findElement(&a.elems[0].tileBase, sizeof(Elem) / 4, a.numElems, tk, 0);
  const Elem el = a.elems[e];
  const uint32_t t = tk - el.tileBase;
  const uint32_t n = a.nDev ? a.nDev[e] : el.n;
  const uint32_t nb = divUp(n, kBlockSyms);
  const uint32_t numTiles = max(1u, divUp(nb, kTileBlocks));
  if (t >= numTiles) return; // device-side size is smaller than the host bound
  uint8_t* ans = (uint8_t*)el.out;
  uint8_t* fbase = nullptr;
  if (a.floatType != kFtNone) {
    fbase = (uint8_t*)el.out + el.sizeBias;
    ans = fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
    if (a.round) ans += *(const uint32_t*)(fbase + 16);
  }
  // ---- exclusive prefix of this tile ----
  const uint32_t firstBlk = t * kTileBlocks;
  const uint32_t grp = firstBlk / kGroupBlocks;
  const uint32_t* groupSum = p.groupWords + el.groupBase;
  uint32_t part = 0;
  for (uint32_t g = tid; g < grp; g += 256) part += groupSum[g];
  for (uint32_t b = grp * kGroupBlocks + tid; b < firstBlk; b += 256)
    part += roundUp(p.blockWords[el.blockBase + b], 8);
  part = warpSum(part);
  if (lane == 0) sWarp[w] = part;
  const uint32_t blk = firstBlk + w;
  const uint32_t gb = el.blockBase + blk;
  uint32_t blkLen = 0, words = 0;
  if (blk < nb) {
    blkLen = min(kBlockSyms, n - blk * kBlockSyms);
    words = p.blockWords[gb];
  }
  if (lane == 0) sWords[w] = roundUp(words, 8);
  __syncthreads();
  uint32_t excl = 0, myOff = 0, tileWords = 0;
#pragma unroll
  for (uint32_t j = 0; j < 8; ++j) {
    excl += sWarp[j];
    uint32_t x = sWords[j];
    myOff += j < w ? x : 0u;
    tileWords += x;
  }
  if (blkLen) { // index entry and payload (dietgpu/ans/GpuANSEncode.cuh:591-623)
    const uint32_t startWord = excl + myOff;
    if (lane == 0) {
      uint2* bw = (uint2*)(ans + kAnsFixedBytes + 128u * nb);
      bw[blk] = make_uint2((blkLen << 16) | words, startWord);
    }
    uint4* dst = (uint4*)(ans + ansOverhead(nb) + 2 * (size_t)startWord);
    const uint4* src = (const uint4*)(p.scratch + (size_t)gb * kSlotBytes);
    const uint32_t nv = roundUp(words, 8) >> 3;
    for (uint32_t i = lane; i < nv; i += 32) dst[i] = src[i];
  }
  if (t == 0) { // pdf table (:568-573); zero for absent symbols and for an empty element
    const uint2 en = p.table[(size_t)e * 256 + tid];
    const uint32_t pdf = (n && en.x) ? (1u << probBits) - (en.y & 0xfffu) : 0u;
    ((uint16_t*)(ans + kAnsHeaderBytes))[tid] = (uint16_t)pdf;
  }
  if (t == numTiles - 1 && tid == 0) {
    const uint32_t totalWords = excl + tileWords;
    uint32_t cs = (a.useChecksum && a.checksum) ? a.checksum[e] : 0u;
    uint4* h = (uint4*)ans; // :533-565; fields the reference leaves undefined are written as 0
    h[0] = make_uint4(kAnsMagic, nb, n, totalWords);
    h[1] = make_uint4((uint32_t)probBits | (a.useChecksum ? 16u : 0u), cs, 0u, 0u);
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

int allocEncodeScratch(Temp& tmp, const EncodeArgs& a, uint32_t totalGroups, EncodeScratch& es, cudaStream_t s) {
  es.table = tmp.alloc<uint2>((size_t)a.numElems * 256);
  es.blockWords = tmp.alloc<uint32_t>(a.totalTiles);
  es.groupWords = tmp.alloc<uint32_t>(totalGroups + a.numElems + 1); // + one run counter per possible launch
  es.scratch = (uint8_t*)tmp.alloc((size_t)a.totalTiles * kSlotBytes);
  if (tmp.failed()) return kErrCuda;
  es.runCounters = es.groupWords + totalGroups;
  DGB_CUDA(cudaMemsetAsync(es.groupWords, 0, (size_t)(totalGroups + a.numElems + 1) * 4, s));
  return kOk;
}

int launchAnsEncode(const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r, cudaStream_t s) {
  if (r.e1 <= r.e0) return kOk;
  if (a.probBits < 9 || a.probBits > 11) return kErrInvalidArg;
  {
    ProfScope ps__("k_normalize", s);
    k_normalize<<<r.e1 - r.e0, 256, 0, s>>>(a.elems, a.nDev, a.hist, a.probBits, es.table, r.e0);
  }
  DGB_CUDA(cudaGetLastError());

  EncParams p;
  p.a = a;
  p.table = es.table;
  p.scratch = es.scratch;
  p.blockWords = es.blockWords;
  p.groupWords = es.groupWords;
  p.runCounter = es.runCounters + r.e0;
  p.tileBegin = r.tile0;
  p.tileEnd = r.tile1;
  const uint32_t runs = (r.tile1 - r.tile0) * (kTileBlocks / kRunBlocks);
  uint32_t perSm = kEncCtasPerSm;
  if (const char* e = getenv("DGB_ENC_CTAS")) perSm = (uint32_t)atoi(e); // experiment knob
  const uint32_t grid = min(divUp(runs, kEncWarps), (uint32_t)smCount() * perSm);
  p.totalWarps = grid * kEncWarps;
  const uint32_t smemBytes = kEncWarps * kEncWarpSmem;
  auto launch = [&](auto kern) -> int {
    DGB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemBytes));
    {
      ProfScope ps__("k_encode", s);
      kern<<<grid, kEncWarps * 32, smemBytes, s>>>(p);
    }
    DGB_CUDA(cudaGetLastError());
    return kOk;
  };
  int rc;
  switch (a.probBits) {
    case 9: rc = launch(k_encode<9>); break;
    case 10: rc = launch(k_encode<10>); break;
    default: rc = launch(k_encode<11>); break;
  }
  if (rc) return rc;
  {
    ProfScope ps__("k_coalesce", s);
    k_coalesce<<<r.tile1 - r.tile0, 256, 0, s>>>(p, a.probBits);
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

} // namespace dgb
