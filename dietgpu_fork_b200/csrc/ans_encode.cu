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
// Encode-table entry (ours): x = magic, y = (2^probBits - pdf) | shift << 12 | cdf << 17 with
// floor(v / pdf) == (umulhi(v, magic) + v) >> shift for v < 2^31. Storing K - pdf serves both the
// renormalisation test (v >= pdf << (31 - probBits)  <=>  bit 31 of v + ((K - pdf) << (31 - probBits)))
// and the state update v' = q * (K - pdf) + v + cdf; shift sits where a wrapping funnel shift can
// take it from y >> 12 without masking. Absent symbols (pdf 0) get an all-zero entry.
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
    uint32_t shift = 32 - __clz(q - 1); // q == 1 -> clz(0) == 32 -> 0
    uint64_t magic = ((1ull << 32) * ((1ull << shift) - q)) / q + 1; // :343-366
    entry.x = (uint32_t)magic;
    entry.y = (weight - q) | (shift << 12) | (cdf << 17);
  }
  table[e * 256 + s] = entry;
}

// ---------------------------------------------------------------------------------------------
// Block encoder + compaction (two kernels).
//
// k_encode: one autonomous warp per 4096-symbol ANS block, no CTA-wide barrier, no waiting on
// other warps.
//   symbols: 512-byte chunks (16 symbol rows) pulled into a 4-deep per-warp shared-memory ring by
//            cp.async.bulk (TMA 1-D) + mbarrier, issued 4 chunks ahead of their use
//   table:   the element's 2 KB encode table, read through L1
//   rows:    rANS step per 32 symbols, ballot/popc compaction into a 1 KB per-warp staging buffer
//   flush:   after every chunk the staged words go, as 16-byte vectors, to the block's slot of a
//            scratch area; final lane states go straight into the archive, the word count into
//            a per-block size array
// k_coalesce: per tile of 8 blocks, a warp-shuffle / decoupled look-back exclusive scan of the
// per-block sizes (roundUp(words, 8)) gives every block its final offset; the payload is copied
// from the scratch slots (L2-resident: just written) to its final place together with the block
// index, the pdf table and the header.
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
  uint32_t totalWarps;
  uint32_t blkBegin, blkEnd; // range of blocks handled by this k_encode launch
  uint32_t tileBegin;        // first 8-block tile handled by this k_coalesce launch
};

constexpr uint32_t kEncWarps = 4;
constexpr uint32_t kGroupBlocks = 256;          // compaction: blocks per size group
constexpr uint32_t kRingSlots = 4;
constexpr uint32_t kChunk = 512;                 // symbols per ring slot = 16 rows
constexpr uint32_t kEncStageBytes = 1024 + 32;   // 16 rows x 64 B + carry
constexpr uint32_t kSlotBytes = kStageWords * 2; // scratch bytes per block (worst case, probBits 11)
constexpr uint32_t kEncWarpSmem = kRingSlots * kChunk + kEncStageBytes + 32;

// One rANS step for 32 symbols (dietgpu/ans/GpuANSEncode.cuh:49-90). e = table entry
// {magic, (K - pdf) | shift << 12 | cdf << 17}; outPtr = shared-memory byte address of the next
// free staging word. Written in PTX to pin the instruction selection (~19 instructions).
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
      "mul.hi.u32 t, %0, %2;\n"
      "add.u32 t, t, %0;\n"
      "shr.u32 c, %3, 12;\n"
      "shf.r.wrap.b32 q, t, 0, c;\n"           // floor(state / pdf)
      "and.b32 k, %3, 0xfff;\n"
      "shr.u32 c, %3, 17;\n"
      "add.u32 c, c, %0;\n"                    // state + cdf
      "mad.lo.u32 %0, q, k, c;\n"              // q * (K - pdf) + state + cdf
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
  const uint32_t q = __funnelshift_r(__umulhi(state, e.x) + state, 0u, e.y >> 12);
  const uint32_t ns = q * (e.y & 0xfffu) + state + (e.y >> 17);
  state = valid ? ns : state;
}

template <int PB>
__global__ void __launch_bounds__(kEncWarps * 32, 10) k_encode(const EncParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const uint32_t lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  uint8_t* ws = smem + w * kEncWarpSmem;
  uint8_t* ring = ws;                        // kRingSlots * kChunk
  uint8_t* stage = ws + kRingSlots * kChunk; // kEncStageBytes
  const uint32_t bar0 = smemAddr(stage + kEncStageBytes); // kRingSlots mbarriers
  const uint32_t ringAddr = smemAddr(ring), stageAddr = smemAddr(stage);
  const uint32_t ltMask = laneMaskLt();
  const EncodeArgs& a = p.a;

  if (lane == 0) {
#pragma unroll
    for (uint32_t s = 0; s < kRingSlots; ++s) mbarInit(bar0 + 8 * s, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncwarp();
  uint32_t phaseBits = 0; // parity expected at the next wait on each slot
  uint32_t guess = 0;

  // blocks are independent here: static striding, so at any moment the resident warps stream one
  // contiguous window of the input (contiguous per-warp ranges measured 7 % slower on B200)
  uint32_t curElem = 0xffffffffu, n = 0, nb = 0;
  Elem el;
  uint8_t* ans = nullptr;
  for (uint32_t gb = p.blkBegin + blockIdx.x * kEncWarps + w; gb < p.blkEnd; gb += p.totalWarps) {
    if (curElem == 0xffffffffu || gb >= a.elems[curElem + 1].blockBase) {
      curElem = findElement(&a.elems[0].blockBase, sizeof(Elem) / 4, a.numElems, gb, guess);
      guess = curElem;
      el = a.elems[curElem];
      n = a.nDev ? a.nDev[curElem] : el.n;
      nb = divUp(n, kBlockSyms);
      ans = (uint8_t*)el.out;
      if (a.floatType != kFtNone) {
        const uint8_t* fbase = (const uint8_t*)el.out + el.sizeBias;
        ans = (uint8_t*)fbase + kFloatHeaderBytes + floatRawBytes(a.floatType, n);
        if (a.round) ans += *(const uint32_t*)(fbase + 16); // firstCompSegmentBytes
      }
    }
    const uint32_t e = curElem;
    const uint32_t blk = gb - el.blockBase;
    if (blk >= nb) { // empty element, or the device-side size is smaller than the host bound
      if (lane == 0) p.blockWords[gb] = 0;
      continue;
    }
    const uint32_t blkLen = min(kBlockSyms, n - blk * kBlockSyms);
    const uint8_t* src =
        (a.planeBase ? a.planeBase + el.planeOff : (const uint8_t*)el.in) + (size_t)blk * kBlockSyms;
    const uint32_t numChunks = divUp(blkLen, kChunk);
    const bool bulk = (((uintptr_t)src) & 15u) == 0;
    uint4* slot = (uint4*)(p.scratch + (size_t)gb * kSlotBytes);
    const uint2* __restrict__ tab = p.table + (size_t)e * 256;

    auto issueChunk = [&](uint32_t c) {
      const uint32_t sl = c & (kRingSlots - 1);
      const uint32_t len = min(kChunk, blkLen - c * kChunk);
      if (bulk) {
        if (lane == 0) {
          const uint32_t bytes = roundUp(len, 16);
          mbarExpectTx(bar0 + 8 * sl, bytes);
          bulkLoad(ringAddr + sl * kChunk, src + c * kChunk, bytes, bar0 + 8 * sl);
        }
      } else { // unaligned input: plain loads (4-byte words when possible)
        const uint8_t* g = src + c * kChunk + lane * 16;
        uint8_t* d = ring + sl * kChunk + lane * 16;
        if ((((uintptr_t)g) & 3u) == 0 && lane * 16 + 16 <= len) {
          const uint32_t* g4 = (const uint32_t*)g;
          *(uint4*)d = make_uint4(g4[0], g4[1], g4[2], g4[3]);
        } else {
#pragma unroll
          for (int k = 0; k < 16; ++k)
            if (lane * 16 + k < len) d[k] = g[k];
        }
      }
    };
    __syncwarp();
    for (uint32_t c = 0; c < min(numChunks, kRingSlots); ++c) issueChunk(c);

    uint32_t state = kStateMin, outPtr = stageAddr, flushedVecs = 0;
    for (uint32_t c = 0; c < numChunks; ++c) {
      const uint32_t sl = c & (kRingSlots - 1);
      if (bulk) {
        mbarWait(bar0 + 8 * sl, (phaseBits >> sl) & 1u);
        phaseBits ^= 1u << sl;
      } else {
        __syncwarp();
      }
      const uint32_t symAddr = ringAddr + sl * kChunk + lane;
      const uint32_t len = min(kChunk, blkLen - c * kChunk);
      if (len == kChunk) {
#pragma unroll
        for (int half = 0; half < 2; ++half) {
          uint2 ent[8];
#pragma unroll
          for (int r = 0; r < 8; ++r) {
            uint32_t sym;
            asm volatile("ld.shared.u8 %0, [%1];" : "=r"(sym) : "r"(symAddr + (half * 8 + r) * 32));
            ent[r] = __ldg(tab + sym);
          }
#pragma unroll
          for (int r = 0; r < 8; ++r) encStep<PB>(state, ent[r], outPtr, ltMask);
        }
      } else {
        const uint32_t rows = len >> 5, rem = len & 31u;
        for (uint32_t r = 0; r < rows; ++r) {
          uint32_t sym;
          asm volatile("ld.shared.u8 %0, [%1];" : "=r"(sym) : "r"(symAddr + r * 32));
          encStep<PB>(state, __ldg(tab + sym), outPtr, ltMask);
        }
        if (rem) {
          const bool valid = lane < rem;
          uint32_t sym = 0;
          if (valid) asm volatile("ld.shared.u8 %0, [%1];" : "=r"(sym) : "r"(symAddr + rows * 32));
          encStepPartial<PB>(state, __ldg(tab + sym), outPtr, ltMask, valid);
        }
      }
      __syncwarp();
      // this ring slot is consumed: refill it with the chunk kRingSlots ahead
      if (c + kRingSlots < numChunks) issueChunk(c + kRingSlots);
      // flush the staged words: whole 16-byte vectors go to the scratch slot, the rest is carried
      uint32_t bytes = outPtr - stageAddr;
      if (c + 1 == numChunks) { // last chunk: pad with zero words to a multiple of 8 words
        const uint32_t padded = roundUp(bytes, 16);
        if (bytes + 2 * lane < padded) *(uint16_t*)(stage + bytes + 2 * lane) = 0;
        __syncwarp();
        if (lane == 0) {
          p.blockWords[gb] = flushedVecs * 8 + (bytes >> 1);
          atomicAdd(&p.groupWords[el.groupBase + blk / kGroupBlocks], flushedVecs * 8 + (padded >> 1));
        }
        bytes = padded;
      }
      const uint32_t nv = bytes >> 4;
      for (uint32_t i = lane; i < nv; i += 32) slot[flushedVecs + i] = ((const uint4*)stage)[i];
      const uint32_t remBytes = bytes & 15u;
      uint32_t carry = 0;
      if (2 * lane < remBytes) carry = *(const uint16_t*)(stage + nv * 16 + 2 * lane);
      __syncwarp();
      if (2 * lane < remBytes) *(uint16_t*)(stage + 2 * lane) = (uint16_t)carry;
      __syncwarp();
      flushedVecs += nv;
      outPtr = stageAddr + remBytes;
    }
    // final lane states go straight to their place in the archive (:584-589)
    ((uint32_t*)(ans + kAnsFixedBytes))[blk * 32 + lane] = state;
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
  es.groupWords = tmp.alloc<uint32_t>(totalGroups);
  es.scratch = (uint8_t*)tmp.alloc((size_t)a.totalTiles * kSlotBytes);
  if (tmp.failed()) return kErrCuda;
  DGB_CUDA(cudaMemsetAsync(es.groupWords, 0, (size_t)totalGroups * 4, s));
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
  p.blkBegin = r.blk0;
  p.blkEnd = r.blk1;
  p.tileBegin = r.tile0;
  const uint32_t blocks = r.blk1 - r.blk0;
  const uint32_t grid = min(divUp(blocks, kEncWarps), (uint32_t)smCount() * 10u);
  p.totalWarps = grid * kEncWarps;
  const uint32_t smemBytes = kEncWarps * kEncWarpSmem;
  auto launch = [&](auto kern) -> int {
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
