// Sparse-float bitmap build / compaction and its inverse for sm_100a.
//
// Archive (dietgpu/float/GpuFloatUtils.cuh:98-128, GpuSparseFloatCompress.cuh:29-247):
//   [0,16)  u32 size = N (+12 bytes the reference leaves undefined, 0 here)
//   [16, 16 + roundUp(ceil(N/8),16))  bitmap, bit (7 - i%8) of byte i/8 = (word i != 0 as an integer)
//   then a complete float archive of the non-zero words in index order.
// The reference builds a byte-per-word bitmap, packs it, runs thrust::exclusive_scan per batch
// element on the default stream between two device synchronisations, and scatters with a u32
// index array. Here a single persistent kernel per direction does it on the caller's stream:
// every warp holds 32 steps x 32 words in registers, one ballot per step is the packed bitmap,
// popc gives the counts, and the tile's output offset comes from a decoupled look-back.
#include <cstdlib>

#include "internal.h"
#include "knobs.h"
#include "prof.h"

namespace dgb {

constexpr uint32_t kSpSteps = 32;                        // steps of 32 words per warp
constexpr uint32_t kSpWarpWords = kSpSteps * 32;         // 1024
constexpr uint32_t kSpTileWords = kSpWarpWords * 8;      // 8192 words per CTA tile

struct SparseParams {
  const Elem* elems;   // tiles sized with kSpTileWords
  uint32_t numElems;
  uint32_t totalTiles;
  uint8_t* dense;      // dense rows at dense + planeOff * wordBytes
  uint32_t* nnz;       // [numElems]
  unsigned long long* tileState;
  uint32_t* ticket;
  uint32_t* errFlag;
  uint8_t* outSuccess; // scatter only, nullable
  uint32_t* outSize;   // scatter only, nullable
  int prefetch;        // compact: L2 prefetch of the tile one grid ahead
};

// ballot (bit l = lane l) -> the 4 bitmap bytes of these 32 words as a little-endian u32
__device__ __forceinline__ uint32_t ballotToBitmap(uint32_t b) { return __byte_perm(__brev(b), 0, 0x0123); }
__device__ __forceinline__ uint32_t bitmapToBallot(uint32_t w) { return __brev(__byte_perm(w, 0, 0x0123)); }

template <typename Word>
__global__ void __launch_bounds__(256, sizeof(Word) == 8 ? 2 : 4) k_sparseCompact(const SparseParams p) {
  __shared__ uint32_t sTicket, sElem, sExcl;
  __shared__ uint32_t sWarpTot[8];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t ltMask = laneMaskLt();
  uint32_t guess = 0, pguess = 0;
  while (true) {
    // (the ticket is taken right before the work, not an iteration ahead: a CTA sitting on a ticket delays the
    // look-back of every later tile - measured 8 % slower)
    if (tid == 0) {
      const uint32_t tk = atomicAdd(p.ticket, 1u);
      sTicket = tk;
      if (tk < p.totalTiles) sElem = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tk, guess);
    }
    __syncthreads();
    const uint32_t tk = sTicket;
    if (tk >= p.totalTiles) break;
    const uint32_t e = sElem;
    guess = e;
    const Elem el = p.elems[e];
    const uint32_t t = tk - el.tileBase;
    const uint32_t n = el.n;
    const uint32_t numTiles = max(1u, divUp(n, kSpTileWords));
    const Word* in = (const Word*)el.in;
    // Tickets are handed out in order and the CTAs advance together, so the tile one grid further on is taken by
    // somebody about one iteration from now: pull it into L2 (one 128-byte line per thread and request). Every
    // tile's exclusive offset waits for its predecessors (look-back), so the slowest load of a whole window of
    // tiles sets the pace; L2 hits are what keeps that tail short.
    {
      const uint32_t pt = tk + gridDim.x;
      if (p.prefetch && pt < p.totalTiles) {
        const uint32_t pe = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, pt, pguess);
        pguess = pe;
        const Elem pel = p.elems[pe];
        const size_t bytes = (size_t)pel.n * sizeof(Word);
        const size_t tileOff = (size_t)(pt - pel.tileBase) * kSpTileWords * sizeof(Word);
        for (uint32_t off = tid * 128; off < kSpTileWords * sizeof(Word); off += 256 * 128)
          if (tileOff + off < bytes) asm volatile("prefetch.global.L2 [%0];" ::"l"((const uint8_t*)pel.in + tileOff + off));
      }
    }
    uint8_t* out = (uint8_t*)el.out;
    uint32_t* bitmap = (uint32_t*)(out + 16);
    Word* dense = (Word*)p.dense + el.planeOff;

    // Reference quirk (GpuSparseFloatCompress.cuh:163-183): the last word goes to idx[N-2] + 1 and
    // nnz = idx[N-2] + bitmap[N-1] + 1, i.e. word N-2 is assumed non-zero. When it is zero the
    // dense array has one extra slot (never written by the reference; 0 here) before the last word.
    const bool quirk = n >= 2 && in[n - 2] == Word(0);

    // ---- phase A: this warp's 1024 words -> 32 ballots (lane k keeps the ballot of words [32k, 32k + 32)) ----
    // Full, 16-byte aligned warp blocks (all but an element's last block and misaligned inputs) load 16 bytes per
    // lane and step - W consecutive words - and keep them in registers; the W-bit masks of the 32 / W lanes that
    // share a 32-word group are OR-ed together with shuffles. (ncu, round 2: with one word per lane and step the
    // kernel was instruction bound, 34 warp instructions per 32 words at 46 % issue activity; this path needs
    // about 10.) Other blocks take one word per lane and step and keep only the ballots: their non-zeros are read
    // again in phase B.
    constexpr uint32_t W = 16 / sizeof(Word);       // words per lane and step
    constexpr uint32_t S = kSpWarpWords / (32 * W); // steps per warp block
    constexpr uint32_t LPG = 32 / W;                // lanes that share a 32-word group
    const uint32_t wbase = t * kSpTileWords + w * kSpWarpWords;
    const bool fast = (((uintptr_t)in) & 15u) == 0 && wbase + kSpWarpWords <= n && !(quirk && wbase + kSpWarpWords == n);
    // (ncu on this version, profiles/r2_k_sparseCompact_v2.txt: 54 M instead of 71 M instructions, but 46 % of the
    // samples now sit at the barrier behind the tile-offset look-up and issue activity is 34 %: in a persistent grid
    // every tile waits for the slowest load among its predecessors. Not keeping the vectors across that wait -
    // reloading the ones with a non-zero in phase B, 32 registers, 8 CTAs per SM - was slower, 0.151 vs 0.135 ms: the
    // eight loads in flight per thread then spill. An aggregate-only offset scheme without any chain was no faster
    // than the look-back either, 0.156 ms.)
    const uint4* v4 = (const uint4*)(in + wbase);
    uint4 x[S];
    uint32_t myBallot = 0;
    if (fast) {
#pragma unroll
      for (uint32_t s0 = 0; s0 < S; ++s0) x[s0] = __ldg(v4 + s0 * 32 + lane);
#pragma unroll
      for (uint32_t s0 = 0; s0 < S; ++s0) {
        uint32_t m; // bit j: word j of this lane's vector is non-zero
        if constexpr (sizeof(Word) == 4) {
          m = (x[s0].x != 0) | ((x[s0].y != 0) << 1) | ((x[s0].z != 0) << 2) | ((x[s0].w != 0) << 3);
        } else if constexpr (sizeof(Word) == 2) {
          m = ((x[s0].x & 0xffffu) != 0) | (((x[s0].x >> 16) != 0) << 1) | (((x[s0].y & 0xffffu) != 0) << 2) |
              (((x[s0].y >> 16) != 0) << 3) | (((x[s0].z & 0xffffu) != 0) << 4) | (((x[s0].z >> 16) != 0) << 5) |
              (((x[s0].w & 0xffffu) != 0) << 6) | (((x[s0].w >> 16) != 0) << 7);
        } else {
          m = ((x[s0].x | x[s0].y) != 0) | (((x[s0].z | x[s0].w) != 0) << 1);
        }
        uint32_t b = m << ((lane % LPG) * W); // in its place inside the group's ballot
#pragma unroll
        for (uint32_t o = 1; o < LPG; o <<= 1) b |= __shfl_xor_sync(0xffffffffu, b, o);
        // this step holds groups [s0 * W, s0 * W + W); lane k of that range takes group k's ballot from its first lane
        const uint32_t got = __shfl_sync(0xffffffffu, b, (lane % W) * LPG);
        if (lane / W == s0) myBallot = got;
      }
    } else {
#pragma unroll 4
      for (uint32_t k = 0; k < kSpSteps; ++k) {
        const uint32_t i = wbase + k * 32 + lane;
        const Word val = i < n ? in[i] : Word(0);
        const uint32_t b = __ballot_sync(0xffffffffu, val != Word(0));
        if (lane == k) myBallot = b;
      }
    }
    const uint32_t myCount = __popc(myBallot);
    uint32_t inc = myCount; // inclusive scan over the 32 groups
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint32_t y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (uint32_t)o) inc += y;
    }
    if (lane == 31) sWarpTot[w] = inc;
    // bitmap words of this warp: group k covers words [wbase + 32k, +32) -> bitmap u32 index (wbase/32 + k)
    {
      uint32_t stepWord = wbase + lane * 32;
      if (stepWord < roundUp(n, 32)) bitmap[stepWord >> 5] = ballotToBitmap(myBallot);
    }
    __syncthreads();
    uint32_t warpOff = 0, tileTot = 0;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) {
      uint32_t y = sWarpTot[j];
      warpOff += j < w ? y : 0u;
      tileTot += y;
    }
    if (w == 0) {
      uint32_t ex = lookbackExclusive(p.tileState + el.tileBase, t, tileTot, p.errFlag);
      if (lane == 0) sExcl = ex;
    }
    __syncthreads();
    const uint32_t excl = sExcl;
    const uint32_t stepOff = excl + warpOff + inc - myCount; // lane k: where group k's non-zeros start

    // ---- phase B: the non-zeros go to their dense positions ----
    if (fast) {
#pragma unroll
      for (uint32_t s0 = 0; s0 < S; ++s0) {
        const uint32_t g = s0 * W + lane / LPG, off = (lane % LPG) * W; // this lane's group and first bit in it
        const uint32_t b = __shfl_sync(0xffffffffu, myBallot, g);
        uint32_t dst = __shfl_sync(0xffffffffu, stepOff, g) + __popc(b & ((1u << off) - 1u));
        const uint32_t m = (b >> off) & ((1u << W) - 1u);
        const uint4 xs = x[s0];
        Word vals[W];
        if constexpr (sizeof(Word) == 4) {
          vals[0] = xs.x; vals[1] = xs.y; vals[2] = xs.z; vals[3] = xs.w;
        } else if constexpr (sizeof(Word) == 2) {
          vals[0] = (Word)xs.x; vals[1] = (Word)(xs.x >> 16); vals[2] = (Word)xs.y; vals[3] = (Word)(xs.y >> 16);
          vals[4] = (Word)xs.z; vals[5] = (Word)(xs.z >> 16); vals[6] = (Word)xs.w; vals[7] = (Word)(xs.w >> 16);
        } else {
          vals[0] = (Word)xs.x | ((Word)xs.y << 32);
          vals[1] = (Word)xs.z | ((Word)xs.w << 32);
        }
#pragma unroll
        for (uint32_t j = 0; j < W; ++j) {
          if ((m >> j) & 1u) dense[dst++] = vals[j];
        }
      }
    } else {
#pragma unroll 4
      for (uint32_t k = 0; k < kSpSteps; ++k) {
        const uint32_t b = __shfl_sync(0xffffffffu, myBallot, k);
        const uint32_t o = __shfl_sync(0xffffffffu, stepOff, k);
        const uint32_t i = wbase + k * 32 + lane;
        uint32_t dst = o + __popc(b & ltMask);
        if (quirk && i == n - 1) {
          dense[dst] = Word(0);
          dst += 1;
        }
        if ((b >> lane) & 1u) dense[dst] = in[i];
      }
    }
    if (t == 0 && tid == 0) ((uint4*)out)[0] = make_uint4(n, 0u, 0u, 0u); // GpuSparseFloatHeader
    if (t == numTiles - 1) {
      if (tid == 0) p.nnz[e] = excl + tileTot + (quirk ? 1u : 0u);
      // zero the bitmap padding up to its 16-byte boundary
      uint32_t firstPad = roundUp(n, 32) >> 5, endPad = sparseBitmapBytes(n) >> 2;
      if (firstPad + tid < endPad) bitmap[firstPad + tid] = 0;
    }
  }
}

// ---- decompress: count -> scan -> scatter, three dependency-free launches ---------------------------------------
// Round 1's single-pass scatter took its tile offsets from a ticket-ordered decoupled look-back: every tile sat on a
// chain of dependent global round trips (1.5 TB/s). The bitmap is 32 x smaller than the data, so counting it first is
// nearly free, and then no CTA ever waits for another: sparse fp32 16 x 16 MiB decompress 1475 -> 1782 GB/s on B200.
// (The same three-launch form for COMPRESS measured slower than k_sparseCompact above, 1049 vs 1231 GB/s: there the
// count pass has to read all the data a second time.)

// sum of `v` over the CTA's 8 warps (all threads get it); sRed: 8 words
__device__ __forceinline__ uint32_t ctaSum8(uint32_t v, uint32_t* sRed, uint32_t lane, uint32_t w) {
  v = warpSum(v);
  __syncthreads();
  if (lane == 0) sRed[w] = v;
  __syncthreads();
  uint32_t t = 0;
#pragma unroll
  for (uint32_t j = 0; j < 8; ++j) t += sRed[j];
  return t;
}

// bitmap -> per-tile non-zero counts: ONE WARP per 8192-word tile (its bitmap is 1 KB = two 16-byte loads per lane), no
// block barrier. (The first version spent a CTA and two __syncthreads per tile: 17 us for 8 MB of bitmap.)
__global__ void __launch_bounds__(256) k_sparseCountBitmap(const SparseParams p, uint32_t* __restrict__ tileOff) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t warp = blockIdx.x * 8u + (threadIdx.x >> 5), numWarps = gridDim.x * 8u;
  uint32_t guess = 0;
  for (uint32_t tk = warp; tk < p.totalTiles; tk += numWarps) {
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tk, guess);
    guess = e;
    const Elem el = p.elems[e];
    const uint32_t t = tk - el.tileBase;
    const uint8_t* arch = (const uint8_t*)el.in;
    const uint32_t n = *(const uint32_t*)arch;
    uint32_t c = 0;
    if (el.n >= n) { // capacity ok (else nothing of this element is decoded)
      // lane l owns bitmap words [8l, 8l + 8) of the tile = data words [t * 8192 + 256 l, + 256); the bitmap is padded
      // with zero bits up to a multiple of 16 bytes, words beyond that are not read
      const uint32_t firstWord = t * (kSpTileWords / 32) + lane * 8u; // in u32 units of the bitmap
      const uint32_t validWords = sparseBitmapBytes(n) >> 2;
      const uint4* bm = (const uint4*)(arch + 16);
      if (firstWord + 4 <= validWords) {
        const uint4 x = bm[firstWord >> 2];
        c += __popc(x.x) + __popc(x.y) + __popc(x.z) + __popc(x.w);
      }
      if (firstWord + 8 <= validWords) {
        const uint4 x = bm[(firstWord >> 2) + 1];
        c += __popc(x.x) + __popc(x.y) + __popc(x.z) + __popc(x.w);
      }
    }
    c = warpSum(c);
    if (lane == 0) tileOff[tk] = c;
  }
}

// One CTA per element: tile counts -> exclusive offsets in place, 1024 tiles per iteration; also reports
// success / size (dietgpu/float/GpuSparseFloatDecompress.cuh:119-121: the reported size is the total N).
__global__ void __launch_bounds__(256) k_sparseScan(const SparseParams p, uint32_t* __restrict__ tileOff) {
  __shared__ uint32_t sWarp[8];
  __shared__ uint32_t sCarry;
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t e = blockIdx.x;
  const Elem el = p.elems[e];
  const uint32_t tb = el.tileBase, T = p.elems[e + 1].tileBase - tb;
  uint32_t* cnt = tileOff + tb;
  if (tid == 0) sCarry = 0;
  __syncthreads();
  for (uint32_t base = 0; base < T; base += 1024) {
    uint32_t c[4];
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) {
      const uint32_t i = base + tid * 4 + j;
      c[j] = i < T ? cnt[i] : 0u;
    }
    const uint32_t mine = c[0] + c[1] + c[2] + c[3];
    uint32_t inc = mine;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (uint32_t)o) inc += x;
    }
    if (lane == 31) sWarp[w] = inc;
    __syncthreads();
    uint32_t off = sCarry + inc - mine, total = 0;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) {
      const uint32_t x = sWarp[j];
      off += j < w ? x : 0u;
      total += x;
    }
#pragma unroll
    for (uint32_t j = 0; j < 4; ++j) {
      const uint32_t i = base + tid * 4 + j;
      if (i < T) cnt[i] = off;
      off += c[j];
    }
    __syncthreads();
    if (tid == 0) sCarry += total;
    __syncthreads();
  }
  if (tid == 0) {
    const uint32_t n = *(const uint32_t*)el.in;
    if (p.outSuccess) p.outSuccess[e] = el.n >= n ? 1 : 0;
    if (p.outSize) p.outSize[e] = n;
  }
}

template <typename Word>
__global__ void __launch_bounds__(256) k_sparseScatter(const SparseParams p, const uint32_t* __restrict__ tileOff) {
  __shared__ uint32_t sWarpTot[8];
  const uint32_t tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
  const uint32_t ltMask = laneMaskLt();
  uint32_t guess = 0;
  for (uint32_t tk = blockIdx.x; tk < p.totalTiles; tk += gridDim.x) {
    const uint32_t e = findElement(&p.elems[0].tileBase, sizeof(Elem) / 4, p.numElems, tk, guess);
    guess = e;
    const Elem el = p.elems[e];
    const uint32_t t = tk - el.tileBase;
    const uint8_t* arch = (const uint8_t*)el.in;
    const uint32_t n = *(const uint32_t*)arch; // total words
    if (el.n < n || t >= divUp(n, kSpTileWords)) continue; // capacity too small (reported by the scan) / padding tile
    const uint32_t* bitmap = (const uint32_t*)(arch + 16);
    const Word* __restrict__ dense = (const Word*)p.dense + el.planeOff;
    Word* __restrict__ out = (Word*)el.out;
    const uint32_t wbase = t * kSpTileWords + w * kSpWarpWords;
    const uint32_t stepWord = wbase + lane * 32;
    uint32_t myBallot = 0; // lane k holds the bitmap of step k (32 words)
    if (stepWord < n) myBallot = bitmapToBallot(bitmap[stepWord >> 5]);
    const uint32_t myCount = __popc(myBallot);
    uint32_t inc = myCount;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t x = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= (uint32_t)o) inc += x;
    }
    __syncthreads(); // previous tile's readers of sWarpTot are done
    if (lane == 31) sWarpTot[w] = inc;
    __syncthreads();
    uint32_t warpOff = 0;
#pragma unroll
    for (uint32_t j = 0; j < 8; ++j) warpOff += j < w ? sWarpTot[j] : 0u;
    const uint32_t stepOff = tileOff[tk] + warpOff + inc - myCount; // lane k: where step k's non-zeros start
    // the last word lives at idx[N-2] + 1 whatever word N-2 was (GpuSparseFloatDecompress.cuh:127-144)
    bool quirk = false;
    if (n >= 2) quirk = ((((const uint8_t*)bitmap)[(n - 2) >> 3] >> (7 - ((n - 2) & 7))) & 1u) == 0;
    // Full, aligned warp block (the common case): every lane produces 16 bytes per step - W consecutive words, their
    // non-zeros consecutive in the dense array - and stores them as one vector. (ncu, round 2: the one-word-per-lane
    // loop below is instruction bound, 25 warp instructions per 32 words at 66 % issue activity.)
    constexpr uint32_t W = 16 / sizeof(Word);
    if ((((uintptr_t)out) & 15u) == 0 && wbase + kSpWarpWords <= n && !(quirk && wbase + kSpWarpWords == n)) {
      uint4* o16 = (uint4*)(out + wbase);
#pragma unroll
      for (uint32_t s = 0; s < kSpWarpWords / (32 * W); ++s) {
        const uint32_t wi = s * 32 * W + lane * W; // first word of this lane inside the warp block
        const uint32_t g = wi >> 5, off = wi & 31u;  // its 32-word group and position in it
        const uint32_t b = __shfl_sync(0xffffffffu, myBallot, g);
        const uint32_t o = __shfl_sync(0xffffffffu, stepOff, g) + __popc(b & ((1u << off) - 1u));
        const uint32_t bits = (b >> off) & ((1u << W) - 1u);
        Word v[W];
#pragma unroll
        for (uint32_t j = 0; j < W; ++j) {
          v[j] = Word(0);
          if ((bits >> j) & 1u) v[j] = __ldg(dense + o + __popc(bits & ((1u << j) - 1u)));
        }
        uint4 x;
        if constexpr (sizeof(Word) == 2) {
          x = make_uint4((uint32_t)v[0] | ((uint32_t)v[1] << 16), (uint32_t)v[2] | ((uint32_t)v[3] << 16),
                         (uint32_t)v[4] | ((uint32_t)v[5] << 16), (uint32_t)v[6] | ((uint32_t)v[7] << 16));
        } else if constexpr (sizeof(Word) == 4) {
          x = make_uint4((uint32_t)v[0], (uint32_t)v[1], (uint32_t)v[2], (uint32_t)v[3]);
        } else {
          x = make_uint4((uint32_t)v[0], (uint32_t)((unsigned long long)v[0] >> 32), (uint32_t)v[1],
                         (uint32_t)((unsigned long long)v[1] >> 32));
        }
        o16[s * 32 + lane] = x;
      }
      continue;
    }
#pragma unroll
    for (uint32_t k0 = 0; k0 < kSpSteps; k0 += 8) { // the loads of 8 steps are issued before their stores
      Word v[8];
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint32_t b = __shfl_sync(0xffffffffu, myBallot, k0 + k);
        const uint32_t o = __shfl_sync(0xffffffffu, stepOff, k0 + k);
        const uint32_t i = wbase + (k0 + k) * 32 + lane;
        v[k] = Word(0);
        if ((b >> lane) & 1u) v[k] = __ldg(dense + o + __popc(b & ltMask) + ((quirk && i == n - 1) ? 1u : 0u));
      }
#pragma unroll
      for (uint32_t k = 0; k < 8; ++k) {
        const uint32_t i = wbase + (k0 + k) * 32 + lane;
        if (i < n) out[i] = v[k];
      }
    }
  }
}

static void sparseParams(SparseParams& p, Temp& tmp, const Elem* elems, uint32_t numElems, uint32_t totalTiles) {
  size_t zeroBytes = (size_t)totalTiles * 8 + 16;
  uint8_t* zero = (uint8_t*)tmp.alloc(zeroBytes);
  p.elems = elems;
  p.numElems = numElems;
  p.totalTiles = totalTiles;
  p.tileState = (unsigned long long*)zero;
  p.ticket = zero ? (uint32_t*)(zero + (size_t)totalTiles * 8) : nullptr;
  p.errFlag = zero ? p.ticket + 1 : nullptr;
  p.outSuccess = nullptr;
  p.outSize = nullptr;
}

int launchSparseCompact(Temp& tmp, int ft, const Elem* elems, uint32_t numElems, uint32_t totalTiles,
                        uint8_t* dense, uint32_t* nnz, cudaStream_t s) {
  if (!numElems) return kOk;
  SparseParams p;
  sparseParams(p, tmp, elems, numElems, totalTiles);
  if (tmp.failed()) return kErrCuda;
  p.dense = dense;
  p.nnz = nnz;
  p.prefetch = DGB_KNOB("DGB_SPARSE_PREFETCH", 1); // experiment knob
  DGB_CUDA(cudaMemsetAsync(p.tileState, 0, (size_t)totalTiles * 8 + 16, s));
  uint32_t grid = min(totalTiles, (uint32_t)smCount() * 4u);
  ProfScope ps__("k_sparseCompact", s);
  switch (floatWordBytes(ft)) {
    case 2: k_sparseCompact<uint16_t><<<grid, 256, 0, s>>>(p); break;
    case 4: k_sparseCompact<uint32_t><<<grid, 256, 0, s>>>(p); break;
    default: k_sparseCompact<unsigned long long><<<grid, 256, 0, s>>>(p); break;
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

template <typename Word>
static void scatterLaunch(const SparseParams& p, uint32_t* tileOff, cudaStream_t s) {
  const uint32_t grid = min(p.totalTiles, (uint32_t)smCount() * 8u);
  {
    ProfScope ps__("k_sparseCountBitmap", s);
    k_sparseCountBitmap<<<min(divUp(p.totalTiles, 8u), (uint32_t)smCount() * 8u), 256, 0, s>>>(p, tileOff);
  }
  {
    ProfScope ps__("k_sparseScan", s);
    k_sparseScan<<<p.numElems, 256, 0, s>>>(p, tileOff);
  }
  {
    ProfScope ps__("k_sparseScatter", s);
    k_sparseScatter<Word><<<grid, 256, 0, s>>>(p, tileOff);
  }
}

int launchSparseScatter(Temp& tmp, int ft, const Elem* elems, uint32_t numElems, uint32_t totalTiles,
                        const uint8_t* dense, uint8_t* outSuccess, uint32_t* outSize, cudaStream_t s) {
  if (!numElems) return kOk;
  SparseParams p;
  p.elems = elems;
  p.numElems = numElems;
  p.totalTiles = totalTiles;
  p.tileState = nullptr;
  p.ticket = nullptr;
  p.errFlag = nullptr;
  p.prefetch = 0;
  p.dense = (uint8_t*)dense;
  p.nnz = nullptr;
  p.outSuccess = outSuccess;
  p.outSize = outSize;
  uint32_t* tileOff = tmp.alloc<uint32_t>((size_t)totalTiles + 4); // counts, then exclusive offsets; fully rewritten per call
  if (tmp.failed()) return kErrCuda;
  switch (floatWordBytes(ft)) {
    case 2: scatterLaunch<uint16_t>(p, tileOff, s); break;
    case 4: scatterLaunch<uint32_t>(p, tileOff, s); break;
    default: scatterLaunch<unsigned long long>(p, tileOff, s); break;
  }
  DGB_CUDA(cudaGetLastError());
  return kOk;
}

} // namespace dgb
