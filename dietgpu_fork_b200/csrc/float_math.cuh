// Bit math of the float codecs: word -> (exponent byte(s), raw bits) and back.
//
// Restates FloatTypeInfo<FT>::split/join (dietgpu/float/GpuFloatUtils.cuh:173-391) and the raw
// section layouts written by the reference's split kernel (dietgpu/float/GpuFloatCompress.cuh:
// 44-165): every word is rotated left by one so the sign lands in bit 0 and the exponent fills the
// top byte(s); the top byte(s) go to the ANS coder, the rest is stored verbatim.
//
//   fp16  comp = w >> 8            raw  u8  [roundUp(N,16)] = w & 0xff
//   bf16  comp = exponent          raw  u8  [roundUp(N,16)] = mantissa7 << 1 | sign
//   fp32  comp = v >> 24           raw  u16 lo[roundUp(N,8)] = v & 0xffff ; u8 hi[roundUp(N,16)] = (v>>16)&0xff
//   fp64  comp0 = v >> 56, comp1 = (v >> 48) & 0xff
//                                  raw  u32 lo[roundUp(N,4)] = v & 0xffffffff ; u16 hi[roundUp(N,8)] = (v>>32)&0xffff
//
// The vector forms work on one 16-byte input vector per thread ("group"): 8 half words, 4 fp32
// words or 2 fp64 words, so that a warp's loads and stores are contiguous.
#pragma once
#include "common.cuh"

namespace dgb {

template <int FT> struct FloatTraits;
template <> struct FloatTraits<kFtF16> { using Word = uint16_t; static constexpr int kGroup = 8; static constexpr int kPlanes = 1; };
template <> struct FloatTraits<kFtBF16> { using Word = uint16_t; static constexpr int kGroup = 8; static constexpr int kPlanes = 1; };
template <> struct FloatTraits<kFtF32> { using Word = uint32_t; static constexpr int kGroup = 4; static constexpr int kPlanes = 1; };
template <> struct FloatTraits<kFtF64> { using Word = uint64_t; static constexpr int kGroup = 2; static constexpr int kPlanes = 2; };

// Addresses of the sections of one float archive's raw area (base = archive + 32).
struct RawSections {
  uint8_t* a; // fp16/bf16: u8 raw ; fp32: u16 lo ; fp64: u32 lo
  uint8_t* b; // fp32: u8 hi ; fp64: u16 hi
};
__device__ __forceinline__ RawSections rawSections(int ft, uint8_t* rawBase, uint32_t n) {
  RawSections r;
  r.a = rawBase;
  r.b = rawBase;
  if (ft == kFtF32) r.b = rawBase + 2u * roundUp(n, 8);
  if (ft == kFtF64) r.b = rawBase + 4u * roundUp(n, 4);
  return r;
}

// per-halfword rotate-left-by-one of two packed 16-bit words
// one bit-select LOP3 ((a & m) | (b & ~m)) instead of two ANDs with different immediates and an OR
__device__ __forceinline__ uint32_t rotl1x2(uint32_t x) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, 0xfffefffe, 0xE4;" : "=r"(r) : "r"(x << 1), "r"(x >> 15));
  return r;
}
__device__ __forceinline__ uint32_t rotr1x2(uint32_t y) {
  return ((y >> 1) & 0x7fff7fffu) | ((y << 15) & 0x80008000u);
}

// ---- scalar forms (unaligned inputs, tails) ------------------------------------------------
template <int FT>
__device__ __forceinline__ void splitScalar(const void* in, uint32_t i, uint8_t* plane0, uint8_t* plane1,
                                            const RawSections& rs, uint32_t& e0, uint32_t& e1) {
  e1 = 0;
  if constexpr (FT == kFtF16) {
    uint32_t w = ((const uint16_t*)in)[i];
    e0 = w >> 8;
    rs.a[i] = (uint8_t)w;
  } else if constexpr (FT == kFtBF16) {
    uint32_t w = ((const uint16_t*)in)[i];
    e0 = (w >> 7) & 0xffu;
    rs.a[i] = (uint8_t)(((w << 1) & 0xfeu) | (w >> 15));
  } else if constexpr (FT == kFtF32) {
    uint32_t v = __funnelshift_l(((const uint32_t*)in)[i], ((const uint32_t*)in)[i], 1);
    e0 = v >> 24;
    ((uint16_t*)rs.a)[i] = (uint16_t)v;
    rs.b[i] = (uint8_t)(v >> 16);
  } else {
    uint64_t w = ((const uint64_t*)in)[i];
    uint64_t v = (w << 1) | (w >> 63);
    e0 = (uint32_t)(v >> 56);
    e1 = (uint32_t)(v >> 48) & 0xffu;
    ((uint32_t*)rs.a)[i] = (uint32_t)v;
    ((uint16_t*)rs.b)[i] = (uint16_t)(v >> 32);
    plane1[i] = (uint8_t)e1;
  }
  plane0[i] = (uint8_t)e0;
}

template <int FT>
__device__ __forceinline__ void joinScalar(void* out, uint32_t i, uint32_t e0, uint32_t e1,
                                           const RawSections& rs) {
  if constexpr (FT == kFtF16) {
    ((uint16_t*)out)[i] = (uint16_t)((e0 << 8) | rs.a[i]);
  } else if constexpr (FT == kFtBF16) {
    uint32_t r = rs.a[i];
    ((uint16_t*)out)[i] = (uint16_t)(((r & 1u) << 15) | (e0 << 7) | (r >> 1));
  } else if constexpr (FT == kFtF32) {
    uint32_t v = (e0 << 24) | ((uint32_t)rs.b[i] << 16) | ((const uint16_t*)rs.a)[i];
    ((uint32_t*)out)[i] = __funnelshift_r(v, v, 1);
  } else {
    uint64_t v = ((uint64_t)e0 << 56) | ((uint64_t)e1 << 48) |
                 ((uint64_t)((const uint16_t*)rs.b)[i] << 32) | ((const uint32_t*)rs.a)[i];
    ((uint64_t*)out)[i] = (v >> 1) | (v << 63);
  }
}

// ---- vector forms: one uint4 of input words per call ------------------------------------------
// g = group index inside the element (word index = g * kGroup). All section bases are 16-B
// aligned, so the stores below are naturally aligned. Returns the exponent bytes packed in
// e0 (and e1 for fp64): kGroup bytes, lowest word in the lowest byte (bf16/fp16 return 8 bytes
// as a uint2 through e0lo/e0hi).
template <int FT>
__device__ __forceinline__ void splitGroup(uint4 x, uint32_t g, uint8_t* plane0, uint8_t* plane1,
                                           const RawSections& rs, uint32_t (&ex)[2], uint32_t (&ex1)[2]) {
  ex1[0] = ex1[1] = 0;
  if constexpr (FT == kFtF16 || FT == kFtBF16) {
    uint32_t y0 = x.x, y1 = x.y, y2 = x.z, y3 = x.w;
    if constexpr (FT == kFtBF16) {
      y0 = rotl1x2(y0); y1 = rotl1x2(y1); y2 = rotl1x2(y2); y3 = rotl1x2(y3);
    }
    ex[0] = __byte_perm(y0, y1, 0x7531);
    ex[1] = __byte_perm(y2, y3, 0x7531);
    uint2 raw = make_uint2(__byte_perm(y0, y1, 0x6420), __byte_perm(y2, y3, 0x6420));
    ((uint2*)plane0)[g] = make_uint2(ex[0], ex[1]);
    ((uint2*)rs.a)[g] = raw;
  } else if constexpr (FT == kFtF32) {
    uint32_t v0 = __funnelshift_l(x.x, x.x, 1), v1 = __funnelshift_l(x.y, x.y, 1);
    uint32_t v2 = __funnelshift_l(x.z, x.z, 1), v3 = __funnelshift_l(x.w, x.w, 1);
    uint32_t a = __byte_perm(v0, v1, 0x7632); // [v0.b2 v0.b3 v1.b2 v1.b3]
    uint32_t b = __byte_perm(v2, v3, 0x7632);
    ex[0] = __byte_perm(a, b, 0x7531);
    ex[1] = 0;
    uint32_t hi = __byte_perm(a, b, 0x6420);
    uint2 lo = make_uint2(__byte_perm(v0, v1, 0x5410), __byte_perm(v2, v3, 0x5410));
    ((uint32_t*)plane0)[g] = ex[0];
    ((uint2*)rs.a)[g] = lo;
    ((uint32_t*)rs.b)[g] = hi;
  } else {
    // two fp64 words: (x.x, x.y) and (x.z, x.w), little endian
    uint32_t l0 = (x.x << 1) | (x.y >> 31), h0 = (x.y << 1) | (x.x >> 31);
    uint32_t l1 = (x.z << 1) | (x.w >> 31), h1 = (x.w << 1) | (x.z >> 31);
    ex[0] = __byte_perm(h0, h1, 0x0073);  // [h0.b3 h1.b3 0 0] -> only 2 valid bytes
    ex[0] &= 0xffffu;
    ex[1] = 0;
    ex1[0] = __byte_perm(h0, h1, 0x0062) & 0xffffu;
    ((uint16_t*)plane0)[g] = (uint16_t)ex[0];
    ((uint16_t*)plane1)[g] = (uint16_t)ex1[0];
    ((uint2*)rs.a)[g] = make_uint2(l0, l1);
    ((uint32_t*)rs.b)[g] = __byte_perm(h0, h1, 0x5410);
  }
}

// Inverse: exponent bytes come from shared memory (sym0/sym1 point at the group's first byte).
template <int FT>
__device__ __forceinline__ uint4 joinGroup(uint32_t g, const uint8_t* sym0, const uint8_t* sym1,
                                           const RawSections& rs) {
  uint4 o;
  if constexpr (FT == kFtF16 || FT == kFtBF16) {
    uint2 e = *(const uint2*)sym0;
    uint2 r = ((const uint2*)rs.a)[g];
    uint32_t y0 = __byte_perm(r.x, e.x, 0x5140); // [r0 e0 r1 e1]
    uint32_t y1 = __byte_perm(r.x, e.x, 0x7362);
    uint32_t y2 = __byte_perm(r.y, e.y, 0x5140);
    uint32_t y3 = __byte_perm(r.y, e.y, 0x7362);
    if constexpr (FT == kFtBF16) {
      y0 = rotr1x2(y0); y1 = rotr1x2(y1); y2 = rotr1x2(y2); y3 = rotr1x2(y3);
    }
    o = make_uint4(y0, y1, y2, y3);
  } else if constexpr (FT == kFtF32) {
    uint32_t e = *(const uint32_t*)sym0;
    uint2 lo = ((const uint2*)rs.a)[g];
    uint32_t hi = ((const uint32_t*)rs.b)[g];
    uint32_t t01 = __byte_perm(hi, e, 0x5140); // [hi0 e0 hi1 e1]
    uint32_t t23 = __byte_perm(hi, e, 0x7362);
    uint32_t v0 = __byte_perm(lo.x, t01, 0x5410);
    uint32_t v1 = __byte_perm(lo.x, t01, 0x7632);
    uint32_t v2 = __byte_perm(lo.y, t23, 0x5410);
    uint32_t v3 = __byte_perm(lo.y, t23, 0x7632);
    o = make_uint4(__funnelshift_r(v0, v0, 1), __funnelshift_r(v1, v1, 1),
                   __funnelshift_r(v2, v2, 1), __funnelshift_r(v3, v3, 1));
  } else {
    uint32_t e0 = *(const uint16_t*)sym0;
    uint32_t e1 = *(const uint16_t*)sym1;
    uint2 lo = ((const uint2*)rs.a)[g];
    uint32_t hi = ((const uint32_t*)rs.b)[g];
    uint32_t h0 = (hi & 0xffffu) | ((e1 & 0xffu) << 16) | ((e0 & 0xffu) << 24);
    uint32_t h1 = (hi >> 16) | ((e1 >> 8) << 16) | ((e0 >> 8) << 24);
    o.x = (lo.x >> 1) | (h0 << 31);
    o.y = (h0 >> 1) | (lo.x << 31);
    o.z = (lo.y >> 1) | (h1 << 31);
    o.w = (h1 >> 1) | (lo.y << 31);
  }
  return o;
}

} // namespace dgb
