/*
 * dietgpu_oracle.c — CPU restatement of the DietGPU stream formats (see dietgpu_oracle.h).
 * TEST INFRASTRUCTURE ONLY; never linked into the product library.
 *
 * Compile: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (no fast-math: the one fp32
 * divide+multiply in dgo_normalize must match CUDA's IEEE div.rn / mul.rn).
 *
 * A warp is emulated as a 32-entry state array; lanes act in ascending order inside a symbol
 * row. Citations are relative to /root/reference/.
 */
#include "dietgpu_oracle.h"

#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define BLOCK 4096u /* dietgpu/ans/GpuANSUtils.cuh:37 */
#define LANES 32u   /* dietgpu/utils/DeviceDefs.cuh:14 */
#define STATE0 (1u << 15) /* kANSStartState == kANSMinState, GpuANSUtils.cuh:41-49 */
#define ANS_MAGIC 0xD00D0001u
#define FLOAT_MAGIC 0xF00F0001u

static inline uint32_t div_up(uint32_t a, uint32_t b) { return (a + b - 1) / b; }
static inline uint32_t round_up(uint32_t a, uint32_t b) { return div_up(a, b) * b; }
static inline size_t round_up_sz(size_t a, size_t b) { return ((a + b - 1) / b) * b; }
static inline void put32(uint8_t* p, uint32_t v) { memcpy(p, &v, 4); }
static inline uint32_t get32(const uint8_t* p) { uint32_t v; memcpy(&v, p, 4); return v; }
static inline void put16(uint8_t* p, uint16_t v) { memcpy(p, &v, 2); }
static inline uint16_t get16(const uint8_t* p) { uint16_t v; memcpy(&v, p, 2); return v; }

/* ANSCoalescedHeader::getCompressedOverhead, dietgpu/ans/GpuANSUtils.cuh:68-82 */
static inline uint32_t ans_overhead(uint32_t nb) {
  return 32u + 512u + 128u * nb + 8u * round_up(nb, 2);
}

/* ------------------------------------------------------------------ sizes */

/* dietgpu/ans/GpuANSEncode.cu:13-25 — note the reference passes the block SIZE (4096) to
 * getCompressedOverhead where a block count is expected; the value is kept. */
uint32_t dgo_ans_max_compressed_size(uint32_t n) {
  uint32_t blocks = div_up(n, BLOCK);
  size_t raw = ans_overhead(BLOCK);
  raw += (size_t)round_up(BLOCK + BLOCK / 4, 16) * blocks; /* GpuANSEncode.cuh:31-36 */
  return (uint32_t)round_up_sz(raw, 16);
}

uint32_t dgo_float_word_bytes(int ft) {
  switch (ft) {
    case DGO_FLOAT16: case DGO_BFLOAT16: return 2;
    case DGO_FLOAT32: return 4;
    case DGO_FLOAT64: return 8;
    default: return 0;
  }
}

/* FloatTypeInfo<FT>::getUncompDataSize, dietgpu/float/GpuFloatUtils.cuh:204-208, :258-262,
 * :303-314, :370-382 */
uint32_t dgo_float_raw_section_bytes(int ft, uint32_t n) {
  switch (ft) {
    case DGO_FLOAT16: case DGO_BFLOAT16: return round_up(n, 16);
    case DGO_FLOAT32: return 2 * round_up(n, 8) + round_up(n, 16);
    case DGO_FLOAT64: return 4 * round_up(n, 4) + 2 * round_up(n, 8);
    default: return 0;
  }
}

/* dietgpu/float/GpuFloatCompress.cu:23-48 */
uint32_t dgo_float_max_compressed_size(int ft, uint32_t n) {
  uint32_t base = 32u + dgo_ans_max_compressed_size(n) + dgo_float_raw_section_bytes(ft, n);
  if (ft == DGO_FLOAT64) base += dgo_ans_max_compressed_size(n);
  return base;
}

/* dietgpu/float/GpuSparseFloatCompress.cu:16-24 */
uint32_t dgo_sparse_float_max_compressed_size(int ft, uint32_t n) {
  return 16u + round_up((n + 7) / 8, 16) + dgo_float_max_compressed_size(ft, n);
}

/* ------------------------------------------------------------- statistics */

void dgo_histogram(const uint8_t* in, uint32_t n, uint32_t counts[256]) {
  memset(counts, 0, 256 * sizeof(uint32_t));
  for (uint32_t i = 0; i < n; ++i) counts[in[i]]++;
}

/* dietgpu/ans/GpuChecksum.cuh:26-93: XOR of all bytes, folded to 8 bits */
uint32_t dgo_checksum(const uint8_t* in, uint32_t n) {
  uint32_t c = 0;
  for (uint32_t i = 0; i < n; ++i) c ^= in[i];
  return c & 0xffu;
}

/* dietgpu/ans/GpuANSStatistics.cuh:178-341 (normalizeProbabilitiesFromHistogram) */
int dgo_normalize(const uint32_t counts[256], uint32_t total, int prob_bits,
                  uint32_t pdf[256], uint32_t cdf[256]) {
  if (total == 0) return 0; /* :192-195 table untouched */
  uint32_t weight = 1u << prob_bits;
  uint32_t q[256];
  int sum = 0;
  for (int s = 0; s < 256; ++s) {
    /* :205-221 — uint32 * float promotes the uint32 to float; fp32 divide then multiply,
     * truncation toward zero on the conversion back */
    volatile float ratio = (float)counts[s] / (float)total;
    volatile float scaled = (float)weight * ratio;
    uint32_t v = (uint32_t)scaled;
    if (counts[s] > 0 && v == 0) v = 1;
    q[s] = v;
    sum += (int)v;
  }
  /* :229-252 — sort (q << 16 | sym) descending; position p holds symbol order[p] */
  uint32_t key[256];
  for (int s = 0; s < 256; ++s) key[s] = (q[s] << 16) | (uint32_t)s;
  for (int i = 1; i < 256; ++i) { /* insertion sort, descending; keys are distinct */
    uint32_t k = key[i];
    int j = i - 1;
    while (j >= 0 && key[j] < k) { key[j + 1] = key[j]; --j; }
    key[j + 1] = k;
  }
  uint32_t order[256], qs[256]; /* sorted-position view */
  for (int p = 0; p < 256; ++p) { order[p] = key[p] & 0xffffu; qs[p] = key[p] >> 16; }

  int diff = (int)weight - sum;
  if (diff > 0) {
    /* :258-273 — the test is on the SYMBOL ID held at the position, not on the rank */
    while (diff > 0) {
      int k = diff < 256 ? diff : 256;
      for (int p = 0; p < 256; ++p)
        if ((int)order[p] < k) qs[p] += 1;
      diff -= k;
    }
  } else if (diff < 0) {
    /* :274-315 — take 1 from the smallest entries that are > 1, by sorted position */
    diff = -diff;
    while (diff > 0) {
      int g = 0;
      for (int p = 0; p < 256; ++p) g += (qs[p] > 1);
      int k = diff < g ? diff : g;
      if (k <= 0) break; /* reference asserts k > 0 */
      for (int p = g - k; p < g; ++p) qs[p] -= 1;
      diff -= k;
    }
  }
  /* :317-341 un-sort, exclusive scan */
  for (int p = 0; p < 256; ++p) pdf[order[p]] = qs[p];
  uint32_t run = 0;
  for (int s = 0; s < 256; ++s) { cdf[s] = run; run += pdf[s]; }
  return 1;
}

/* ------------------------------------------------------- per-block rANS */

/* Encode one block of n <= 4096 symbols (dietgpu/ans/GpuANSEncode.cuh:49-211).
 * states[32] receives the final lane states; words[] the emitted u16 words in emission
 * order. Returns the word count. */
static uint32_t encode_block(const uint8_t* sym, uint32_t n, const uint32_t* pdf,
                             const uint32_t* cdf, int pb, uint32_t states[LANES],
                             uint16_t* words) {
  uint32_t st[LANES];
  for (uint32_t l = 0; l < LANES; ++l) st[l] = STATE0;
  uint32_t nw = 0;
  uint32_t rows = div_up(n, LANES);
  for (uint32_t r = 0; r < rows; ++r) {
    uint32_t base = r * LANES;
    uint32_t valid = n - base < LANES ? n - base : LANES;
    /* emit decisions use the pre-update state of each lane; lanes ascending (:49-90) */
    for (uint32_t l = 0; l < valid; ++l) {
      uint32_t p = pdf[sym[base + l]];
      uint32_t s = st[l];
      if (s >= (p << (31 - pb))) {
        words[nw++] = (uint16_t)(s & 0xffffu);
        s >>= 16;
      }
      uint32_t c = cdf[sym[base + l]];
      st[l] = ((s / p) << pb) + (s % p) + c;
    }
  }
  memcpy(states, st, sizeof(st));
  return nw;
}

/* Decode one block (dietgpu/ans/GpuANSDecode.cuh:55-217). lut entry = (slot - cdf) << 20 |
 * pdf << 8 | sym (:34-41). */
static void decode_block(const uint32_t states[LANES], const uint16_t* words, uint32_t nw,
                         uint32_t n, const uint32_t* lut, int pb, uint8_t* out) {
  uint32_t st[LANES];
  memcpy(st, states, sizeof(st));
  uint32_t mask = (1u << pb) - 1u;
  uint32_t p = nw;
  uint32_t rows = div_up(n, LANES);
  for (uint32_t r = rows; r-- > 0;) {
    uint32_t base = r * LANES;
    uint32_t valid = n - base < LANES ? n - base : LANES;
    for (uint32_t l = 0; l < valid; ++l) {
      uint32_t e = lut[st[l] & mask];
      out[base + l] = (uint8_t)(e & 0xffu);
      st[l] = ((e >> 8) & 0xfffu) * (st[l] >> pb) + (e >> 20);
    }
    /* readers consume from the end of the word stream, highest lane first (:87-101) */
    for (uint32_t l = valid; l-- > 0;) {
      if (st[l] < STATE0) {
        p -= 1;
        st[l] = (st[l] << 16) + words[p];
      }
    }
  }
}

/* dietgpu/ans/GpuANSDecode.cuh:405-476 */
static void build_decode_lut(const uint16_t* pdf16, int pb, uint32_t* lut) {
  uint32_t c = 0;
  for (uint32_t s = 0; s < 256; ++s) {
    uint32_t p = pdf16[s];
    for (uint32_t j = 0; j < p && c + j < (1u << pb); ++j)
      lut[c + j] = (j << 20) | (p << 8) | s;
    c += p;
  }
}

/* ------------------------------------------------------------ ANS archive */

/* Layout (dietgpu/ans/GpuANSEncode.cuh:511-624):
 *  [0,32) header | [32,544) u16 pdf[256] | states u32[nb][32] | uint2 blockWords[roundUp(nb,2)]
 *  | block payloads, each padded to 8 words. */
static uint32_t ans_encode_impl(const uint8_t* in, uint32_t n, int pb, int use_checksum,
                                uint32_t checksum, const uint32_t* hist, uint8_t* out,
                                int parallel) {
  uint32_t nb = div_up(n, BLOCK);
  uint32_t counts[256], pdf[256], cdf[256];
  memset(pdf, 0, sizeof(pdf));
  memset(cdf, 0, sizeof(cdf));
  if (hist) memcpy(counts, hist, sizeof(counts)); else dgo_histogram(in, n, counts);
  dgo_normalize(counts, n, pb, pdf, cdf);

  uint32_t overhead = ans_overhead(nb);
  memset(out, 0, overhead);
  uint8_t* states = out + 544;
  uint8_t* block_words = states + 128u * nb;
  uint8_t* payload = out + overhead;

  /* encode every block into a private slot, then compact in block order */
  uint16_t* tmp = NULL;
  uint32_t* nwords = NULL;
  const uint32_t slot = 2816; /* 88 words/lane worst case at probBits 11 */
  if (nb) {
    tmp = (uint16_t*)malloc((size_t)nb * slot * sizeof(uint16_t));
    nwords = (uint32_t*)malloc((size_t)nb * sizeof(uint32_t));
  }
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (parallel && nb > 8)
#endif
  for (int64_t b = 0; b < (int64_t)nb; ++b) {
    uint32_t start = (uint32_t)b * BLOCK;
    uint32_t len = n - start < BLOCK ? n - start : BLOCK;
    uint32_t st[LANES];
    nwords[b] = encode_block(in + start, len, pdf, cdf, pb, st, tmp + (size_t)b * slot);
    memcpy(states + 128u * (uint32_t)b, st, sizeof(st));
  }
  uint32_t total_words = 0;
  for (uint32_t b = 0; b < nb; ++b) {
    uint32_t start = b * BLOCK;
    uint32_t len = n - start < BLOCK ? n - start : BLOCK;
    put32(block_words + 8u * b, (len << 16) | nwords[b]);
    put32(block_words + 8u * b + 4, total_words);
    uint32_t padded = round_up(nwords[b], 8);
    memcpy(payload + 2u * (size_t)total_words, tmp + (size_t)b * slot, 2u * nwords[b]);
    memset(payload + 2u * ((size_t)total_words + nwords[b]), 0, 2u * (padded - nwords[b]));
    total_words += padded;
  }
  free(tmp);
  free(nwords);

  put32(out + 0, ANS_MAGIC);
  put32(out + 4, nb);
  put32(out + 8, n);
  put32(out + 12, total_words);
  put32(out + 16, (uint32_t)pb | ((use_checksum ? 1u : 0u) << 4));
  put32(out + 20, use_checksum ? checksum : 0u);
  for (uint32_t s = 0; s < 256; ++s) put16(out + 32 + 2 * s, (uint16_t)pdf[s]);
  return overhead + 2u * total_words;
}

uint32_t dgo_ans_encode(const uint8_t* in, uint32_t n, int pb, int use_checksum,
                        const uint32_t* hist, uint8_t* out) {
  uint32_t cs = use_checksum ? dgo_checksum(in, n) : 0;
  return ans_encode_impl(in, n, pb, use_checksum, cs, hist, out, 1);
}

static int ans_decode_impl(const uint8_t* a, uint8_t* out, uint32_t capacity,
                           uint32_t* out_size, int parallel) {
  uint32_t nb = get32(a + 4), n = get32(a + 8);
  int pb = (int)(get32(a + 16) & 0xfu);
  if (out_size) *out_size = n;
  if (capacity < n) return 0; /* dietgpu/ans/GpuANSDecode.cuh:325-341 */
  if (n == 0) return 1;
  uint32_t* lut = (uint32_t*)malloc(sizeof(uint32_t) << pb);
  build_decode_lut((const uint16_t*)(a + 32), pb, lut);
  const uint8_t* states = a + 544;
  const uint8_t* block_words = states + 128u * nb;
  const uint8_t* payload = a + ans_overhead(nb);
#ifdef _OPENMP
#pragma omp parallel for schedule(static) if (parallel && nb > 8)
#endif
  for (int64_t b = 0; b < (int64_t)nb; ++b) {
    uint32_t x = get32(block_words + 8u * (uint32_t)b);
    uint32_t start_word = get32(block_words + 8u * (uint32_t)b + 4);
    uint32_t st[LANES];
    memcpy(st, states + 128u * (uint32_t)b, sizeof(st));
    decode_block(st, (const uint16_t*)(payload + 2u * (size_t)start_word), x & 0xffffu,
                 x >> 16, lut, pb, out + (size_t)b * BLOCK);
  }
  free(lut);
  return 1;
}

int dgo_ans_decode(const uint8_t* a, uint8_t* out, uint32_t capacity, uint32_t* out_size) {
  return ans_decode_impl(a, out, capacity, out_size, 1);
}

/* Bytes the reference never writes: options bits 5-31, unused0/1,
 * checksum when the flag is off, the odd blockWords padding entry, each block's tail padding
 * words, and pdf[] when N == 0. */
uint32_t dgo_ans_defined_mask(const uint8_t* a, uint8_t* mask) {
  uint32_t nb = get32(a + 4), n = get32(a + 8), tw = get32(a + 12);
  uint32_t opt = get32(a + 16);
  uint32_t overhead = ans_overhead(nb);
  uint32_t total = overhead + 2u * tw;
  memset(mask, 0, total);
  memset(mask, 1, 16);          /* magic, numBlocks, sizes */
  mask[16] = 1;                 /* only bits 0-4 of options are defined */
  if (opt & 0x10u) memset(mask + 20, 1, 4);
  if (n) memset(mask + 32, 1, 512);
  memset(mask + 544, 1, 128u * nb + 8u * nb);
  const uint8_t* bw = a + 544 + 128u * nb;
  for (uint32_t b = 0; b < nb; ++b) {
    uint32_t w = get32(bw + 8u * b) & 0xffffu, s = get32(bw + 8u * b + 4);
    memset(mask + overhead + 2u * (size_t)s, 1, 2u * w);
  }
  return total;
}

/* ----------------------------------------------------------- float codecs */

static inline uint32_t rotl32(uint32_t v, int s) { return (v << s) | (v >> (32 - s)); }
static inline uint32_t rotr32(uint32_t v, int s) { return (v >> s) | (v << (32 - s)); }
static inline uint64_t rotl64(uint64_t v, int s) { return (v << s) | (v >> (64 - s)); }
static inline uint64_t rotr64(uint64_t v, int s) { return (v >> s) | (v << (64 - s)); }

/* Split every word into exponent plane(s) + raw section (dietgpu/float/GpuFloatUtils.cuh:
 * 177-391 bit math; dietgpu/float/GpuFloatCompress.cuh:44-165, :214-245 section layout).
 * raw points at archive+32 and is pre-zeroed by the caller (padding = 0). */
static void float_split(int ft, const void* in, uint32_t n, uint8_t* plane0, uint8_t* plane1,
                        uint8_t* raw) {
  switch (ft) {
    case DGO_FLOAT16: {
      const uint16_t* w = (const uint16_t*)in;
      for (uint32_t i = 0; i < n; ++i) { plane0[i] = (uint8_t)(w[i] >> 8); raw[i] = (uint8_t)w[i]; }
    } break;
    case DGO_BFLOAT16: {
      const uint16_t* w = (const uint16_t*)in;
      for (uint32_t i = 0; i < n; ++i) {
        uint32_t v = rotl32(((uint32_t)w[i] << 16) | w[i], 1); /* :236-242 */
        plane0[i] = (uint8_t)(v >> 24);
        raw[i] = (uint8_t)(v & 0xffu);
      }
    } break;
    case DGO_FLOAT32: {
      const uint32_t* w = (const uint32_t*)in;
      uint8_t* lo = raw;                          /* u16 lo[roundUp(n,8)] */
      uint8_t* hi = raw + 2u * round_up(n, 8);    /* u8  hi[roundUp(n,16)] */
      for (uint32_t i = 0; i < n; ++i) {
        uint32_t v = rotl32(w[i], 1);
        plane0[i] = (uint8_t)(v >> 24);
        put16(lo + 2u * i, (uint16_t)(v & 0xffffu));
        hi[i] = (uint8_t)((v >> 16) & 0xffu);
      }
    } break;
    case DGO_FLOAT64: {
      const uint64_t* w = (const uint64_t*)in;
      uint8_t* lo = raw;                          /* u32 lo[roundUp(n,4)] */
      uint8_t* hi = raw + 4u * round_up(n, 4);    /* u16 hi[roundUp(n,8)] */
      for (uint32_t i = 0; i < n; ++i) {
        uint64_t v = rotl64(w[i], 1);
        plane0[i] = (uint8_t)(v >> 56);
        plane1[i] = (uint8_t)((v >> 48) & 0xffu);
        put32(lo + 4u * i, (uint32_t)(v & 0xffffffffu));
        put16(hi + 2u * i, (uint16_t)((v >> 32) & 0xffffu));
      }
    } break;
  }
}

static void float_join(int ft, const uint8_t* plane0, const uint8_t* plane1, const uint8_t* raw,
                       uint32_t n, void* out) {
  switch (ft) {
    case DGO_FLOAT16: {
      uint16_t* w = (uint16_t*)out;
      for (uint32_t i = 0; i < n; ++i) w[i] = (uint16_t)(((uint32_t)plane0[i] << 8) | raw[i]);
    } break;
    case DGO_BFLOAT16: {
      uint16_t* w = (uint16_t*)out;
      for (uint32_t i = 0; i < n; ++i) {
        /* inverse of the rotate: raw = mantissa7 << 1 | sign (:244-256) */
        uint32_t r = raw[i];
        w[i] = (uint16_t)(((r & 1u) << 15) | ((uint32_t)plane0[i] << 7) | (r >> 1));
      }
    } break;
    case DGO_FLOAT32: {
      uint32_t* w = (uint32_t*)out;
      const uint8_t* lo = raw;
      const uint8_t* hi = raw + 2u * round_up(n, 8);
      for (uint32_t i = 0; i < n; ++i) {
        uint32_t v = ((uint32_t)plane0[i] << 24) | ((uint32_t)hi[i] << 16) | get16(lo + 2u * i);
        w[i] = rotr32(v, 1);
      }
    } break;
    case DGO_FLOAT64: {
      uint64_t* w = (uint64_t*)out;
      const uint8_t* lo = raw;
      const uint8_t* hi = raw + 4u * round_up(n, 4);
      for (uint32_t i = 0; i < n; ++i) {
        uint64_t v = ((uint64_t)plane0[i] << 56) | ((uint64_t)plane1[i] << 48) |
                     ((uint64_t)get16(hi + 2u * i) << 32) | get32(lo + 4u * i);
        w[i] = rotr64(v, 1);
      }
    } break;
  }
}

/* Archive: hdr16 | hdr2_16 | raw section | ANS(plane0) [| ANS(plane1)] */
static uint32_t float_compress_impl(int ft, const void* in, uint32_t n, int pb,
                                    int use_checksum, uint8_t* out, int parallel) {
  uint32_t R = dgo_float_raw_section_bytes(ft, n);
  memset(out, 0, 32u + R);
  uint8_t* plane0 = (uint8_t*)malloc((size_t)n * 2 + 16);
  uint8_t* plane1 = plane0 + n;
  float_split(ft, in, n, plane0, plane1, out + 32);

  put32(out + 0, FLOAT_MAGIC);
  put32(out + 4, n);
  put32(out + 8, (uint32_t)ft | ((use_checksum ? 1u : 0u) << 4));
  /* the float-level checksum covers only the first N BYTES of the input: the provider
   * reports words (dietgpu/float/GpuFloatCompress.cuh:708-710, GpuChecksum.cuh:95-102) */
  put32(out + 12, use_checksum ? dgo_checksum((const uint8_t*)in, n) : 0u);

  uint32_t c0 = ans_encode_impl(plane0, n, pb, 0, 0, NULL, out + 32 + R, parallel);
  put32(out + 16, c0); /* firstCompSegmentBytes, written for every type (:656-667) */
  uint32_t total = 32u + R + c0;
  if (ft == DGO_FLOAT64)
    total += ans_encode_impl(plane1, n, pb, 0, 0, NULL, out + 32 + R + c0, parallel);
  free(plane0);
  return total;
}

uint32_t dgo_float_compress(int ft, const void* in, uint32_t n, int pb, int use_checksum,
                            uint8_t* out) {
  return float_compress_impl(ft, in, n, pb, use_checksum, out, 1);
}

static int float_decompress_impl(const uint8_t* a, void* out, uint32_t cap, uint32_t* out_words,
                                 int parallel) {
  uint32_t n = get32(a + 4);
  int ft = (int)(get32(a + 8) & 0xfu);
  if (out_words) *out_words = n;
  if (cap < n) return 0;
  uint32_t R = dgo_float_raw_section_bytes(ft, n);
  uint8_t* plane0 = (uint8_t*)malloc((size_t)n * 2 + 16);
  uint8_t* plane1 = plane0 + n;
  uint32_t sz;
  ans_decode_impl(a + 32 + R, plane0, n, &sz, parallel);
  if (ft == DGO_FLOAT64) ans_decode_impl(a + 32 + R + get32(a + 16), plane1, n, &sz, parallel);
  float_join(ft, plane0, plane1, a + 32, n, out);
  free(plane0);
  return 1;
}

int dgo_float_decompress(const uint8_t* a, void* out, uint32_t cap, uint32_t* out_words) {
  return float_decompress_impl(a, out, cap, out_words, 1);
}

/* Undefined in the reference: header options bits 5-31, checksum when off, Header2 bytes
 * 4-15, raw-section padding. */
uint32_t dgo_float_defined_mask(const uint8_t* a, uint8_t* mask) {
  uint32_t n = get32(a + 4), opt = get32(a + 8);
  int ft = (int)(opt & 0xfu);
  uint32_t R = dgo_float_raw_section_bytes(ft, n);
  memset(mask, 0, 32u + R);
  memset(mask, 1, 9);
  if (opt & 0x10u) memset(mask + 12, 1, 4);
  memset(mask + 16, 1, 4);
  uint8_t* m = mask + 32;
  switch (ft) {
    case DGO_FLOAT16: case DGO_BFLOAT16: memset(m, 1, n); break;
    case DGO_FLOAT32: memset(m, 1, 2u * n); memset(m + 2u * round_up(n, 8), 1, n); break;
    case DGO_FLOAT64: memset(m, 1, 4u * n); memset(m + 4u * round_up(n, 4), 1, 2u * n); break;
  }
  uint32_t c0 = dgo_ans_defined_mask(a + 32 + R, mask + 32 + R);
  uint32_t total = 32u + R + c0;
  if (ft == DGO_FLOAT64) total += dgo_ans_defined_mask(a + 32 + R + c0, mask + 32 + R + c0);
  return total;
}

/* ------------------------------------------------------------ sparse float */

static int word_is_nonzero(int ft, const void* in, uint32_t i) {
  /* integer compare on the raw word: -0.0 counts as non-zero
   * (dietgpu/float/GpuSparseFloatCompress.cuh:56) */
  switch (ft) {
    case DGO_FLOAT16: case DGO_BFLOAT16: return ((const uint16_t*)in)[i] != 0;
    case DGO_FLOAT32: return ((const uint32_t*)in)[i] != 0;
    default: return ((const uint64_t*)in)[i] != 0;
  }
}

/* Archive: u32 size + 12 pad | bitmap roundUp(ceil(N/8),16), MSB-first
 * (dietgpu/float/GpuSparseFloatCompress.cuh:108-110) | float archive of the non-zero words.
 *
 * Quirk kept from the reference (GpuSparseFloatCompress.cuh:163-183, mirrored by the decoder at
 * GpuSparseFloatDecompress.cuh:127-144): the exclusive scan stops one element short, and the last
 * word is placed at idx[N-2] + 1 with nnz = idx[N-2] + bitmap[N-1] + 1, which assumes word N-2 is
 * non-zero. When word N-2 IS zero the dense array therefore has one extra slot (at index
 * idx[N-2]) that the reference never writes (uninitialised temp memory); this restatement
 * writes 0 there. The reference is undefined for N < 2 (reads index -1); here N < 2 is plain
 * compaction. */
uint32_t dgo_sparse_float_compress(int ft, const void* in, uint32_t n, int pb,
                                   int use_checksum, uint8_t* out) {
  uint32_t wb = dgo_float_word_bytes(ft);
  uint32_t bm = round_up((n + 7) / 8, 16);
  memset(out, 0, 16u + bm);
  put32(out, n);
  uint8_t* nz = (uint8_t*)calloc((size_t)n * wb + 32, 1);
  uint32_t nnz = 0;
  for (uint32_t i = 0; i < n; ++i) {
    int b = word_is_nonzero(ft, in, i);
    if (b) out[16 + (i >> 3)] |= (uint8_t)(0x80u >> (i & 7));
    if (n >= 2 && i == n - 2 && !b) { ++nnz; continue; } /* the never-written slot (0 here) */
    if (b) {
      memcpy(nz + (size_t)nnz * wb, (const uint8_t*)in + (size_t)i * wb, wb);
      ++nnz;
    }
  }
  uint32_t c = float_compress_impl(ft, nz, nnz, pb, use_checksum, out + 16 + bm, 1);
  free(nz);
  return 16u + bm + c;
}

int dgo_sparse_float_decompress(int ft, const uint8_t* a, void* out, uint32_t cap,
                                uint32_t* out_words) {
  uint32_t n = get32(a);
  uint32_t wb = dgo_float_word_bytes(ft);
  uint32_t bm = round_up((n + 7) / 8, 16);
  if (out_words) *out_words = n; /* dietgpu/float/GpuSparseFloatDecompress.cuh:119-121 */
  if (cap < n) return 0;
  const uint8_t* fa = a + 16 + bm;
  uint32_t nnz = get32(fa + 4);
  uint8_t* nz = (uint8_t*)malloc((size_t)nnz * wb + 16);
  uint32_t got;
  float_decompress_impl(fa, nz, nnz, &got, 1);
  uint32_t k = 0;
  for (uint32_t i = 0; i < n; ++i) {
    uint8_t* dst = (uint8_t*)out + (size_t)i * wb;
    int bit = (a[16 + (i >> 3)] & (0x80u >> (i & 7))) != 0;
    /* the last word lives at idx[N-2] + 1 whatever word N-2 was (:127-144) */
    if (n >= 2 && i == n - 1) {
      int b2 = (a[16 + ((n - 2) >> 3)] & (0x80u >> ((n - 2) & 7))) != 0;
      if (!b2) ++k;
    }
    if (bit) {
      memcpy(dst, nz + (size_t)k * wb, wb);
      ++k;
    } else {
      memset(dst, 0, wb);
    }
  }
  free(nz);
  return 1;
}

/* When word N-2 is zero the reference's nested float archive depends on an uninitialised word
 * (see above): only the sparse header and the bitmap are then defined. */
uint32_t dgo_sparse_float_defined_mask(int ft, const uint8_t* a, uint8_t* mask) {
  (void)ft;
  uint32_t n = get32(a);
  uint32_t bm = round_up((n + 7) / 8, 16);
  memset(mask, 0, 16u + bm);
  memset(mask, 1, 4);
  memset(mask + 16, 1, (n + 7) / 8);
  uint32_t inner = dgo_float_defined_mask(a + 16 + bm, mask + 16 + bm);
  if (n >= 2) {
    int b2 = (a[16 + ((n - 2) >> 3)] & (0x80u >> ((n - 2) & 7))) != 0;
    if (!b2) {
      memset(mask + 16 + bm, 0, inner);
      memset(mask + 16 + bm, 1, 8); /* magic + nnz stay defined */
    }
  }
  return 16u + bm + inner;
}

/* -------------------------------------------- batched drivers (CPU baseline) */

int dgo_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

void dgo_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* Batch elements are independent (README.md:110): parallelise over elements; a single
 * large element is parallelised over its 4 KiB blocks inside *_impl. */
void dgo_ans_encode_batch(uint32_t batch, const uint8_t* const* in, const uint32_t* n, int pb,
                          uint8_t* const* out, uint32_t* out_bytes) {
  int inner = batch < (uint32_t)dgo_num_threads();
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) if (!inner)
#endif
  for (int64_t b = 0; b < (int64_t)batch; ++b)
    out_bytes[b] = ans_encode_impl(in[b], n[b], pb, 0, 0, NULL, out[b], inner);
}

void dgo_ans_decode_batch(uint32_t batch, const uint8_t* const* in, uint8_t* const* out,
                          const uint32_t* capacity, uint32_t* out_size) {
  int inner = batch < (uint32_t)dgo_num_threads();
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) if (!inner)
#endif
  for (int64_t b = 0; b < (int64_t)batch; ++b)
    ans_decode_impl(in[b], out[b], capacity[b], &out_size[b], inner);
}

void dgo_float_compress_batch(int ft, uint32_t batch, const void* const* in, const uint32_t* n,
                              int pb, uint8_t* const* out, uint32_t* out_bytes) {
  int inner = batch < (uint32_t)dgo_num_threads();
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) if (!inner)
#endif
  for (int64_t b = 0; b < (int64_t)batch; ++b)
    out_bytes[b] = float_compress_impl(ft, in[b], n[b], pb, 0, out[b], inner);
}

void dgo_float_decompress_batch(uint32_t batch, const uint8_t* const* in, void* const* out,
                                const uint32_t* capacity, uint32_t* out_words) {
  int inner = batch < (uint32_t)dgo_num_threads();
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 1) if (!inner)
#endif
  for (int64_t b = 0; b < (int64_t)batch; ++b)
    float_decompress_impl(in[b], out[b], capacity[b], &out_words[b], inner);
}
