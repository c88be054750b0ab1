/*
 * dietgpu_oracle.h — CPU restatement of the DietGPU (NSagan271/dietgpu_fork) stream formats.
 *
 * TEST INFRASTRUCTURE ONLY. Nothing in the product path (dietgpu_fork_b200/) may include, link
 * or call this. It is used by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs as the checker and as the timed CPU baseline.
 *
 * Parity status: the reference holds no golden byte streams (every reference test is a self
 * round trip). This oracle is pinned (a) against the reference's known-answer
 * tests for table normalisation (dietgpu/ans/ANSStatisticsTest.cu:127-167) and (b) on the GPU
 * box against the reference's own CUDA build (oracle/_ref/libdietgpu_ref.so, built by
 * oracle/build_ref.sh from the sources in /root/reference) under the defined-byte mask below.
 *
 * All citations are relative to /root/reference/.
 */
#ifndef DIETGPU_ORACLE_H
#define DIETGPU_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* dietgpu/float/GpuFloatCodec.h:18-24 */
enum { DGO_FLOAT16 = 1, DGO_BFLOAT16 = 2, DGO_FLOAT32 = 3, DGO_FLOAT64 = 4 };

/* ---- sizes (dietgpu/ans/GpuANSEncode.cu:13-25, float/GpuFloatCompress.cu:23-48,
 *      float/GpuSparseFloatCompress.cu:16-24) ---- */
uint32_t dgo_ans_max_compressed_size(uint32_t uncompressed_bytes);
uint32_t dgo_float_max_compressed_size(int float_type, uint32_t num_words);
uint32_t dgo_sparse_float_max_compressed_size(int float_type, uint32_t num_words);
uint32_t dgo_float_word_bytes(int float_type);
/* bytes of the raw (non-compressed) section, FloatTypeInfo<FT>::getUncompDataSize */
uint32_t dgo_float_raw_section_bytes(int float_type, uint32_t num_words);

/* ---- statistics (dietgpu/ans/GpuANSStatistics.cuh:21-143, :178-367) ---- */
void dgo_histogram(const uint8_t* in, uint32_t n, uint32_t counts[256]);
/* returns 0 and leaves pdf/cdf untouched when total == 0 */
int dgo_normalize(const uint32_t counts[256], uint32_t total, int prob_bits,
                  uint32_t pdf[256], uint32_t cdf[256]);
/* XOR-fold 8-bit checksum, dietgpu/ans/GpuChecksum.cuh:26-93 */
uint32_t dgo_checksum(const uint8_t* in, uint32_t n);

/* ---- ANS byte codec; one archive ---- */
/* hist may be NULL (computed from the input). Returns archive bytes. Undefined bytes of the
 * reference's stream are written as 0. */
uint32_t dgo_ans_encode(const uint8_t* in, uint32_t n, int prob_bits, int use_checksum,
                        const uint32_t* hist, uint8_t* out);
/* returns 1 on success (capacity >= size), 0 otherwise; *out_size = uncompressed size */
int dgo_ans_decode(const uint8_t* archive, uint8_t* out, uint32_t capacity, uint32_t* out_size);
/* mask[i] = 1 where the reference's stream byte i is defined (always written), 0 where the
 * reference leaves it uninitialised. Returns archive bytes. */
uint32_t dgo_ans_defined_mask(const uint8_t* archive, uint8_t* mask);

/* ---- float codecs (rows F1-F4) ---- */
uint32_t dgo_float_compress(int float_type, const void* in, uint32_t num_words, int prob_bits,
                            int use_checksum, uint8_t* out);
int dgo_float_decompress(const uint8_t* archive, void* out, uint32_t capacity_words,
                         uint32_t* out_words);
uint32_t dgo_float_defined_mask(const uint8_t* archive, uint8_t* mask);

/* ---- sparse float (rows S1-S2) ---- */
uint32_t dgo_sparse_float_compress(int float_type, const void* in, uint32_t num_words,
                                   int prob_bits, int use_checksum, uint8_t* out);
int dgo_sparse_float_decompress(int float_type, const uint8_t* archive, void* out,
                                uint32_t capacity_words, uint32_t* out_words);
uint32_t dgo_sparse_float_defined_mask(int float_type, const uint8_t* archive, uint8_t* mask);

/* ---- batched, multithreaded (OpenMP) drivers: the timed CPU baseline ---- */
int dgo_num_threads(void);
void dgo_set_num_threads(int n);
void dgo_ans_encode_batch(uint32_t batch, const uint8_t* const* in, const uint32_t* n,
                          int prob_bits, uint8_t* const* out, uint32_t* out_bytes);
void dgo_ans_decode_batch(uint32_t batch, const uint8_t* const* in, uint8_t* const* out,
                          const uint32_t* capacity, uint32_t* out_size);
void dgo_float_compress_batch(int float_type, uint32_t batch, const void* const* in,
                              const uint32_t* n, int prob_bits, uint8_t* const* out,
                              uint32_t* out_bytes);
void dgo_float_decompress_batch(uint32_t batch, const uint8_t* const* in, void* const* out,
                                const uint32_t* capacity, uint32_t* out_words);

#ifdef __cplusplus
}
#endif
#endif
