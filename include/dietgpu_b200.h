/*
 * dietgpu_b200.h — C ABI of libdietgpu_b200.so, the B200-native (sm_100a) DietGPU codec.
 *
 * Every entry point is the C face of one function of the reference's C++ API
 * (namespace dietgpu, /root/reference/dietgpu/ans/GpuANSCodec.h and
 * /root/reference/dietgpu/float/GpuFloatCodec.h); argument meaning, units and host/device
 * placement are the reference's:
 *   - `in`, `in_size`, `out`, `out_capacity`, split-size arrays: HOST arrays, read during the call;
 *   - `*_dev`: DEVICE arrays, written asynchronously on `stream`;
 *   - ANS sizes are bytes; float sizes / capacities / out sizes are float WORDS; compressed
 *     sizes are always bytes;
 *   - everything is enqueued on `stream` (a cudaStream_t) of the current device; calls return
 *     without synchronising, except decode with use_checksum != 0 (the reference blocks there
 *     too, dietgpu/ans/GpuANSDecode.cuh:558-591).
 * The reference's `StackDeviceMemory& res` becomes (`temp_dev`, `temp_bytes`): a caller-owned
 * device scratch buffer. If it is NULL or too small the library takes the remainder from the
 * stream-ordered CUDA pool (the reference falls back to cudaMalloc,
 * dietgpu/utils/StackDeviceMemory.cpp:119-139). include/dietgpu_b200.hpp rebuilds the C++
 * signatures (StackDeviceMemory, ANSCodecConfig, status structs) on top of this file.
 *
 * Archives must be 16-byte aligned (as the reference requires); inputs/outputs need only the
 * natural alignment of their element type, 16-byte alignment enables the vector paths.
 *
 * Return value: 0 = ok, 1 = checksum mismatch (decode only), < 0 = error (DGB_ERR_*).
 * There is no CPU fallback: without a CUDA device every compute entry point fails.
 */
#ifndef DIETGPU_B200_H
#define DIETGPU_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGB_OK 0
#define DGB_CHECKSUM_MISMATCH 1
#define DGB_ERR_INVALID_ARG (-1)
#define DGB_ERR_CUDA (-2)
#define DGB_ERR_ALIGNMENT (-3)
#define DGB_ERR_INTERNAL (-4)
#define DGB_ERR_TOO_LARGE (-5) /* dgb_decode_batch_noalloc only: batch beyond DGB_NOALLOC_MAX_BATCH */

/* dietgpu::FloatType, dietgpu/float/GpuFloatCodec.h:18-24 */
#define DGB_FLOAT16 1
#define DGB_BFLOAT16 2
#define DGB_FLOAT32 3
#define DGB_FLOAT64 4

/* ---- sizes ------------------------------------------------------------------------------ */
/* getMaxCompressedSize, dietgpu/ans/GpuANSCodec.h:22 (dietgpu/ans/GpuANSEncode.cu:13-25) */
uint32_t dgb_ans_max_compressed_size(uint32_t uncompressed_bytes);
/* getMaxFloatCompressedSize, dietgpu/float/GpuFloatCodec.h:27 (GpuFloatCompress.cu:23-48) */
uint32_t dgb_float_max_compressed_size(int float_type, uint32_t num_words);
/* getMaxSparseFloatCompressedSize, dietgpu/float/GpuFloatCodec.h:30 (GpuSparseFloatCompress.cu:16-24) */
uint32_t dgb_sparse_float_max_compressed_size(int float_type, uint32_t num_words);
/* Upper bound on the scratch an operation takes from temp_dev for a batch of `num_in_batch`
 * elements of at most `max_size` bytes (ANS) / words (float). op: 0 ans encode, 1 ans decode,
 * 2 float compress, 3 float decompress, 4 sparse compress, 5 sparse decompress. */
size_t dgb_temp_bytes_required(int op, int float_type, uint32_t num_in_batch, uint32_t max_size);

/* ---- ANS byte codec --------------------------------------------------------------------- */
/* ansEncodeBatchPointer, dietgpu/ans/GpuANSCodec.h:157-176 */
int dgb_ans_encode_batch_pointer(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                 uint32_t num_in_batch, const void* const* in, const uint32_t* in_size,
                                 const uint32_t* histogram_dev, void* const* out,
                                 uint32_t* out_size_dev, void* stream);
/* ansEncodeBatchStride, dietgpu/ans/GpuANSCodec.h:98-131 */
int dgb_ans_encode_batch_stride(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                uint32_t num_in_batch, const void* in_dev, uint32_t in_per_batch_size,
                                uint32_t in_per_batch_stride, const uint32_t* histogram_dev,
                                void* out_dev, uint32_t out_per_batch_stride,
                                uint32_t* out_batch_size_dev, void* stream);
/* ansEncodeBatchSplitSize, dietgpu/ans/GpuANSCodec.h:178-214 */
int dgb_ans_encode_batch_split_size(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                    uint32_t num_in_batch, const void* in_dev,
                                    const uint32_t* in_split_sizes, const uint32_t* histogram_dev,
                                    void* out_dev, uint32_t out_stride, uint32_t* out_size_dev,
                                    void* stream);

/* ansDecodeBatchPointer, dietgpu/ans/GpuANSCodec.h:259-279. On DGB_CHECKSUM_MISMATCH the first
 * min(*num_mismatch, mismatch_capacity) failing batch indices are stored in mismatch_idx
 * (both nullable); this replaces ANSDecodeStatus::errorInfo. */
int dgb_ans_decode_batch_pointer(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                 uint32_t num_in_batch, const void* const* in, void* const* out,
                                 const uint32_t* out_capacity, uint8_t* out_success_dev,
                                 uint32_t* out_size_dev, void* stream, uint32_t* mismatch_idx,
                                 uint32_t mismatch_capacity, uint32_t* num_mismatch);
/* ansDecodeBatchStride, dietgpu/ans/GpuANSCodec.h:222-257 */
int dgb_ans_decode_batch_stride(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                uint32_t num_in_batch, const void* in_dev, uint32_t in_per_batch_stride,
                                void* out_dev, uint32_t out_per_batch_stride,
                                uint32_t out_per_batch_capacity, uint8_t* out_success_dev,
                                uint32_t* out_size_dev, void* stream, uint32_t* mismatch_idx,
                                uint32_t mismatch_capacity, uint32_t* num_mismatch);
/* ansDecodeBatchSplitSize, dietgpu/ans/GpuANSCodec.h:281-313 */
int dgb_ans_decode_batch_split_size(void* temp_dev, size_t temp_bytes, int prob_bits, int use_checksum,
                                    uint32_t num_in_batch, const void* const* in, void* out_dev,
                                    const uint32_t* out_split_sizes, uint8_t* out_success_dev,
                                    uint32_t* out_size_dev, void* stream, uint32_t* mismatch_idx,
                                    uint32_t mismatch_capacity, uint32_t* num_mismatch);

/* ansGetCompressedInfo / ansGetCompressedInfoDevice, dietgpu/ans/GpuANSCodec.h:315-341.
 * in_is_device != 0: `in` is a device array of device pointers. Outputs nullable. */
int dgb_ans_get_compressed_info(const void* const* in, int in_is_device, uint32_t num_in_batch,
                                uint32_t* out_sizes_dev, uint32_t* out_checksum_dev, void* temp_dev,
                                size_t temp_bytes, void* stream);

/* ---- float codecs ----------------------------------------------------------------------- */
/* floatCompress, dietgpu/float/GpuFloatCodec.h:105-140 */
int dgb_float_compress(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                       int use_checksum, uint32_t num_in_batch, const void* const* in,
                       const uint32_t* in_size, void* const* out, uint32_t* out_size_dev, void* stream);
/* floatCompressSplitSize, dietgpu/float/GpuFloatCodec.h:142-164 */
int dgb_float_compress_split_size(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                                  int use_checksum, uint32_t num_in_batch, const void* in_dev,
                                  const uint32_t* in_split_sizes, void* out_dev, uint32_t out_stride,
                                  uint32_t* out_size_dev, void* stream);
/* floatCompressSparse, dietgpu/float/GpuFloatCodec.h:166-176 */
int dgb_float_compress_sparse(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                              int use_checksum, uint32_t num_in_batch, const void* const* in,
                              const uint32_t* in_size, void* const* out, uint32_t* out_size_dev,
                              void* stream);

/* floatDecompress, dietgpu/float/GpuFloatCodec.h:190-224 */
int dgb_float_decompress(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                         int use_checksum, uint32_t num_in_batch, const void* const* in,
                         void* const* out, const uint32_t* out_capacity, uint8_t* out_success_dev,
                         uint32_t* out_size_dev, void* stream, uint32_t* mismatch_idx,
                         uint32_t mismatch_capacity, uint32_t* num_mismatch);
/* floatDecompressSplitSize, dietgpu/float/GpuFloatCodec.h:226-260 */
int dgb_float_decompress_split_size(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                                    int use_checksum, uint32_t num_in_batch, const void* const* in,
                                    void* out_dev, const uint32_t* out_split_sizes,
                                    uint8_t* out_success_dev, uint32_t* out_size_dev, void* stream,
                                    uint32_t* mismatch_idx, uint32_t mismatch_capacity,
                                    uint32_t* num_mismatch);
/* floatDecompressSparse, dietgpu/float/GpuFloatCodec.h:262-273 */
int dgb_float_decompress_sparse(void* temp_dev, size_t temp_bytes, int float_type, int prob_bits,
                                int use_checksum, uint32_t num_in_batch, const void* const* in,
                                void* const* out, const uint32_t* out_capacity,
                                uint8_t* out_success_dev, uint32_t* out_size_dev, void* stream);

/* floatGetCompressedInfo / floatGetCompressedInfoDevice, dietgpu/float/GpuFloatCodec.h:275-322 */
int dgb_float_get_compressed_info(const void* const* in, int in_is_device, uint32_t num_in_batch,
                                  uint32_t* out_sizes_dev, uint32_t* out_types_dev,
                                  uint32_t* out_checksum_dev, void* temp_dev, size_t temp_bytes,
                                  void* stream);

/* ---- host buffers ---------------------------------------------------------------------------- */
/* The reference's API takes device pointers only; a caller whose tensors live in HOST memory wraps it in
 * cudaMemcpy in / call / cudaMemcpy the sizes / cudaMemcpy out. These entry points are that sequence as one
 * call, pipelined over groups of batch elements on three internal streams with their own device staging, so
 * that the host-to-device copy of one group, the codec kernels of the next and the device-to-host copy of a
 * third overlap (PCIe is full duplex; csrc/host_pipe.cu). Archives are byte-identical to the device calls'.
 *   - every pointer is a HOST pointer (pinned memory for the copies to overlap; pageable works, serialised);
 *   - in_size / out_capacity of the float calls are WORDS, of the ANS calls BYTES; archive sizes are bytes;
 *   - compress: out_capacity (nullable) = bytes available at out[i]; an archive that does not fit is not
 *     copied, its size is still reported and the call returns DGB_ERR_INVALID_ARG;
 *   - decompress: in_bytes[i] = size of archive i as reported by compress;
 *   - the calls are ordered after the work already enqueued on `stream` and return when every result is in
 *     host memory. Device staging is cached per device until dgb_host_release();
 *   - thread safe: concurrent callers on the same device take turns (one pipeline and one lock per device), callers
 *     on different devices do not wait for each other. */
int dgb_float_compress_host(int float_type, int prob_bits, int use_checksum, uint32_t num_in_batch,
                            const void* const* in, const uint32_t* in_size, void* const* out,
                            const uint32_t* out_capacity, uint32_t* out_size, void* stream);
int dgb_float_decompress_host(int float_type, int prob_bits, int use_checksum, uint32_t num_in_batch,
                              const void* const* in, const uint32_t* in_bytes, void* const* out,
                              const uint32_t* out_capacity, uint8_t* out_success, uint32_t* out_size,
                              void* stream);
int dgb_ans_encode_host(int prob_bits, int use_checksum, uint32_t num_in_batch, const void* const* in,
                        const uint32_t* in_size, void* const* out, const uint32_t* out_capacity,
                        uint32_t* out_size, void* stream);
int dgb_ans_decode_host(int prob_bits, int use_checksum, uint32_t num_in_batch, const void* const* in,
                        const uint32_t* in_bytes, void* const* out, const uint32_t* out_capacity,
                        uint8_t* out_success, uint32_t* out_size, void* stream);
/* Frees the cached staging buffers and streams of the host-buffer calls (all devices). */
void dgb_host_release(void);

/* ---- decode without any allocation -------------------------------------------------------------- */
/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Decodes a batch exactly as dgb_ans_decode_batch_pointer (float_type 0 = raw ANS bytes, sizes in bytes) or
 * dgb_float_decompress (float_type 1..4, sizes in words) do - same `in` / `out` / `out_capacity` host arrays, same
 * out_success_dev / out_size_dev device arrays (nullable), bit-identical output - but the decoder's scratch
 * (descriptor table, per-element records, decode tables, work counter) lives in one fixed `static __device__`
 * array inside the library (about 0.5 MiB, resident on every GPU that loads the library until the process exits),
 * and the call performs no cudaMalloc / cudaFree / cudaMallocAsync / cudaMallocHost and no allocating library call.
 *   - at most DGB_NOALLOC_MAX_BATCH batch elements per call, of any size; more returns DGB_ERR_TOO_LARGE (with a
 *     message on stderr) and launches nothing;
 *   - no checksum verification (it needs a read-back; use the ordinary entry points for that);
 *   - all calls on one device share the scratch array: calls ordered on one stream are safe, calls that OVERLAP on
 *     the GPU (other streams, unordered host threads) race on it and must be serialised by the caller. */
#define DGB_NOALLOC_MAX_BATCH 32
int dgb_decode_batch_noalloc(int float_type, int prob_bits, uint32_t num_in_batch, const void* const* in,
                             void* const* out, const uint32_t* out_capacity, uint8_t* out_success_dev,
                             uint32_t* out_size_dev, void* stream);

/* ---- peer buffers (compressed collectives inside one NVLink / NVSwitch node) ------------------- */
/* A device buffer that the other ranks (processes) of the node can map: cudaMalloc + CUDA IPC. A rank
 * compresses into its own buffer; after a stream-ordered barrier every rank passes the PEER pointers as `in`
 * of dgb_float_decompress / dgb_ans_decode_batch_pointer, and the decoder pulls exactly the compressed bytes
 * over NVLink while it decodes (dietgpu_fork_b200/dist.py: all_gather_compressed_peer,
 * all_to_all_compressed_peer). No reference counterpart (the reference is single-GPU). */
#define DGB_PEER_HANDLE_BYTES 64
/* allocates `bytes` on the current device; writes the device pointer and its 64-byte IPC handle */
int dgb_peer_alloc(size_t bytes, void** dev_ptr, void* handle64);
/* maps another process's buffer (peer access is enabled on first use); *peer_ptr is valid on this device */
int dgb_peer_open(const void* handle64, void** peer_ptr);
int dgb_peer_close(void* peer_ptr);
int dgb_peer_free(void* dev_ptr);
/* Copies `num` float archives from (peer or local) device memory to local device memory on `stream`; each archive's
 * length is taken from its own headers ON THE DEVICE (no size exchange, no host synchronisation). src_dev, dst_dev and
 * dst_capacity_dev are DEVICE arrays of `num` entries (16-byte aligned archives / destinations, capacities in bytes);
 * max_bytes (host value, an upper bound of any one archive) only sizes the launch. An archive whose headers are not
 * sane or that does not fit its capacity arrives as a zeroed header, which the decoder reports as failed.
 * dist.py pulls peer k + 1 on a copy stream while it decodes peer k. No reference counterpart. */
int dgb_peer_pull_float(uint32_t num, const void* const* src_dev, void* const* dst_dev,
                        const uint32_t* dst_capacity_dev, uint32_t max_bytes, void* stream);

/* Scratch bytes the last finished call on this thread took (from temp_dev plus the pool): what the
 * reference reports through StackDeviceMemory::getMaxMemoryUsage(),
 * dietgpu/utils/StackDeviceMemory.h:196-200. */
size_t dgb_last_temp_bytes_used(void);

/* ---- tuning ----------------------------------------------------------------------------------- */
/* Compress chooses between two bit-identical schedules: one persistent fused launch (split or byte
 * histogram + table + encode + compaction) for batches of at least `tiles` 8-block tiles (32768 symbols
 * each), separate launches below. Default: 1024 tiles with at least 16 elements, 2048 with at least 2,
 * 8192 for a single element (measured cross-overs on B200); 0 = always fused,
 * a huge value = never, negative = back to the default. Process-wide; used by the parity tests to
 * cover both schedules at every size. */
void dgb_set_fused_min_tiles(long long tiles);

/* ---- launch accounting (bench.py) --------------------------------------------------------- */
/* Total kernel launches the library has enqueued in this process. */
unsigned long long dgb_profile_launches(void);
/* When enabled, every launch is bracketed by CUDA events on its stream. */
void dgb_profile_enable(int on);
/* Waits for the recorded launches and writes "kernel count total_ms\n" lines into buf; clears the
 * records; returns the number of distinct kernels. */
int dgb_profile_report(char* buf, size_t cap);

/* library version / build tag, e.g. "dietgpu_b200 r2 sm_100a" */
const char* dgb_version(void);

#ifdef __cplusplus
}
#endif
#endif /* DIETGPU_B200_H */
