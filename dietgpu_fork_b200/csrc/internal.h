// Host-side plumbing shared by the translation units of libdietgpu_b200.so: the per-element
// descriptor that every kernel of a call indexes, the stream-ordered temp allocator, status codes.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <vector>

#include "common.cuh"

namespace dgb {

// status codes of the C ABI (include/dietgpu_b200.h)
enum Status : int {
  kOk = 0,
  kErrInvalidArg = -1,
  kErrCuda = -2,
  kErrAlignment = -3,
  kErrInternal = -4,
  kErrTooLarge = -5, // dgb_decode_batch_noalloc: the batch exceeds the fixed scratch array
  kChecksumMismatch = 1,
};

// One per batch element (+1 sentinel carrying totalTiles in tileBase). Built on the host from the
// caller's host-side pointer/size arrays, shipped with one H2D copy, indexed by every kernel.
struct Elem {
  const void* in;    // compress: caller's input        decompress: archive
  void* out;         // compress: archive               decompress: caller's output
  uint32_t n;        // compress: symbols/words (host)  decompress: capacity (symbols/words)
  uint32_t tileBase; // first tile of this element in the call's work list
  uint64_t planeOff; // byte offset of this element's row in the exponent-plane temp (16-B aligned)
  uint32_t sizeBias; // added to the reported compressed size (sparse: header + bitmap bytes)
  uint32_t blockBase; // first 4096-symbol block of this element in the call's block list (encoder)
  uint32_t groupBase; // first 256-block size group of this element (compaction)
  uint32_t pad;
};
static_assert(sizeof(Elem) == 48, "Elem layout");

// Per (element, plane) decode descriptor produced on the device by k_decodePrep.
struct DecInfo {
  const uint8_t* ans;   // ANS archive base (16-B aligned)
  const uint8_t* fbase; // float archive base (float modes), else null
  uint32_t nb;          // blocks
  uint32_t n;         // symbols
  uint32_t ok;        // capacity >= n and header sane
  uint32_t totalWords; // payload length in 16-bit words (header): bounds the block index entries
};

// Bump allocator over the caller's temp buffer; anything that does not fit is taken from the
// stream-ordered pool (cudaMallocAsync) and returned to it when the call has been enqueued, so a
// short temp buffer never synchronises the host (the reference warns and cudaMallocs,
// dietgpu/utils/StackDeviceMemory.cpp:119-139).
class Temp {
 public:
  Temp(void* base, size_t bytes, cudaStream_t s) : base_((uint8_t*)base), cap_(base ? bytes : 0), s_(s) {
    // align the base
    size_t mis = (size_t)((uintptr_t)base_ & 255u);
    if (base_ && mis) {
      size_t skip = 256 - mis;
      if (skip > cap_) skip = cap_;
      base_ += skip;
      cap_ -= skip;
    }
  }
  ~Temp() {
    for (void* p : overflow_) cudaFreeAsync(p, s_);
    lastHighWater() = highWater();
  }
  // high-water mark of the last finished call on this thread (getMaxMemoryUsage() analogue,
  // dietgpu/utils/StackDeviceMemory.cpp:218-224)
  static size_t& lastHighWater() {
    static thread_local size_t v = 0;
    return v;
  }
  void* alloc(size_t bytes) {
    bytes = roundUpSz(bytes ? bytes : 1, 256);
    if (off_ + bytes <= cap_) {
      void* p = base_ + off_;
      off_ += bytes;
      used_ = off_ > used_ ? off_ : used_;
      return p;
    }
    void* p = nullptr;
    if (cudaMallocAsync(&p, bytes, s_) != cudaSuccess) {
      failed_ = true;
      return nullptr;
    }
    overflow_.push_back(p);
    overflowBytes_ += bytes;
    return p;
  }
  template <typename T>
  T* alloc(size_t count) { return (T*)alloc(count * sizeof(T)); }
  bool failed() const { return failed_; }
  size_t highWater() const { return used_ + overflowBytes_; }

 private:
  uint8_t* base_;
  size_t cap_;
  size_t off_ = 0, used_ = 0, overflowBytes_ = 0;
  cudaStream_t s_;
  std::vector<void*> overflow_;
  bool failed_ = false;
};

#define DGB_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t e__ = (expr);                                                            \
    if (e__ != cudaSuccess) {                                                            \
      fprintf(stderr, "dietgpu_b200: CUDA error %s at %s:%d\n", cudaGetErrorString(e__), \
              __FILE__, __LINE__);                                                       \
      return kErrCuda;                                                                   \
    }                                                                                    \
  } while (0)

// current device, clamped to [0, kMaxDevices)
constexpr int kMaxDevices = 64;
int currentDevice();
// SM count of the current device (cached per device)
int smCount();

// "do this once per device" guard for per-kernel attribute calls (one static instance per call site)
struct DeviceOnce {
  unsigned long long mask = 0;
  // true exactly for the first caller per device (benign race: a second thread may repeat the idempotent call)
  bool first(int dev) {
    const unsigned long long bit = 1ull << (dev & 63);
    const unsigned long long seen = __atomic_load_n(&mask, __ATOMIC_ACQUIRE);
    if (seen & bit) return false;
    return true;
  }
  void done(int dev) { __atomic_fetch_or(&mask, 1ull << (dev & 63), __ATOMIC_RELEASE); }
};

// ---- launchers implemented in ans.cu ------------------------------------------------------
struct EncodeArgs {
  const Elem* elems;        // device, numElems + 1
  uint32_t numElems;
  uint32_t totalTiles;      // total 4096-symbol blocks (Elem::blockBase of the sentinel)
  const uint32_t* nDev;     // nullable: device-side symbol counts overriding Elem::n (sparse nnz)
  const uint8_t* planeBase; // nullable: symbols come from planeBase + Elem::planeOff instead of Elem::in
  const uint32_t* hist;     // device [numElems][256]
  const uint32_t* checksum; // nullable device [numElems]
  uint32_t* outSize;        // nullable device [numElems]
  int floatType;            // kFtNone: Elem::out is the ANS archive; else the float archive
  int round;                // 0, or 1 for the second fp64 exponent plane
  int probBits;
  int useChecksum;          // ANS-level checksum flag stored in the ANS header
};
// Enqueues normalise + encode for one exponent plane / byte stream. Scratch comes from `tmp`.
// A slice of the work list: elements [e0, e1) with their 8-block tiles and 4096-symbol blocks.
// fused kernel control words: [0] split counter (fine prefix) | [1, 513) per-SM role tickets | [513] compaction
// claim counter | [514] split counter (coarse claims after the prefix) | [515] unused
constexpr uint32_t kFusedCtrlFixed = 516;
constexpr uint32_t kFusedCtrlCoal = 513, kFusedCtrlSplitB = 514;

struct WorkRange {
  uint32_t e0, e1, tile0, tile1, blk0, blk1;
};
// Scratch of one encode pass (one exponent plane / byte stream) over the whole work list.
struct EncodeScratch {
  uint2* table;
  uint32_t* blockWords;
  uint32_t* groupWords;
  uint32_t* runCounters;
  uint32_t* fusedCtrl; // [kFusedCtrlFixed + 3 * numElems + tiles]: split counter, SM tickets, compaction claims | tilesDone | ready | runsDone | tileWords
  uint8_t* scratch;
};
// totalGroups = number of 256-block size groups of the work list. Allocates and zeroes.
int allocEncodeScratch(Temp& tmp, const EncodeArgs& a, uint32_t totalGroups, EncodeScratch& es, cudaStream_t s);
// Enqueues normalise + encode + compaction for the elements of `r` on stream s.
// tablesReady: the encode tables were already built by the histogram / split kernel (no k_normalize launch)
int launchAnsEncode(const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r, cudaStream_t s,
                    bool tablesReady = false);
// compaction only (k_coalesce), for callers that produced the block slots themselves
int launchCoalesce(const EncodeArgs& a, const EncodeScratch& es, const WorkRange& r, cudaStream_t s);
// float split (ft == kFtNone: byte histogram of the raw input) + table normalisation + block encoding +
// compaction in ONE persistent launch. Single-plane float types or raw ANS, host-known sizes, dense input (fused.cu).
// foldChecksum: `checksum` is a zeroed array the kernel accumulates into (no k_checksum pass beforehand)
int launchFusedCompress(int ft, const EncodeArgs& a, const EncodeScratch& es, uint32_t totalTiles,
                             const uint32_t* checksum, int useChecksum, cudaStream_t s, bool foldChecksum = false);
// normTilesDone (nullable, zeroed [numElems]): the CTA completing an element's histogram also builds its table
// checksumAcc (nullable, zeroed [numElems]): the histogram pass also accumulates the checksum of every element
int launchHistogram(const Elem* elems, uint32_t numElems, uint32_t totalTiles, uint32_t* hist, cudaStream_t s,
                    uint32_t* normTilesDone = nullptr, uint2* normTable = nullptr, int normProbBits = 0,
                    uint32_t* checksumAcc = nullptr);
// XOR-fold checksum of the first n * bytesPerUnit bytes of every element's input (useOut = 0) or
// output (useOut = 1); n = nDev[e] when given, else Elem::n. maxUnits bounds n.
int launchChecksum(const Elem* elems, uint32_t numElems, uint32_t maxUnits, uint32_t bytesPerUnit,
                   const uint32_t* nDev, int useOut, uint32_t* checksum, cudaStream_t s);

int launchFloatSplit(int ft, const Elem* elems, uint32_t numElems, uint32_t tileBegin, uint32_t tileEnd,
                     const uint32_t* nDev,
                     const uint8_t* denseIn, uint8_t* plane0, uint8_t* plane1, uint32_t* hist0,
                     uint32_t* hist1, const uint32_t* checksum, int useChecksum, cudaStream_t s,
                     uint32_t* normTilesDone = nullptr, uint2* normTable0 = nullptr, uint2* normTable1 = nullptr,
                     int normProbBits = 0, uint32_t minElemTiles = 0, bool foldChecksum = false);

struct DecodeArgs {
  const Elem* elems; // device, numElems + 1 (tiles sized from capacities)
  uint32_t numElems;
  uint32_t totalTiles;
  int floatType;     // kFtNone raw bytes, else fused float join
  int sparseNested;  // float archive sits behind a sparse header + bitmap inside Elem::in
  int probBits;
  uint8_t* outSuccess; // nullable
  uint32_t* outSize;   // nullable
  uint8_t* denseBase;  // nullable: write to denseBase + Elem::planeOff instead of Elem::out (sparse)
};
int launchDecode(Temp& tmp, const DecodeArgs& a, cudaStream_t s);

} // namespace dgb
