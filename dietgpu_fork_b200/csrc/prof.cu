#include "prof.h"

#include <stdio.h>
#include <string.h>

#include <atomic>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/dietgpu_b200.h"

namespace dgb {
namespace {
std::atomic<unsigned long long> gLaunches{0};
std::atomic<int> gEnabled{0};
struct Rec {
  const char* name;
  cudaEvent_t a, b;
};
std::mutex gMu;
std::vector<Rec> gRecs;
} // namespace

ProfScope::ProfScope(const char* name, cudaStream_t s) : name_(name), s_(s) {
  gLaunches.fetch_add(1, std::memory_order_relaxed);
  if (gEnabled.load(std::memory_order_relaxed)) {
    cudaEventCreate(&start_);
    cudaEventRecord(start_, s_);
  }
}

ProfScope::~ProfScope() {
  if (start_) {
    cudaEvent_t stop;
    cudaEventCreate(&stop);
    cudaEventRecord(stop, s_);
    std::lock_guard<std::mutex> g(gMu);
    gRecs.push_back({name_, start_, stop});
  }
}
} // namespace dgb

using namespace dgb;

extern "C" {

void dgb_profile_enable(int on) { gEnabled.store(on ? 1 : 0); }

unsigned long long dgb_profile_launches(void) { return gLaunches.load(); }

// Waits for the recorded launches, writes "name count total_ms\n" lines, clears the records.
int dgb_profile_report(char* buf, size_t cap) {
  std::vector<Rec> recs;
  {
    std::lock_guard<std::mutex> g(gMu);
    recs.swap(gRecs);
  }
  std::map<std::string, std::pair<unsigned, double>> agg;
  for (auto& r : recs) {
    float ms = 0.f;
    cudaEventSynchronize(r.b);
    cudaEventElapsedTime(&ms, r.a, r.b);
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
    auto& e = agg[r.name];
    e.first += 1;
    e.second += ms;
  }
  size_t off = 0;
  if (cap) buf[0] = 0;
  for (auto& kv : agg) {
    int n = snprintf(buf + off, off < cap ? cap - off : 0, "%s %u %.6f\n", kv.first.c_str(), kv.second.first,
                     kv.second.second);
    if (n < 0 || off + (size_t)n >= cap) break;
    off += (size_t)n;
  }
  return (int)agg.size();
}

} // extern "C"
