/*
 * Minimal stand-in for <glog/logging.h>, used ONLY to compile the unmodified reference sources
 * (in place under /root/reference) into oracle/_ref/libdietgpu_ref.so. The reference's vendored
 * glog submodule needs cmake-generated headers; the reference uses nothing from glog except
 * CHECK / CHECK_{EQ,NE,LE,GE,LT,GT} with optional `<<` streaming, so this header provides those.
 * TEST INFRASTRUCTURE ONLY (see oracle/build_ref.sh).
 */
#pragma once
#include <cstdlib>
#include <iostream>
#include <sstream>

namespace dgo_glog_stub {
struct Fatal {
  std::ostringstream ss;
  Fatal(const char* file, int line, const char* expr) {
    ss << "CHECK failed: " << expr << " at " << file << ":" << line << " ";
  }
  template <typename T>
  Fatal& operator<<(const T& v) {
    ss << v;
    return *this;
  }
  [[noreturn]] ~Fatal() {
    std::cerr << ss.str() << std::endl;
    std::abort();
  }
};
struct Voidify {
  void operator&(Fatal&) {}
  void operator&(const Fatal&) {}
};
} // namespace dgo_glog_stub

#define DGO_CHECK_IMPL(cond, text) \
  (cond) ? (void)0 : ::dgo_glog_stub::Voidify() & ::dgo_glog_stub::Fatal(__FILE__, __LINE__, text)

#define CHECK(x) DGO_CHECK_IMPL(!!(x), #x)
#define CHECK_EQ(a, b) DGO_CHECK_IMPL((a) == (b), #a " == " #b)
#define CHECK_NE(a, b) DGO_CHECK_IMPL((a) != (b), #a " != " #b)
#define CHECK_LE(a, b) DGO_CHECK_IMPL((a) <= (b), #a " <= " #b)
#define CHECK_LT(a, b) DGO_CHECK_IMPL((a) < (b), #a " < " #b)
#define CHECK_GE(a, b) DGO_CHECK_IMPL((a) >= (b), #a " >= " #b)
#define CHECK_GT(a, b) DGO_CHECK_IMPL((a) > (b), #a " > " #b)
