// Minimal stand-in for <glog/logging.h> for translation units that include it only for CHECK-style
// macros (the reference's DietGpu.cpp:10): failures throw instead of aborting.
#pragma once
#include <sstream>
#include <stdexcept>
namespace dgb_compat {
struct CheckFail {
  std::ostringstream os;
  CheckFail(const char* file, int line, const char* what) { os << file << ":" << line << " CHECK failed: " << what << " "; }
  template <typename T> CheckFail& operator<<(const T& v) { os << v; return *this; }
  [[noreturn]] ~CheckFail() noexcept(false) { throw std::runtime_error(os.str()); }
};
struct Voidify { void operator&(const CheckFail&) {} };
} // namespace dgb_compat
#ifndef CHECK
#define CHECK(c) (c) ? (void)0 : dgb_compat::Voidify() & dgb_compat::CheckFail(__FILE__, __LINE__, #c)
#define CHECK_EQ(a, b) CHECK((a) == (b))
#define CHECK_NE(a, b) CHECK((a) != (b))
#define CHECK_LE(a, b) CHECK((a) <= (b))
#define CHECK_LT(a, b) CHECK((a) < (b))
#define CHECK_GE(a, b) CHECK((a) >= (b))
#define CHECK_GT(a, b) CHECK((a) > (b))
#endif
