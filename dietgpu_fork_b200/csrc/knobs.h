// Experiment knobs (DGB_* environment variables). Each call site reads its variable ONCE per process (function-local
// static, thread-safe initialisation), so no getenv runs on a launch path after the first call. kKnobUnset marks an
// absent variable where presence matters.
#pragma once
#include <limits.h>
#include <stdlib.h>

namespace dgb {
constexpr int kKnobUnset = INT_MIN;
inline int envInt(const char* name, int dflt) {
  const char* e = getenv(name);
  return (e && *e) ? atoi(e) : dflt;
}
} // namespace dgb

#define DGB_KNOB(name, dflt)                      \
  ([]() -> int {                                  \
    static const int dgbKnobValue_ = ::dgb::envInt(name, dflt); \
    return dgbKnobValue_;                         \
  }())
