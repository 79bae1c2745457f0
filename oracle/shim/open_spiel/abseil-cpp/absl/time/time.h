// Stand-in for absl/time/time.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_TIME_H_
#define BSG_SHIM_ABSL_TIME_H_
#include <chrono>
#include <cstdint>
namespace absl {
struct Duration { int64_t ns = 0; };
struct Time { int64_t ns = 0; };
inline Time UnixEpoch() { return Time{0}; }
inline Duration operator-(Time a, Time b) { return Duration{a.ns - b.ns}; }
inline int64_t ToInt64Nanoseconds(Duration d) { return d.ns; }
inline double ToDoubleSeconds(Duration d) { return d.ns * 1e-9; }
}  // namespace absl
#endif
