// Stand-in for absl/time/clock.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_CLOCK_H_
#define BSG_SHIM_ABSL_CLOCK_H_
#include <chrono>
#include "open_spiel/abseil-cpp/absl/time/time.h"
namespace absl {
inline Time Now() {
  return Time{std::chrono::duration_cast<std::chrono::nanoseconds>(
                  std::chrono::system_clock::now().time_since_epoch()).count()};
}
}  // namespace absl
#endif
