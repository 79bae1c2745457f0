// Stand-in for absl/random/uniform_real_distribution.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_UNIFORM_REAL_H_
#define BSG_SHIM_ABSL_UNIFORM_REAL_H_
#include <random>
namespace absl {
template <typename T = double>
using uniform_real_distribution = std::uniform_real_distribution<T>;
}
#endif
