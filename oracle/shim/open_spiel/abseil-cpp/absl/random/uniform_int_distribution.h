// Stand-in for absl/random/uniform_int_distribution.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_UNIFORM_INT_H_
#define BSG_SHIM_ABSL_UNIFORM_INT_H_
#include <random>
namespace absl {
template <typename T = int>
using uniform_int_distribution = std::uniform_int_distribution<T>;
}
#endif
