// Stand-in for absl/random/distributions.h. Test infrastructure only.
#ifndef BSG_SHIM_ABSL_DISTRIBUTIONS_H_
#define BSG_SHIM_ABSL_DISTRIBUTIONS_H_
#include "open_spiel/abseil-cpp/absl/random/bit_gen_ref.h"
#endif
